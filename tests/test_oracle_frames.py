"""CPU tests of the round-2 oracle pieces: the per-frame fixtures (tests/golden/*_frames.npz, dumped from the
UNMODIFIED reference by replaying generate() through its own methods), the restatement of the codebook-0
post-processing / sampling against them, and the product's restatement of libstdc++'s introselect
(csrc/qtts_nth.cuh, compiled for the host inside the oracle harness) against std::nth_element itself."""
import os

import numpy as np
import pytest

from test_oracle import load_gold

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EOS, V = 2150, 3072


def frames(preset):
    p = os.path.join(GOLD, f"ref_{preset}_seed1234_frames.npz")
    if not os.path.exists(p):
        pytest.skip(f"{p} missing")
    return np.load(p)


@pytest.mark.parametrize("preset", ["tiny", "tiny-mtp", "0.6b", "1.7b"])
def test_frames_fixture_consistent_with_base_fixture(preset):
    g, gf = load_gold(preset), frames(preset)
    assert np.array_equal(gf["g_codes"], g["codes"])
    assert np.array_equal(gf["g_raw"][0], g["pre_logits"]) and np.array_equal(gf["g_hidden"][0], g["pre_hidden"])
    # greedy id of every stored frame = argmax of the post-processed stored logits
    import restate
    seen = set()
    for f in range(gf["g_raw"].shape[0]):
        lg = restate.cb0_postprocess(gf["g_raw"][f], seen, f, len(g["tokens"]), 1.05, EOS, V)
        assert restate.argmax_first(lg) == gf["g_codes"][f, 0]
        seen.add(int(gf["g_codes"][f, 0]))


@pytest.mark.parametrize("preset", ["tiny", "tiny-mtp", "0.6b", "1.7b"])
def test_cb0_postprocess_bit_exact_vs_reference_ramp(preset):
    """restate.cb0_postprocess == the reference's processed logits (suppress, repetition penalty over the set,
    EOS ramp values inside the ramp) bit for bit, and the reference stops where the ramp makes EOS the argmax."""
    import restate
    gf = frames(preset)
    toks, codes, stop = gf["e_tokens"], gf["e_codes"], int(gf["e_eos_frame"])
    expected = max(20, 4 * len(toks))
    assert stop == len(codes) and expected <= stop <= 2 * expected + 1
    for i, f in enumerate(gf["e_frames"]):
        want = restate.cb0_postprocess(gf["e_raw"][i], set(int(c) for c in codes[:f, 0]), int(f), len(toks), 1.05, EOS, V)
        assert np.array_equal(want, gf["e_proc"][i]), f"frame {f}"
        tok = restate.argmax_first(want)
        assert tok == (EOS if f == stop else codes[f, 0])
    assert (gf["e_frames"] >= expected).sum() >= 3      # the fixture does reach into the ramp


@pytest.mark.parametrize("preset", ["tiny", "tiny-mtp", "0.6b", "1.7b"])
def test_sampled_fixture_cb0_vs_reference_sampler(preset, ref_mod):
    """The seeded reference's codebook-0 ids = ref_sample_cb0 (same libstdc++ calls as generate()) on the stored
    processed logits and uniforms; the exact-top-k restatement agrees whenever nth_element kept all top_k."""
    import restate
    gf = frames(preset)
    n_dev = 0
    for f in range(gf["s_proc"].shape[0]):
        u = float(np.float32(gf["s_u"][f, 0]))
        tok, kept = ref_mod.sample_cb0(gf["s_proc"][f], EOS, 0.9, 50, u)
        if f < len(gf["s_codes"]):
            assert tok == gf["s_codes"][f, 0]
        lg = (gf["s_proc"][f] / np.float32(0.9)).astype(np.float32)
        thr = np.sort(lg)[-50]
        n_exact = int((lg >= thr).sum()) + (0 if lg[EOS] >= thr else 1)      # exact top-50 (ties kept) + EOS re-inserted
        assert kept <= n_exact
        if kept == n_exact:
            assert restate.sample_topk(gf["s_proc"][f], 0.9, 50, u, eos_id=EOS) == tok
        else:
            n_dev += 1
    print(f"{preset}: reference kept < 50 candidates in {n_dev} of {gf['s_proc'].shape[0]} frames")


def test_introselect_restatement_equals_std_nth_element(ref_mod):
    r = np.random.default_rng(0)
    for trial in range(600):
        n = int(r.integers(4, 3200))
        k = int(r.integers(1, n))
        kind = trial % 6
        if kind == 0:
            v = r.standard_normal(n)
        elif kind == 1:
            v = np.round(r.standard_normal(n) * 2)                           # many ties
        elif kind == 2:
            v = np.sort(r.standard_normal(n))
        elif kind == 3:
            v = -np.sort(r.standard_normal(n))
        elif kind == 4:
            v = np.where(r.random(n) < 0.3, -np.inf, r.standard_normal(n))    # suppressed ids
        else:
            h = n // 2
            v = np.concatenate([np.arange(h), np.arange(n - h)[::-1]]).astype(float)   # organ pipe
        a, b = ref_mod.nth_check(v.astype(np.float32), k)
        assert np.array_equal(a, b), (trial, n, k)
    # the call the reference makes: 3072 logits / temperature, top_k = 50
    for trial in range(50):
        v = (r.standard_normal(V) / 0.9).astype(np.float32)
        v[2048:] = -np.inf
        v[EOS] = 0.1
        a, b = ref_mod.nth_check(v, 50)
        assert np.array_equal(a, b)


def test_trace_replay_equals_generate(pkg, make_gguf, ref_mod):
    """ref_generate_trace (the harness' frame-by-frame replay through the reference's own methods) reproduces the
    real generate(): greedy and sampled with the same seed."""
    path, cfg = make_gguf("tiny", 1234)
    toks = pkg.synth.synth_tokens(cfg, "Hi there")
    R = ref_mod.Reference(path)
    c, _ = R.generate(toks, None, max_len=30, temperature=0.0)
    t = R.generate_trace(toks, None, max_len=30, temperature=0.0)
    assert np.array_equal(c, t["codes"]) and t["raw"].shape[0] == len(c) + (1 if t["eos_frame"] >= 0 else 0)
    for seed in (1, 2):
        R.seed_rng(seed)
        c, _ = R.generate(toks, None, max_len=20, temperature=0.9, top_k=50)
        R.seed_rng(seed)
        t = R.generate_trace(toks, None, max_len=20, temperature=0.9, top_k=50)
        assert np.array_equal(c, t["codes"])
        R.seed_rng(seed + 10)
        c2, _ = R.generate(toks, None, max_len=20, temperature=0.9, top_k=50)
        assert not np.array_equal(c, c2) or len(c) < 3
    R.close()
