"""CPU tests: the numpy restatement (oracle/restate.py) against the golden vectors produced by the
UNMODIFIED reference (tests/golden/make_golden.py), and — when oracle/_ref is built — against the
compiled reference live.

Tolerances.  The restatement reproduces the reference's rounding POINTS (F16 activations, F16 KV,
F16 softmax probabilities / flash accumulator) but not the AVX2 summation ORDER of
ggml_vec_dot_f16, so an F32 sum that lands within ~1e-7 of an F16 rounding boundary can round the
other way.  With random-init weights (residual stream ~0.02, RMSNorm gain ~50x) such flips grow
to ~1e-3 absolute on logits of magnitude ~1.  Stated tolerances: NMSE <= 1e-5 and
max|d| <= 4e-3 for teacher-forced logits; exact ids wherever the reference's own top-2 margin
exceeds 2*max|d|.
"""
import os

import numpy as np
import pytest

from conftest import nmse

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NMSE_TOL = 1e-5
ABS_TOL = 4e-3


def load_gold(preset):
    p = os.path.join(GOLD, f"ref_{preset}_seed1234.npz")
    if not os.path.exists(p):
        pytest.skip(f"{p} missing")
    return np.load(p)


@pytest.fixture(scope="module", params=["tiny", "tiny-mtp"])
def setup(request, pkg, make_gguf):
    import restate
    path, cfg = make_gguf(request.param, 1234)
    model = restate.Model.from_gguf(path, pkg.gguf_io.parse_config)
    return request.param, restate, model, load_gold(request.param)


def test_build_prefill(setup):
    _, restate, model, g = setup
    S = restate.Restatement(model, n_ctx=64)
    pre, trail, pad = S.build_prefill(g["tokens"], g["speaker"])
    assert pre.shape == g["prefill"].shape and trail.shape == g["trailing"].shape
    for a, b in ((pre, g["prefill"]), (trail, g["trailing"]), (pad, g["pad"])):
        assert np.abs(a - b).max() <= 1e-6


def test_forward_prefill_and_steps(setup):
    _, restate, model, g = setup
    S = restate.Restatement(model, n_ctx=64)
    hid, logits = S.forward_prefill(g["prefill"])
    assert nmse(logits, g["pre_logits"]) <= NMSE_TOL and np.abs(logits - g["pre_logits"]).max() <= ABS_TOL
    assert nmse(hid, g["pre_hidden"]) <= NMSE_TOL
    assert int(logits[:2048].argmax()) == int(g["cp_cb0"])
    n_past = g["prefill"].shape[0]
    for i in range(g["step_embd"].shape[0]):
        lg, hd = S.forward_step(g["step_embd"][i], n_past + i)
        assert nmse(lg, g["step_logits"][i]) <= NMSE_TOL
        assert np.abs(lg - g["step_logits"][i]).max() <= ABS_TOL
        assert nmse(hd, g["step_hidden"][i]) <= NMSE_TOL


def test_code_predictor_teacher_forced(setup):
    _, restate, model, g = setup
    S = restate.Restatement(model, n_ctx=64)
    codes, logits = S.predict_codes(g["pre_hidden"], int(g["cp_cb0"]), forced=g["cp_codes"])
    for s in range(15):
        d = np.abs(logits[s] - g["cp_logits"][s]).max()
        assert nmse(logits[s], g["cp_logits"][s]) <= NMSE_TOL and d <= ABS_TOL
        top2 = np.sort(g["cp_logits"][s])[-2:]
        if top2[1] - top2[0] > 2 * d:
            assert codes[s] == g["cp_codes"][s]


def test_generate_prefix(setup):
    """Free-running greedy ids: identical until the first near-tie; the first frame must match in
    full and every code that differs must sit at a frame where an earlier code already diverged
    or be explained by the teacher-forced tolerance (checked above)."""
    _, restate, model, g = setup
    S = restate.Restatement(model, n_ctx=128)
    n = 12
    out = S.generate(g["tokens"], g["speaker"], max_len=n)
    ref = g["codes"][:n]
    assert out.shape == ref.shape
    assert np.array_equal(out[0], ref[0])
    bad = (out != ref).any(axis=1)
    first_bad = int(np.argmax(bad)) if bad.any() else n
    assert first_bad >= 4, f"restatement diverges from the reference already at frame {first_bad}"


# ---- sampling restatement: known answers worked by hand from src/tts_transformer.cpp:2417-2495 ----
def test_cb0_postprocess_known_answers():
    import restate
    V, eos = 3072, 2150
    lg = np.zeros(V, np.float32)
    lg[5], lg[7], lg[2100], lg[eos], lg[3000] = 2.0, -1.0, 9.0, 0.5, 50.0
    out = restate.cb0_postprocess(lg, {5, 7}, frame=0, n_text_tokens=21, rep_pen=1.05, eos_id=eos, vocab=V)
    assert np.isneginf(out[2048:eos]).all() and np.isneginf(out[eos + 1:]).all()
    assert out[eos] == np.float32(0.5)
    assert out[5] == np.float32(2.0) / np.float32(1.05) and out[7] == np.float32(-1.0) * np.float32(1.05)
    # EOS boost: expected = max(20, 84) = 84; frame 126 -> ramp 0.5; max = 2/1.05
    out = restate.cb0_postprocess(lg, {5}, frame=126, n_text_tokens=21, rep_pen=1.05, eos_id=eos, vocab=V)
    mx = np.float32(2.0) / np.float32(1.05)
    want = np.float32(0.5) + np.float32(0.5) * (np.float32(mx + np.float32(5.0)) - np.float32(0.5))
    assert out[eos] == np.float32(want)
    # ramp saturates at 1 -> eos = max + 5 -> argmax is EOS
    out = restate.cb0_postprocess(lg, set(), frame=400, n_text_tokens=21, rep_pen=1.05, eos_id=eos, vocab=V)
    assert restate.argmax_first(out) == eos


def test_sample_topk_known_answers():
    import restate
    lg = np.array([1.0, 3.0, 3.0, 2.0, -5.0, 0.0], np.float32)
    # top-2 threshold = 3.0, ties kept -> support {1,2}, p = .5/.5
    assert restate.sample_topk(lg, 1.0, 2, 0.49) == 1
    assert restate.sample_topk(lg, 1.0, 2, 0.51) == 2
    assert restate.sample_topk(lg, 1.0, 2, 1.0) in (2, 5)     # cum reaches 1.0 at index 2
    # EOS (index 4) survives top-k for codebook 0 even though it is below the threshold
    hits = {restate.sample_topk(lg, 1.0, 2, u, eos_id=4) for u in np.linspace(1e-6, 0.99999, 4001)}
    assert hits == {1, 2, 4}
    # a draw of exactly 0 selects index 0 even at probability 0: the scan starts with cumsum = probs[0] >= 0 (:2490-2494)
    assert restate.sample_topk(lg, 1.0, 2, 0.0, eos_id=4) == 0
    # argmax: first maximum wins
    assert restate.argmax_first(lg) == 1


def test_flash_matches_softmax_attention_loosely():
    """The two attention restatements compute the same function up to F16 accumulator error."""
    import restate
    r = np.random.default_rng(3)
    q = r.standard_normal((4, 128)).astype(np.float32)
    K = restate.f16r(r.standard_normal((40, 2, 128)))
    V = restate.f16r(r.standard_normal((40, 2, 128)))
    a = restate.attn_flash(q, K, V, 1 / np.sqrt(128.0))
    b = restate.attn_softmax_p16(q, K, V, 1 / np.sqrt(128.0))
    assert nmse(a, b) < 1e-4
    # split-KV path (>= 512 keys) agrees with the single-chunk path
    K = restate.f16r(r.standard_normal((600, 2, 128)))
    V = restate.f16r(r.standard_normal((600, 2, 128)))
    assert nmse(restate.attn_flash(q, K, V, 0.088, n_threads=4), restate.attn_flash(q, K, V, 0.088, n_threads=1)) < 1e-4


# ---- live check against the compiled reference, when it is built (CPU container) ------------------
def test_restatement_vs_live_reference(pkg, make_gguf, ref_mod):
    import restate
    path, cfg = make_gguf("tiny", 4321)      # a seed with no committed fixture
    toks = pkg.synth.synth_tokens(cfg, n_text=17, seed=5)
    spk = pkg.synth.synth_speaker(cfg, seed=9)
    R = ref_mod.Reference(path)
    S = restate.Restatement(restate.Model.from_gguf(path, pkg.gguf_io.parse_config), n_ctx=64)
    pre_r, _, _ = R.build_prefill(toks, spk)
    pre_s, _, _ = S.build_prefill(toks, spk)
    assert np.abs(pre_r - pre_s).max() <= 1e-6
    R.init_kv(64)
    hr, lr = R.forward_prefill(pre_r)
    hs, ls = S.forward_prefill(pre_r)
    assert nmse(ls, lr) <= NMSE_TOL and nmse(hs, hr) <= NMSE_TOL
    R.close()
