"""GPU tests of the BATCHED frame kernel (qtts_batched.cu, tcgen05 / TMEM): (1) one projection
stage on caller-supplied operands against numpy (operand layouts, descriptors, epilogues, ring and
barrier recycling), (2) whole frames: streams advanced in lock step by the batched kernel against
the same streams advanced one by one by the batch-1 kernel (which is itself pinned to the
reference's golden vectors in test_gpu_parity.py) and against the golden ids directly.

Tolerances: a projection output is an FP32 tensor-core accumulation of exact F16 x BF16 products, so
against a float64 dot product of the same rounded operands |d| <= 2e-5 * sum|terms|; whole-frame
logits against the batch-1 kernel NMSE <= 1e-5 (the stated BF16-level tolerance of
test_gpu_parity.py); greedy ids exact over the first frames.
"""
import numpy as np
import pytest

from conftest import nmse
from test_oracle import load_gold

pytestmark = pytest.mark.gpu


def f16r(x):
    return np.asarray(x, np.float32).astype(np.float16).astype(np.float32)


def ref_gemm(pkg, W, X, n_tok, epi, in_f16, norm_w, res, bias, eps):
    Wb = pkg.synth.bf16_round(W.copy()) if True else W
    X = np.asarray(X, np.float32)
    if norm_w is not None and not in_f16:
        ss = (X.astype(np.float32) ** 2).astype(np.float64).sum(axis=1)
        scale = (1.0 / np.sqrt((ss / X.shape[1]).astype(np.float32) + np.float32(eps))).astype(np.float32)
        X = (X * scale[:, None]) * norm_w[None, :].astype(np.float32)
    A = f16r(X).astype(np.float64)
    out = A @ Wb.astype(np.float64).T
    mag = np.abs(A) @ np.abs(Wb.astype(np.float64)).T
    if epi == 1:
        out = out + res
    elif epi == 3:
        out = out + bias[None, :]
    elif epi == 2:
        g, u = out[:, 0::2], out[:, 1::2]
        out = f16r(g / (1.0 + np.exp(-g)) * u).astype(np.float64)
    return out, mag


CASES = [
    # n_streams, n_tok, rows, K, epi, in_f16, norm
    (64, 1, 1024, 512, 0, False, True),
    (100, 2, 256, 1024, 0, False, True),      # two m-tiles (200 activation rows)
    (64, 1, 512, 3072, 1, True, False),       # down-projection shape: F16 input, + residual, multi-piece tiles
    (37, 1, 2048, 512, 2, False, True),       # gate/up with SwiGLU -> F16 output
    (8, 2, 512, 256, 3, False, False),        # + bias (mtp_proj)
    (128, 1, 4096, 1024, 0, False, True),     # QKV shape of the 0.6B model, full group
]


@pytest.mark.parametrize("fmt", ["f16"])
@pytest.mark.parametrize("case", CASES, ids=[f"s{c[0]}x{c[1]}_r{c[2]}_k{c[3]}_e{c[4]}" for c in CASES])
def test_bgemm_stage(pkg, case, fmt):
    n_streams, n_tok, rows, K, epi, in_f16, norm = case
    r = np.random.default_rng(rows + K + n_streams)
    W = pkg.synth.bf16_round((0.05 * r.standard_normal((rows, K))).astype(np.float32))
    X = r.standard_normal((n_streams * n_tok, K)).astype(np.float32)
    if in_f16:
        X = f16r(X)
    norm_w = (1.0 + 0.1 * r.standard_normal(K)).astype(np.float32) if norm else None
    res = r.standard_normal((n_streams * n_tok, rows)).astype(np.float32) if epi == 1 else None
    bias = r.standard_normal(rows).astype(np.float32) if epi == 3 else None
    out, ssq = pkg.debug_bgemm(W, X, n_tok=n_tok, epi=epi, in_f16=in_f16, norm_w=norm_w, res=res, bias=bias,
                               eps=1e-6, weight_fmt=pkg.BF16 if fmt == "bf16" else pkg.F16, want_ssq=True)
    want, mag = ref_gemm(pkg, W, X, n_tok, epi, in_f16, norm_w, res, bias, 1e-6)
    if epi == 2:      # F16-rounded output of a product of two accumulations
        assert np.allclose(out, want, rtol=3e-3, atol=3e-4), np.abs(out - want).max()
        return
    err = np.abs(out - want)
    assert (err <= 2e-5 * mag + 1e-6).all(), f"max err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}"
    if True:
        want_ssq = (want.astype(np.float32) ** 2).reshape(want.shape[0], rows // 16, 16).sum(axis=2)
        assert np.allclose(ssq, want_ssq, rtol=1e-3, atol=1e-4)


@pytest.fixture(scope="module")
def engines(pkg, make_gguf):
    cache = {}

    def get(preset, max_streams, max_ctx=256, fmt=None):
        key = (preset, max_streams, max_ctx, fmt)
        if key not in cache:
            path, cfg = make_gguf(preset, 1234)
            T = pkg.TTSTransformer(device=0, max_streams=max_streams, max_ctx=max_ctx)
            if fmt is not None:
                T.batched_weight_format = fmt
            assert T.load_model(path), T.get_error()
            cache[key] = (T, cfg)
        return cache[key]

    yield get
    for T, _ in cache.values():
        T.unload_model()


def run_streams(T, toks, spk, max_len, temperature=0.0, seeds=None, chunk=4, logits_at=None):
    n = len(toks)
    for i in range(n):
        assert T.stream_begin(i, toks[i], spk[i], max_len=max_len, temperature=temperature,
                              seed=0 if seeds is None else seeds[i]), T.get_error()
    got = [[] for _ in range(n)]
    for _ in range(4 * max_len):
        out = T.step_frames(chunk)
        assert out is not None, T.get_error()
        codes, produced, active = out
        for i in range(n):
            got[i].append(codes[i, :produced[i]])
        if not active.any():
            break
    return [np.concatenate(g) for g in got]


def first_diff(a, b):
    m = min(len(a), len(b))
    bad = (a[:m] != b[:m]).any(axis=1)
    return int(np.argmax(bad)) if bad.any() else m


@pytest.mark.parametrize("preset,n_streams,min_exact", [("tiny", 5, 4), ("tiny-mtp", 3, 4), ("0.6b", 4, 2)])
def test_batched_matches_sequential(engines, pkg, preset, n_streams, min_exact):
    T, cfg = engines(preset, n_streams)
    texts = ["Hello, world!", "A second, longer utterance for slot one.", "Third", "Numbers 1 2 3 4 5 6 7 8 9",
             "The quick brown fox jumps over the lazy dog."][:n_streams]
    toks = [pkg.synth.synth_tokens(cfg, t) for t in texts]
    spk = [pkg.synth.synth_speaker(cfg, seed=20 + i) for i in range(n_streams)]
    T.set_numerics(1, 4)      # batch-1 kernel in FAST numerics: the batched kernel's attention arithmetic
    T.set_batched(False)
    seq = run_streams(T, toks, spk, 12)
    T.set_batched(True)
    bat = run_streams(T, toks, spk, 12)
    T.set_numerics(0, 4)
    fd = []
    for i in range(n_streams):
        assert bat[i].shape[1] == 16 and len(bat[i]) > 0
        assert (bat[i] >= 0).all() and (bat[i][:, 1:] < cfg.cp_vocab_size).all()
        fd.append(first_diff(bat[i], seq[i]))
    # random-init weights give near-ties: one stream may flip early (the logit-level check with the
    # top-2 margin classification is test_batched_first_frame_logits_and_golden_ids)
    print("first differing frame per stream:", fd)
    assert sorted(fd)[1] >= min_exact and bat[0][0, 0] == seq[0][0, 0], fd


@pytest.mark.parametrize("preset,min_exact", [("tiny", 4), ("tiny-mtp", 4), ("0.6b", 2)])
def test_batched_first_frame_logits_and_golden_ids(engines, pkg, preset, min_exact):
    """The golden utterance in slot 1 of a 3-slot engine (other slots busy with other texts)."""
    T, cfg = engines(preset, 3)
    g = load_gold(preset)
    toks = [pkg.synth.synth_tokens(cfg, "Other text"), g["tokens"], pkg.synth.synth_tokens(cfg, "x y z w")]
    spk = [pkg.synth.synth_speaker(cfg, seed=3), g["speaker"], None]
    T.set_batched(True)
    for i in range(3):
        assert T.stream_begin(i, toks[i], spk[i], max_len=int(g["max_len"]), temperature=0.0), T.get_error()
    codes, produced, active = T.step_frames(1)
    cb0, cp = T.stream_logits(1)
    # frame 0: codebook-0 logits come from the (batch-1) prefill; the 15 code-predictor logit vectors
    # are the batched kernel's, free-running == teacher-forced while the ids agree
    ref_codes = g["codes"]
    assert codes[1, 0, 0] == ref_codes[0, 0]
    for s in range(15):
        d = np.abs(cp[s] - g["cp_logits"][s]).max()
        assert nmse(cp[s], g["cp_logits"][s]) <= 1e-5 and d <= 4e-3, (s, d)
        if codes[1, 0, s + 1] != ref_codes[0, s + 1]:
            top2 = np.sort(g["cp_logits"][s])[-2:]
            assert top2[1] - top2[0] <= 2 * d
            break
    rest = [codes[1, :produced[1]]]
    while active[1]:
        codes, produced, active = T.step_frames(8)
        rest.append(codes[1, :produced[1]])
    out = np.concatenate(rest)
    assert first_diff(out, ref_codes) >= min_exact


def test_batched_sampling_reproducible_and_ragged(engines, pkg):
    """temperature 0.9 / top-k 50 with per-stream seeds; ragged lengths; EOS by the ramp."""
    T, cfg = engines("tiny", 5)
    texts = ["Hi", "Hello, world!", "A somewhat longer sentence here.", "Hi", "Four"]
    toks = [pkg.synth.synth_tokens(cfg, t) for t in texts]
    spk = [None] * 5
    T.set_batched(True)
    # FAST numerics = exact top-k.  (REFERENCE numerics reproduce the reference's std::nth_element threshold, which on
    # these random-init logits often keeps a handful of candidates + EOS: utterances then end within a few frames,
    # as the seeded reference's do - tests/test_gpu_frames.py covers that mode against the reference.)
    T.set_numerics(1, 4)
    a = run_streams(T, toks, spk, 120, temperature=0.9, seeds=[1, 2, 3, 1, 5], chunk=16)
    b = run_streams(T, toks, spk, 120, temperature=0.9, seeds=[1, 2, 3, 1, 5], chunk=7)
    for i in range(5):
        assert np.array_equal(a[i], b[i])
        assert 0 < len(a[i]) <= 120
    assert np.array_equal(a[0], a[3])            # same text, same seed, different slot
    assert not np.array_equal(a[0][:8], a[4][:8])
    T.set_numerics(0, 4)


@pytest.mark.gpu
@pytest.mark.parametrize("n,probe", [(140, (0, 63, 127, 128, 139)), (128, (0, 64, 127)), (72, (0, 31, 71))])
def test_batched_large_group_matches_batch1(pkg, make_gguf, n, probe):
    """140: more than 128 streams in ONE lock-step group (two 128-row M tiles per GEMM, four in the merged
    code-predictor pass, split-M on the narrow projections).  128 / 72: one M tile, the O and down projections
    split over K with the second half's partial added by the following norm stage (two M tiles in the merged
    code-predictor pass).  First-frame logits of slots on both sides of the tile boundaries against the
    batch-1 kernel, and greedy ids of all streams over the first frames."""
    path, cfg = make_gguf("tiny", 1234)
    toks = [pkg.synth.synth_tokens(cfg, n_text=8 + (i % 5), seed=i) for i in range(n)]
    spks = [pkg.synth.synth_speaker(cfg, seed=i) for i in range(n)]
    T = pkg.TTSTransformer(device=0, max_streams=n, max_ctx=128)
    assert T.load_model(path), T.get_error()
    res = {}
    for batched in (True, False):
        assert T.set_batched(batched)
        for i in range(n):
            assert T.stream_begin(i, toks[i], spks[i], max_len=3, temperature=0.0, seed=i), T.get_error()
        codes, produced, active = T.step_frames(1)
        assert (produced == 1).all()
        logits = {i: T.stream_logits(i, want_cp=False)[0].copy() for i in probe}
        codes2, produced2, _ = T.step_frames(2)
        res[batched] = (codes[:, 0].copy(), logits, codes2[:, :2].copy())
    for i, l in res[True][1].items():
        ref = res[False][1][i]
        keep = np.isfinite(ref)
        assert nmse(l[keep], ref[keep]) <= 1e-5, i
    same0 = (res[True][0] == res[False][0]).all(axis=1)
    assert same0.mean() >= 0.95, f"only {same0.sum()} of {n} streams agree on the first frame"
    T.unload_model()


@pytest.mark.gpu
@pytest.mark.parametrize("preset,n", [("tiny", 7), ("tiny-mtp", 4), ("0.6b", 5), ("tiny", 140)])
def test_batched_prefill_matches_batch1_prefill(pkg, make_gguf, preset, n):
    """stream_prepare + prefill_pending (forward_prefill of all streams as lock-step passes of the batched kernel)
    against stream_begin (the batch-1 kernel's token-by-token prefill): codebook-0 logits of the first frame
    (they are the prefill's output) and greedy ids.  Ragged prefill lengths: every third stream has no speaker
    row (9 instead of 10 positions)."""
    path, cfg = make_gguf(preset, 1234)
    toks = [pkg.synth.synth_tokens(cfg, n_text=8 + (i % 5), seed=i) for i in range(n)]
    spks = [None if i % 3 == 2 else pkg.synth.synth_speaker(cfg, seed=i) for i in range(n)]
    T = pkg.TTSTransformer(device=0, max_streams=n, max_ctx=128)
    assert T.load_model(path), T.get_error()
    T.set_numerics(1, 4)
    res = {}
    for mode in ("batch1", "batched"):
        for i in range(n):
            if mode == "batch1":
                assert T.stream_begin(i, toks[i], spks[i], max_len=4, temperature=0.0, seed=i), T.get_error()
            else:
                assert T.stream_prepare(i, toks[i], spks[i], max_len=4, temperature=0.0, seed=i), T.get_error()
        if mode == "batched":
            assert T.prefill_pending(), T.get_error()
        codes, produced, active = T.step_frames(4)
        assert (produced == 4).all() and not active.any()
        res[mode] = (codes.copy(), [T.stream_logits(i, want_cp=False)[0].copy() for i in range(min(n, 8))])
    T.unload_model()
    ca, la = res["batch1"]
    cb, lb = res["batched"]
    for a, b in zip(la, lb):
        assert not np.isnan(b).any() and (np.isfinite(a) == np.isfinite(b)).all()
    # stream_logits holds the LAST frame's processed codebook-0 logits; compare where both runs took the same path
    same_first = (ca[:, 0, :] == cb[:, 0, :]).all(axis=1)
    print("streams equal on frame 0:", int(same_first.sum()), "of", n)
    assert same_first.mean() >= 0.8
    assert (ca[:, 0, 0] == cb[:, 0, 0]).mean() >= 0.9


@pytest.mark.gpu
def test_batched_prefill_first_frame_logits(pkg, make_gguf):
    """Prefill logits themselves: one frame only, so stream_logits is the prefill's codebook-0 logit vector after
    the suppress / repetition stage, equal processing on both paths."""
    path, cfg = make_gguf("0.6b", 1234)
    n = 3
    toks = [pkg.synth.synth_tokens(cfg, n_text=12 + 7 * i, seed=i) for i in range(n)]
    spks = [pkg.synth.synth_speaker(cfg, seed=i) for i in range(n)]
    T = pkg.TTSTransformer(device=0, max_streams=n, max_ctx=128)
    assert T.load_model(path), T.get_error()
    T.set_numerics(1, 4)
    out = {}
    for mode in ("batch1", "batched"):
        for i in range(n):
            fn = T.stream_begin if mode == "batch1" else T.stream_prepare
            assert fn(i, toks[i], spks[i], max_len=2, temperature=0.0, seed=i), T.get_error()
        codes, produced, active = T.step_frames(1)      # pending streams are prefilled implicitly
        assert (produced == 1).all()
        out[mode] = [T.stream_logits(i, want_cp=False)[0].copy() for i in range(n)]
    T.unload_model()
    for a, b in zip(out["batch1"], out["batched"]):
        fin = np.isfinite(a) & np.isfinite(b)
        assert (np.isfinite(a) == np.isfinite(b)).all()
        assert nmse(b[fin], a[fin]) <= 1e-5 and np.abs(a[fin] - b[fin]).max() <= 4e-3


@pytest.mark.gpu
def test_batched_prefill_joins_running_group(pkg, make_gguf):
    """Continuous batching: streams prepared while others of the same lock-step group are mid-utterance.  The
    running streams' ids must not change (the prefill works on private rows), the joiner's ids equal a solo run."""
    path, cfg = make_gguf("tiny", 1234)
    toks = [pkg.synth.synth_tokens(cfg, n_text=9 + i, seed=i) for i in range(4)]
    spks = [pkg.synth.synth_speaker(cfg, seed=i) for i in range(4)]
    T = pkg.TTSTransformer(device=0, max_streams=4, max_ctx=128)
    assert T.load_model(path), T.get_error()
    T.set_numerics(1, 4)

    def drain(got):
        while True:
            codes, produced, active = T.step_frames(5)
            for i in range(4):
                got[i].append(codes[i, :produced[i]])
            if not active.any():
                break

    # A: slots 0, 1 alone; B: slot 3 alone; C: 0, 1 start, slot 3 (+2) join after 6 frames
    runs = {}
    for name, first, late in (("A", (0, 1), ()), ("B", (3,), ()), ("C", (0, 1), (3, 2))):
        got = [[] for _ in range(4)]
        for i in first:
            assert T.stream_prepare(i, toks[i], spks[i], max_len=20, temperature=0.0, seed=i), T.get_error()
        assert T.prefill_pending(), T.get_error()
        codes, produced, active = T.step_frames(6)
        for i in range(4):
            got[i].append(codes[i, :produced[i]])
        for i in late:
            assert T.stream_prepare(i, toks[i], spks[i], max_len=20, temperature=0.0, seed=i), T.get_error()
        drain(got)
        runs[name] = [np.concatenate(g) for g in got]
    T.unload_model()
    for i in (0, 1):
        assert len(runs["A"][i]) == 20 and np.array_equal(runs["A"][i], runs["C"][i])
    assert len(runs["B"][3]) == 20 and np.array_equal(runs["B"][3], runs["C"][3])
    assert len(runs["C"][2]) == 20
