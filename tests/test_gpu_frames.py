"""GPU parity of the FRAME LOOP itself, kernel by kernel and numerics mode by numerics mode, against the
UNMODIFIED reference (round-1 verdict: the benchmarked paths — FAST numerics, the batched tcgen05 kernel's
Talker step, >= 512 keys, the 1.7B preset in the batched kernel, the EOS ramp and temperature sampling —
had no direct oracle comparison).

Method: teacher forcing at frame level (qtts_stream_force).  The engine runs its own frame loop, but the
codes fed back after every frame are the reference's, so frame f sees exactly the reference's history and
its logits are comparable with the reference's frame f for EVERY frame, not only until the first near-tie
flips an id.  What is compared per frame:
  * processed codebook-0 logits (qtts_stream_logits) vs restate.cb0_postprocess(reference raw logits)
    — restate.cb0_postprocess is pinned bit-exact to the reference's processed logits in test_oracle_frames.py;
  * the 15 code-predictor logit vectors vs the reference's;
  * the engine's own choice of every code vs the reference's, wherever the reference's top-2 margin
    exceeds 2 * max|d| (the classification the teacher-forced tests of round 1 used).
Reference values: tests/golden/ref_<preset>_seed1234_frames.npz (first frames, committed) and, when
oracle/_ref/libqtts_ref.so travelled with the snapshot, the reference run live for all frames.

Tolerances (stated): REFERENCE numerics NMSE <= 1e-5, max|d| <= 4e-3 * sqrt(H / 1024) on Talker logits;
FAST numerics (FP32 P.V accumulation instead of the reference's sequential F16 accumulator) NMSE <= 2e-5,
max|d| <= 6e-3 * sqrt(H / 1024).
"""
import os

import numpy as np
import pytest

from conftest import nmse
from test_oracle import load_gold

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = {0: (1e-5, 4e-3), 1: (2e-5, 6e-3)}         # numerics mode -> (NMSE, max|d| at H = 1024)
EOS, V = 2150, 3072


def load_frames(preset):
    p = os.path.join(GOLD, f"ref_{preset}_seed1234_frames.npz")
    if not os.path.exists(p):
        pytest.skip(f"{p} missing (tests/golden/make_golden.py frames)")
    return np.load(p)


_TRACES = {}


def live_trace(ref_available, path, toks, spk, n, temperature=0.0, seed=None):
    """The reference's per-frame intermediates, live (None when oracle/_ref did not travel); cached per input."""
    if not ref_available:
        return None
    key = (path, tuple(int(t) for t in toks), None if spk is None else spk.tobytes(), n, temperature, seed)
    if key not in _TRACES:
        _TRACES[key] = _live_trace(path, toks, spk, n, temperature, seed)
    return _TRACES[key]


def _live_trace(path, toks, spk, n, temperature, seed):
    import ref
    R = ref.Reference(path)
    if seed is not None:
        R.seed_rng(seed)
    tr = R.generate_trace(toks, spk, max_len=n, temperature=temperature, top_k=50)
    cpl = {}
    if temperature <= 0.0:
        for f in range(min(tr["n_frames"], 4)):
            _, cpl[f] = R.predict_codes_logits(tr["hidden"][f], int(tr["codes"][f, 0]))
    tr["cp_logits"] = cpl
    R.close()
    return tr


@pytest.fixture(scope="module")
def ref_available():
    import ref
    return ref.available()


def finite_cmp(a, b):
    """(nmse, max|d|) over the entries finite in both; the -inf patterns must be identical."""
    fa, fb = np.isfinite(a), np.isfinite(b)
    assert np.array_equal(fa, fb), "suppressed-id pattern differs"
    return nmse(a[fa], b[fb]), float(np.abs(a[fa] - b[fb]).max())


def check_forced_run(T, pkg, cfg, slot, toks, spk, ref_codes, ref_raw, ref_cp, mode, n_frames, n_text, batched,
                     prepare=False, label=""):
    """Runs `slot` teacher-forced over ref_codes and compares frame by frame.  Returns per-frame stats."""
    import restate
    nm_tol, abs_tol = TOL[mode]
    abs_tol *= max(1.0, (cfg.hidden_size / 1024.0) ** 0.5)
    # code-predictor logits: 4e-3 in the batch-1 kernel; 6e-3 in the batched kernel, whose tensor-core accumulation
    # order (K split over issuers / M tiles) moved single outliers of 2048 logits to 4.4e-3 at NMSE 2e-6 (measured)
    cp_abs = (6e-3 if batched else 4e-3) * max(1.0, (cfg.hidden_size / 1024.0) ** 0.5)
    T.set_batched(batched)
    ok = (T.stream_prepare if prepare else T.stream_begin)(slot, toks, spk, max_len=n_frames, temperature=0.0)
    assert ok, T.get_error()
    if prepare:
        assert T.prefill_pending(), T.get_error()
    assert T.stream_force(slot, forced_codes=ref_codes[:n_frames]), T.get_error()
    seen = set()
    worst = (0.0, 0.0)
    n_cmp = exact = 0
    for f in range(n_frames):
        out = T.step_frames(1)
        assert out is not None, T.get_error()
        codes, produced, active = out
        assert produced[slot] == 1, (label, f, produced[slot])
        cb0, cp = T.stream_logits(slot)
        own = codes[slot, 0]
        if f < ref_raw.shape[0]:
            want = restate.cb0_postprocess(ref_raw[f], seen, f, n_text, 1.05, EOS, V)
            e, d = finite_cmp(cb0, want)
            assert e <= nm_tol and d <= abs_tol, f"{label} frame {f}: codebook-0 logits NMSE {e:.2e} max|d| {d:.2e}"
            worst = (max(worst[0], e), max(worst[1], d))
            n_cmp += 1
            top2 = np.sort(want[np.isfinite(want)])[-2:]
            if top2[1] - top2[0] > 2 * d:
                assert own[0] == ref_codes[f, 0], f"{label} frame {f}: codebook 0 {own[0]} != {ref_codes[f, 0]}"
        if f in ref_cp:
            for s in range(15):
                e = nmse(cp[s], ref_cp[f][s])
                d = float(np.abs(cp[s] - ref_cp[f][s]).max())
                assert e <= 1e-5 and d <= cp_abs, f"{label} frame {f} code {s + 1}: NMSE {e:.2e} max|d| {d:.2e}"
                top2 = np.sort(ref_cp[f][s])[-2:]
                if top2[1] - top2[0] > 2 * d:
                    assert own[s + 1] == ref_codes[f, s + 1], f"{label} frame {f} code {s + 1}"
        exact += int(np.array_equal(own, ref_codes[f]))
        seen.add(int(ref_codes[f, 0]))
    print(f"{label}: {n_cmp} frames of logits compared, worst NMSE {worst[0]:.2e} max|d| {worst[1]:.2e}; "
          f"{exact}/{n_frames} frames with all 16 own ids equal to the reference's")
    return worst, exact


def reference_frames(preset, ref_available, path, toks, spk, n_frames):
    """(codes, raw logits [rows, V], {frame: cp logits}) from the live reference if present, else the fixture."""
    g, gf = load_gold(preset), load_frames(preset)
    cp = {0: g["cp_logits"], 1: gf["g_cp_logits"][0], 2: gf["g_cp_logits"][1]}
    tr = live_trace(ref_available, path, toks, spk, n_frames)
    if tr is not None:
        assert np.array_equal(tr["codes"][:len(gf["g_codes"])], gf["g_codes"][:tr["n_frames"]]), \
            "live reference disagrees with its committed fixture"
        k = gf["g_raw"].shape[0]
        assert np.abs(tr["raw"][:k] - gf["g_raw"]).max() == 0.0, "live reference logits differ from the fixture"
        cp.update(tr["cp_logits"])
        return tr["codes"], tr["raw"], cp
    return gf["g_codes"], gf["g_raw"], cp


@pytest.mark.parametrize("preset", ["tiny", "tiny-mtp", "0.6b", "1.7b"])
@pytest.mark.parametrize("mode", [0, 1], ids=["reference", "fast"])
def test_batch1_kernel_frames_vs_reference(pkg, make_gguf, ref_available, preset, mode):
    """qtts_run_program (the kernel bench.py's headline runs, FAST numerics included): every golden frame."""
    path, cfg = make_gguf(preset, 1234)
    g = load_gold(preset)
    n = int(g["max_len"])
    codes, raw, cp = reference_frames(preset, ref_available, path, g["tokens"], g["speaker"], n)
    T = pkg.TTSTransformer(device=0, max_streams=1, max_ctx=128)
    assert T.load_model(path), T.get_error()
    T.set_numerics(mode, int(g["threads"]))
    T.set_debug(1)
    n = min(n, len(codes))
    worst, exact = check_forced_run(T, pkg, cfg, 0, g["tokens"], g["speaker"], codes, raw, cp, mode, n, len(g["tokens"]),
                                    batched=False, label=f"batch-1 {preset} mode {mode}")
    T.unload_model()
    assert exact >= n - 2      # near-ties may flip an own id in at most two frames (each one is margin-classified above)


@pytest.mark.parametrize("preset,n_slots,slot", [("tiny", 3, 1), ("tiny-mtp", 5, 2), ("0.6b", 64, 33), ("0.6b", 140, 131),
                                                 ("0.6b", 256, 200), ("1.7b", 64, 5), ("1.7b", 140, 70), ("1.7b", 256, 255)])
def test_batched_kernel_frames_vs_reference(pkg, make_gguf, ref_available, preset, n_slots, slot):
    """qtts_run_batched: the golden utterance in one slot of a FULL lock-step group (all other slots busy with
    other utterances), M = 64 / two M tiles / 256 rows; Talker step + code predictor of every golden frame."""
    path, cfg = make_gguf(preset, 1234)
    g = load_gold(preset)
    n = min(int(g["max_len"]), 16 if n_slots > 5 else 48)
    codes, raw, cp = reference_frames(preset, ref_available, path, g["tokens"], g["speaker"], n)
    n = min(n, len(codes))
    T = pkg.TTSTransformer(device=0, max_streams=n_slots, max_ctx=96)
    assert T.load_model(path), T.get_error()
    T.set_numerics(1, 4)
    T.set_batched(True)
    for i in range(n_slots):
        if i == slot:
            continue
        t = pkg.synth.synth_tokens(cfg, n_text=8 + (i % 23), seed=100 + i)
        s = None if i % 4 == 3 else pkg.synth.synth_speaker(cfg, seed=100 + i)
        assert T.stream_prepare(i, t, s, max_len=n + 4, temperature=0.0, seed=i), T.get_error()
    worst, exact = check_forced_run(T, pkg, cfg, slot, g["tokens"], g["speaker"], codes, raw, cp, 1, n, len(g["tokens"]),
                                    batched=True, prepare=True, label=f"batched {preset} {n_slots} slots")
    T.unload_model()
    assert exact >= n - 2


@pytest.mark.parametrize("kernel", ["batch1-reference", "batch1-fast", "batched-3", "batched-140"])
def test_long_context_frames_vs_live_reference(pkg, make_gguf, ref_available, kernel):
    """>= 512 keys, where the reference's CPU flash attention splits the keys over its 4 threads
    (ggml/src/ggml-cpu/ops.cpp:8740-8772), the batch-1 FAST path merges split-KV partials across CTAs and the
    batched kernel splits keys over warps: 640 frames of one utterance, teacher-forced on the reference's codes."""
    if not ref_available:
        pytest.skip("needs the compiled reference (oracle/_ref) for a 640-frame trace")
    path, cfg = make_gguf("tiny", 1234)
    toks = pkg.synth.synth_tokens(cfg, n_text=200, seed=77)        # EOS ramp would start at frame 832
    spk = pkg.synth.synth_speaker(cfg, seed=77)
    n = 640
    tr = live_trace(True, path, toks, spk, n)
    if tr["n_frames"] < 600:
        pytest.skip(f"reference stopped by EOS at frame {tr['eos_frame']}")
    n = tr["n_frames"]
    batched = kernel.startswith("batched")
    n_slots = int(kernel.split("-")[1]) if batched else 1
    mode = 0 if kernel.endswith("reference") else 1
    T = pkg.TTSTransformer(device=0, max_streams=n_slots, max_ctx=n + 32)
    assert T.load_model(path), T.get_error()
    T.set_numerics(mode, 4)
    T.set_batched(batched)
    slot = n_slots // 2
    for i in range(n_slots):
        if i != slot:
            assert T.stream_prepare(i, pkg.synth.synth_tokens(cfg, n_text=150 + i, seed=i), None, max_len=n, temperature=0.0), T.get_error()
    ok = (T.stream_prepare if batched else T.stream_begin)(slot, toks, spk, max_len=n, temperature=0.0)
    assert ok, T.get_error()
    if batched:
        assert T.prefill_pending(), T.get_error()
    assert T.stream_force(slot, forced_codes=tr["codes"]), T.get_error()
    import restate
    nm_tol, abs_tol = TOL[mode]
    check = set(range(0, 4)) | set(range(126, 132)) | set(range(498, 516)) | set(range(n - 24, n))
    seen, done, worst = set(), 0, (0.0, 0.0, -1)
    while done < n:
        nxt = min(f for f in check | {n} if f >= done)
        if nxt > done:                                       # run up to the next checked frame in one call
            out = T.step_frames(min(nxt - done, 64))
            assert out is not None, T.get_error()
            k = int(out[1][slot])
            assert k > 0
            for f in range(done, done + k):
                seen.add(int(tr["codes"][f, 0]))
            done += k
            continue
        out = T.step_frames(1)
        assert out is not None and out[1][slot] == 1, T.get_error()
        cb0, _ = T.stream_logits(slot, want_cp=False)
        want = restate.cb0_postprocess(tr["raw"][done], seen, done, len(toks), 1.05, EOS, V)
        e, d = finite_cmp(cb0, want)
        if e > worst[0]:
            worst = (e, d, done)
        assert e <= 4 * nm_tol and d <= 2 * abs_tol, f"{kernel} frame {done} ({done + 10} keys): NMSE {e:.2e} max|d| {d:.2e}"
        seen.add(int(tr["codes"][done, 0]))
        done += 1
    print(f"{kernel}: worst NMSE {worst[0]:.2e} (max|d| {worst[1]:.2e}) at frame {worst[2]}")
    T.unload_model()


@pytest.mark.parametrize("preset", ["tiny", "0.6b"])
@pytest.mark.parametrize("kernel", ["batch1-reference", "batch1-fast", "batched"])
def test_eos_ramp_and_stop_frame(pkg, make_gguf, preset, kernel):
    """"Hi" (10 tokens): the EOS ramp starts at frame 40 (src/tts_transformer.cpp:2439-2445) and ends the
    utterance.  Teacher-forced on the reference's codes, the engine must (1) produce the reference's processed
    logits around and inside the ramp, (2) choose EOS itself at exactly the reference's stop frame."""
    import restate
    path, cfg = make_gguf(preset, 1234)
    gf = load_frames(preset)
    toks, codes, stop = gf["e_tokens"], gf["e_codes"], int(gf["e_eos_frame"])
    assert stop == len(codes) and stop > 40
    batched = kernel == "batched"
    mode = 0 if kernel.endswith("reference") else 1
    T = pkg.TTSTransformer(device=0, max_streams=2 if batched else 1, max_ctx=stop + 64)
    assert T.load_model(path), T.get_error()
    T.set_numerics(mode, 4)
    T.set_batched(batched)
    if batched:
        assert T.stream_prepare(1, pkg.synth.synth_tokens(cfg, "Another one"), None, max_len=stop + 8, temperature=0.0)
    assert T.stream_begin(0, toks, None, max_len=stop + 8, temperature=0.0), T.get_error()
    assert T.stream_force(0, forced_codes=codes), T.get_error()
    nm_tol, abs_tol = TOL[mode]
    frames = {int(f): i for i, f in enumerate(gf["e_frames"])}
    own_all = []
    for f in range(stop + 1):
        out = T.step_frames(1)
        assert out is not None, T.get_error()
        cdz, produced, active = out
        if f in frames:
            cb0, _ = T.stream_logits(0, want_cp=False)
            e, d = finite_cmp(cb0, gf["e_proc"][frames[f]])
            assert e <= nm_tol and d <= abs_tol, f"{kernel} {preset} frame {f}: processed logits NMSE {e:.2e} max|d| {d:.2e}"
            # and the restatement of the post-processing on the reference's raw logits is what the reference computed
            want = restate.cb0_postprocess(gf["e_raw"][frames[f]], set(int(c) for c in codes[:f, 0]), f, len(toks), 1.05, EOS, V)
            assert np.array_equal(want, gf["e_proc"][frames[f]])
        if f < stop:
            assert produced[0] == 1 and active[0] == 1, f"{kernel} {preset}: stopped at frame {f}, the reference at {stop}"
            own_all.append(cdz[0, 0])
        else:
            assert produced[0] == 0 and active[0] == 0, f"{kernel} {preset}: no EOS at the reference's stop frame {stop}"
    own = np.stack(own_all)
    print(f"{kernel} {preset}: EOS at frame {stop} like the reference; {(own == codes).all(axis=1).sum()}/{stop} frames with all own ids equal")
    T.unload_model()


@pytest.mark.parametrize("preset", ["tiny", "0.6b"])
@pytest.mark.parametrize("batched", [False, True], ids=["batch1", "batched"])
def test_sampled_frames_vs_seeded_reference(pkg, make_gguf, preset, batched):
    """temperature 0.9 / top-k 50 against the reference with its mt19937 re-seeded (oracle harness): the engine gets
    the uniforms the reference's sampler consumed (qtts_stream_force) and is teacher-forced on the reference's codes.
    REFERENCE numerics: codebook 0 uses the reference's std::nth_element threshold.  Every own id must be what
    the samplers of the reference give on the ENGINE's logits with the same uniform (exact), and must equal the
    reference's id in the large majority of draws (a draw within the logit tolerance of a CDF step or of the
    top-k threshold may land on the neighbouring id)."""
    import restate
    import ref
    path, cfg = make_gguf(preset, 1234)
    g, gf = load_gold(preset), load_frames(preset)
    codes, u = gf["s_codes"], gf["s_u"].astype(np.float32)
    n = len(codes)
    assert n >= 8
    T = pkg.TTSTransformer(device=0, max_streams=2 if batched else 1, max_ctx=128)
    assert T.load_model(path), T.get_error()
    T.set_numerics(0, 4)
    T.set_debug(1)
    T.set_batched(batched)
    if batched:
        assert T.stream_prepare(1, pkg.synth.synth_tokens(cfg, "Another one"), None, max_len=n, temperature=0.9, seed=4)
    assert T.stream_begin(0, g["tokens"], g["speaker"], max_len=n, temperature=0.9, top_k=50), T.get_error()
    assert T.stream_force(0, forced_codes=codes, uniforms=u[:n]), T.get_error()
    same = total = 0
    for f in range(n):
        out = T.step_frames(1)
        assert out is not None and out[1][0] == 1, T.get_error()
        own = out[0][0, 0]
        cb0, cp = T.stream_logits(0)
        if ref.available():
            tok, kept = ref.sample_cb0(cb0, EOS, 0.9, 50, float(u[f, 0]))
            assert own[0] == tok, f"frame {f}: codebook 0 {own[0]} but the reference's sampler gives {tok} on these logits (kept {kept})"
        if f < gf["s_proc"].shape[0]:
            e, d = finite_cmp(cb0, gf["s_proc"][f])
            assert e <= 1e-5 and d <= 4e-3
        for s in range(15):
            assert own[s + 1] == restate.sample_topk(cp[s], 0.9, 50, float(u[f, s + 1])), (f, s)
        same += int((own == codes[f]).sum())
        total += 16
    print(f"sampled {preset} {'batched' if batched else 'batch-1'}: {same}/{total} own ids equal the seeded reference's")
    T.unload_model()
    assert same >= 0.9 * total


def test_sample_cb0_entry_vs_reference_sampler(pkg, make_gguf):
    """qtts_sample_cb0 (the inline block of generate(), src/tts_transformer.cpp:2417-2499) on random logits:
    processed logits bit-equal to restate.cb0_postprocess; sampled id equal to the reference's sampler
    (std::nth_element threshold) in REFERENCE numerics and to the exact top-k restatement in FAST numerics;
    u = 0 selects index 0 (:2490-2494)."""
    import restate
    import ref
    path, cfg = make_gguf("tiny", 1234)
    T = pkg.TTSTransformer(device=0, max_streams=1, max_ctx=256)
    assert T.load_model(path), T.get_error()
    r = np.random.default_rng(5)
    for trial in range(24):
        raw = (r.standard_normal(V) * (0.3 + 0.2 * (trial % 4))).astype(np.float32)
        seen = set(int(x) for x in r.integers(0, 2048, size=trial))
        frame = [0, 39, 40, 41, 60, 79, 80, 200][trial % 8]
        n_text = 10
        u = float(r.random()) if trial % 6 else 0.0
        temp, top_k = (0.9, 50) if trial % 3 else (1.3, 7)
        want = restate.cb0_postprocess(raw, seen, frame, n_text, 1.05, EOS, V)
        for mode in (0, 1):
            T.set_numerics(mode, 4)
            tok, proc = T.sample_cb0(raw, frame, n_text, seen, 1.05, temp, top_k, u)
            assert np.array_equal(proc, want), (trial, mode)
            if mode == 1:
                assert tok == restate.sample_topk(want, temp, top_k, u, eos_id=EOS), (trial, "fast")
            elif ref.available():
                assert tok == ref.sample_cb0(want, EOS, temp, top_k, u)[0], (trial, "reference")
            if u == 0.0:
                assert tok == 0
        tok, _ = T.sample_cb0(raw, frame, n_text, seen, 1.05, 0.0, 0, 0.5)
        assert tok == restate.argmax_first(want)
    T.unload_model()
