"""Dev benchmark of the batched frame kernel: B concurrent streams on one GPU.
usage: python tools/bench_batched.py [preset] [streams] [frames] [n_text] [profile]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import qtts_pkg
pkg = qtts_pkg.load()
preset = sys.argv[1] if len(sys.argv) > 1 else "0.6b"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 80
n_text = int(sys.argv[4]) if len(sys.argv) > 4 else 64
prof = int(sys.argv[5]) if len(sys.argv) > 5 else 0
import tempfile
path = os.path.join(os.environ.get("QTTS_GGUF_CACHE") or tempfile.gettempdir(), f"synth-{preset}-1234.gguf")
if not os.path.exists(path):
    pkg.synth.write_gguf(path, preset, 1234)
cfg = pkg.synth.PRESETS[preset]
T = pkg.TTSTransformer(device=0, max_streams=B, max_ctx=frames + 64 + 40)
t0 = time.time()
assert T.load_model(path), T.get_error()
print(f"load {time.time() - t0:.1f}s", flush=True)
T.set_numerics(pkg.NUMERICS_FAST, 4)
T.set_ignore_eos(True)
r = np.random.default_rng(0)
for rep in range(2):
    t0 = time.time()
    for i in range(B):
        toks = pkg.synth.synth_tokens(cfg, n_text=int(r.integers(max(8, n_text // 2), n_text + 1)), seed=i)
        fn = T.stream_begin if os.environ.get("QTTS_SEQ_PREFILL") else T.stream_prepare
        assert fn(i, toks, pkg.synth.synth_speaker(cfg, seed=i), max_len=frames, temperature=0.9, top_k=50, seed=i), T.get_error()
    t_prep = time.time() - t0
    assert T.prefill_pending(), T.get_error()
    print(f"  prepare {t_prep / B * 1e3:.2f} ms/stream, batched prefill {T.last_loop_ms():.1f} ms device for {B} streams", flush=True)
    t_begin = time.time() - t0
    if prof and rep == 1:
        T.batched_profile(True)
    tot = 0; ms = 0.0
    t0 = time.time()
    while True:
        codes, produced, active = T.step_frames(40)
        tot += int(produced.sum()); ms += T.last_loop_ms()
        if not active.any():
            break
    wall = time.time() - t0
    print(f"rep {rep}: B={B} frames={tot} device {ms:.1f} ms -> {tot / ms * 1e3:.0f} frames/s, {ms / (tot / B):.3f} ms/frame-step; "
          f"wall {tot / wall:.0f} frames/s; stream_begin {t_begin / B * 1e3:.1f} ms/stream", flush=True)
if prof:
    p = T.batched_profile(False)
    clk = 1.965e3  # cycles per us at max clock (approx)
    for k, v in p.items():
        if v["count"]:
            print(f"  {k:12s} n={v['count']:6d} stage {v['stage_cycles'] / v['count'] / clk:7.2f} us  barrier {v['barrier_cycles'] / v['count'] / clk:6.2f} us"
                  f"  total {(v['stage_cycles'] + v['barrier_cycles']) / clk / 1e3:8.2f} ms")
