"""GPU box: batch-1 frame loop time of the library QTTS_LIB_PATH names, on the bench workload (0.6B, 208 text tokens,
512 frames, FAST numerics): median of 5 generate() calls.  usage: QTTS_LIB_PATH=... python tools/ab_b1.py [preset] [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import qtts_pkg
m = qtts_pkg.load()
preset = sys.argv[1] if len(sys.argv) > 1 else "0.6b"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 512
path = os.path.join(os.environ.get("QTTS_GGUF_CACHE", "/tmp"), f"synth-{preset}-1234.gguf")
if not os.path.exists(path):
    m.synth.write_gguf(path, preset, 1234)
cfg = m.synth.PRESETS[preset]
T = m.TTSTransformer(max_ctx=1024)
assert T.load_model(path), T.get_error()
T.set_numerics(1, 4); T.set_ignore_eos(True)
toks = m.synth.synth_tokens(cfg, n_text=208); spk = m.synth.synth_speaker(cfg)
ms = []
for i in range(6):
    out = T.generate(toks, spk, max_len=frames, temperature=0.9, seed=1 + i)
    ms.append(T.last_loop_ms() / len(out))
print(f"{os.environ.get('QTTS_LIB_PATH', 'in-tree')}: {preset} {frames} frames: median {np.median(ms[1:]):.4f} ms/frame  all {[round(x, 4) for x in ms]}", flush=True)
