"""GPU box: per-stage cycle profile of the frame kernel (CTA 0's view) on the bench workload."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import qtts_pkg
m = qtts_pkg.load()
preset = sys.argv[1] if len(sys.argv) > 1 else "0.6b"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 120
numerics = int(sys.argv[3]) if len(sys.argv) > 3 else 1
path = os.path.join(os.environ.get("QTTS_GGUF_CACHE", "/tmp"), f"synth-{preset}-1234.gguf")
if not os.path.exists(path):
    m.synth.write_gguf(path, preset, 1234)
cfg = m.synth.PRESETS[preset]
T = m.TTSTransformer(max_ctx=1024)
assert T.load_model(path), T.get_error()
T.set_numerics(numerics, 4); T.set_ignore_eos(True)
toks = m.synth.synth_tokens(cfg, n_text=200); spk = m.synth.synth_speaker(cfg)
T.generate(toks, spk, max_len=40, temperature=0.9, seed=1)
print(f"warm: {T.last_loop_ms()/40:.3f} ms/frame")
cta = int(sys.argv[4]) if len(sys.argv) > 4 else 0
T.profile(1 + cta)
out = T.generate(toks, spk, max_len=frames, temperature=0.9, seed=2)
ms = T.last_loop_ms()
p = T.profile(0)
print("recording CTA", cta)
print(f"profiled run: {len(out)} frames, {ms/len(out):.3f} ms/frame")
GHZ = 1.965
tot = 0


per = p.pop('per_cta')[:148].astype(float) / GHZ / 1e3 / len(out)
import numpy as np
for j, nm in enumerate(['gemv stage', 'gemv barrier', 'attn stage', 'attn barrier']):
    c = per[:, j]
    print(f'per-CTA {nm:12s} us/frame: min {c.min():8.1f} (cta {c.argmin()}) median {np.median(c):8.1f} max {c.max():8.1f} (cta {c.argmax()})  top5 {np.argsort(-c)[:5].tolist()}')
for k, v in p.items():
    if v[0] == 0: continue
    n = v[0]
    print(f"{k:11s} n/frame={n/len(out):7.1f} stage={v[1]/n/GHZ/1e3:6.2f}us barrier={v[2]/n/GHZ/1e3:6.2f}us (cta-sync {v[7]/n/GHZ/1e3:5.2f}, prefetch-issue {v[3]/n/GHZ/1e3:5.2f})"
          + (f" | load={v[3]/n/GHZ/1e3:5.2f} norm={v[4]/n/GHZ/1e3:5.2f} rows={v[5]/n/GHZ/1e3:5.2f} (wait {v[6]/n/GHZ/1e3:5.2f})" if k == "gemv" else ""))
    tot += (v[1] + v[2])
print(f"sum = {tot/len(out)/GHZ/1e6:.3f} ms/frame (cycles at {GHZ} GHz)")
