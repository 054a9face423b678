"""GPU box: stage timeline of the batch-1 frame kernel (qtts_trace) on the bench workload: where one CTA's thread 0
spends a frame, split into input poll / norm + staging / rows / barrier, per stage class."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import qtts_pkg
m = qtts_pkg.load()
preset = sys.argv[1] if len(sys.argv) > 1 else "0.6b"
cta = int(sys.argv[2]) if len(sys.argv) > 2 else 0
path = os.path.join(os.environ.get("QTTS_GGUF_CACHE", "/tmp"), f"synth-{preset}-1234.gguf")
if not os.path.exists(path):
    m.synth.write_gguf(path, preset, 1234)
cfg = m.synth.PRESETS[preset]
T = m.TTSTransformer(max_ctx=1024)
assert T.load_model(path), T.get_error()
T.set_numerics(1, 4); T.set_ignore_eos(True)
toks = m.synth.synth_tokens(cfg, n_text=200); spk = m.synth.synth_speaker(cfg)
T.generate(toks, spk, max_len=40, temperature=0.9, seed=1)
print(f"warm: {T.last_loop_ms() / 40:.3f} ms/frame")
T.trace(True, cta)
out = T.generate(toks, spk, max_len=40, temperature=0.9, seed=2)
print(f"traced run: {T.last_loop_ms() / 40:.3f} ms/frame (CTA {cta})")
tr = T.trace(False, cta)
tr = tr[tr[:, 1] != 0]
GHZ = 1.965
typ = tr[:, 0] & 255; opi = tr[:, 0] >> 8
d = lambda a, b: ((tr[:, a].astype(np.int64) - tr[:, b].astype(np.int64)) & 0xffffffff) / GHZ / 1e3
n_ops = int(opi.max()) + 1
frames = len(tr) / n_ops
print(f"{len(tr)} stages = {frames:.1f} frames of {n_ops} ops")
names = {0: "gemv", 1: "attn", 2: "sample_cb0", 3: "sample_cp", 4: "frame_end", 5: "copy"}
gap = np.zeros(len(tr)); gap[1:] = ((tr[1:, 1].astype(np.int64) - tr[:-1, 5].astype(np.int64)) & 0xffffffff) / GHZ / 1e3
tot = 0.0
for t in sorted(set(typ.tolist())):
    sel = typ == t
    n = sel.sum() / frames
    if t == 0:
        poll, prep, rows, bar = d(2, 1)[sel], d(3, 2)[sel], d(4, 3)[sel], d(5, 4)[sel]
        print(f"{names[t]:10s} n/frame={n:6.1f}  poll {poll.mean():5.2f}  norm+stage {prep.mean():5.2f}  rows {rows.mean():5.2f}  end/barrier {bar.mean():5.2f}  "
              f"gap-before {gap[sel].mean():5.2f}  total {(d(5, 1)[sel] + gap[sel]).mean():5.2f} us")
    else:
        print(f"{names[t]:10s} n/frame={n:6.1f}  stage {d(4, 1)[sel].mean():5.2f}  end/barrier {d(5, 4)[sel].mean():5.2f}  gap-before {gap[sel].mean():5.2f}  "
              f"total {(d(5, 1)[sel] + gap[sel]).mean():5.2f} us")
    tot += (d(5, 1)[sel] + gap[sel]).sum() / frames
print(f"sum {tot / 1e3:.3f} ms/frame")
# per GEMV op class (rows x K) of the frame program: index within the frame -> mean split
g = typ == 0
import collections
by = collections.defaultdict(list)
for k in np.nonzero(g)[0]:
    by[int(opi[k])].append(k)
rows = []
for o, ks in by.items():
    ks = np.array(ks)
    rows.append((o, d(2, 1)[ks].mean(), d(3, 2)[ks].mean(), d(4, 3)[ks].mean(), d(5, 4)[ks].mean()))
rows.sort()
print("first 40 GEMV ops of the frame (op index: poll, norm+stage, rows, end):")
for r in rows[:40]:
    print(f"  op {r[0]:4d}: {r[1]:5.2f} {r[2]:5.2f} {r[3]:5.2f} {r[4]:5.2f}")
print("last 12 GEMV ops of the frame:")
for r in rows[-12:]:
    print(f"  op {r[0]:4d}: {r[1]:5.2f} {r[2]:5.2f} {r[3]:5.2f} {r[4]:5.2f}")
