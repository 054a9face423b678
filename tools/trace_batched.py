"""GPU box: stage timeline of the batched frame kernel (qtts_trace, 16 store-only stamps per stage of one CTA).
slots: 1 stage start | 2 A producer: inputs ready (row counter) | 3 A producer: all chunk copies issued |
11 issuer 0: first weight piece landed | 4 issuer 0: first A chunk landed | 5 issuer 0: last commit issued |
6 epilogue: accumulator ready | 7 epilogue: stores issued | 8 stage function returned |
9 grid barrier: arrived | 10 grid barrier: released"""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import qtts_pkg
m = qtts_pkg.load()
preset = sys.argv[1] if len(sys.argv) > 1 else "0.6b"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
cta = int(sys.argv[3]) if len(sys.argv) > 3 else 0
path = os.path.join(os.environ.get("QTTS_GGUF_CACHE", "/tmp"), f"synth-{preset}-1234.gguf")
if not os.path.exists(path):
    m.synth.write_gguf(path, preset, 1234)
cfg = m.synth.PRESETS[preset]
T = m.TTSTransformer(max_streams=B, max_ctx=160)
assert T.load_model(path), T.get_error()
T.set_numerics(1, 4); T.set_ignore_eos(True)
def start():
    for i in range(B):
        assert T.stream_prepare(i, m.synth.synth_tokens(cfg, n_text=40 + i % 20, seed=i), m.synth.synth_speaker(cfg, seed=i), max_len=24, temperature=0.9, seed=i)
    assert T.prefill_pending()
start()
T.step_frames(24)
print(f"warm: {T.last_loop_ms() / 24:.3f} ms per frame-step")
start()
T.step_frames(8)
T.trace(True, cta)
T.step_frames(4)
print(f"traced: {T.last_loop_ms() / 4:.3f} ms per frame-step, CTA {cta}")
tr = T.trace(False, cta, max_stages=4096).reshape(-1, 16)
tr = tr[tr[:, 1] != 0]
GHZ = 1.965
d = lambda a, b: np.where((tr[:, a] != 0) & (tr[:, b] != 0), ((tr[:, a].astype(np.int64) - tr[:, b].astype(np.int64)) & 0xffffffff) / GHZ / 1e3, np.nan)
typ = tr[:, 0] & 15; f16 = (tr[:, 0] >> 4) & 1; flags = (tr[:, 0] >> 8) & 255
frames = 4.0
names = {0: "gemm", 1: "attn_cp", 2: "attn_tk", 3: "sample", 4: "norm"}
nxt = np.full(len(tr), np.nan); nxt[:-1] = ((tr[1:, 1].astype(np.int64) - tr[:-1, 8].astype(np.int64)) & 0xffffffff) / GHZ / 1e3
print(f"{len(tr)} stages recorded = {len(tr) / frames:.0f} per frame-step")
tot = 0.0
groups = collections.OrderedDict()
for k in range(len(tr)):
    key = (names[int(typ[k])], int(tr[k, 12]) if typ[k] in (0, 4) else 0, int(tr[k, 13]) if typ[k] == 0 else 0, bool(flags[k] & 8), bool(tr[k, 2] != 0))
    groups.setdefault(key, []).append(k)
print("class (K, rows, waits-for-row-counter, this CTA has a tile): n/frame | stage | ->next stage start (barrier + loop) | inside a GEMM stage: inputs ready, copies issued, weights, first A chunk, MMAs issued, acc ready, stores")
for key, ks in groups.items():
    ks = np.array(ks)
    st = np.nanmean(d(8, 1)[ks]); nx = np.nanmean(nxt[ks])
    line = f"{key[0]:8s} K={key[1]:5d} rows={key[2]:5d} wait={int(key[3])} tile={int(key[4])}: n={len(ks) / frames:6.1f} stage {st:6.2f} next {nx:6.2f}"
    if key[0] == "gemm" and key[4]:
        line += (f" | in-ready {np.nanmean(d(2, 1)[ks]):5.2f} copies {np.nanmean(d(3, 1)[ks]):5.2f} w {np.nanmean(d(11, 1)[ks]):5.2f} A0 {np.nanmean(d(4, 1)[ks]):5.2f}"
                 f" mma {np.nanmean(d(5, 1)[ks]):5.2f} acc {np.nanmean(d(6, 1)[ks]):5.2f} stored {np.nanmean(d(7, 1)[ks]):5.2f}")
    bar = d(10, 8)[ks]; arr = d(9, 8)[ks]; syn = d(15, 8)[ks]
    if np.isfinite(bar).any():
        line += f" | barrier: cta-sync {np.nanmean(syn):5.2f} arrive {np.nanmean(arr):5.2f} release {np.nanmean(bar):5.2f}"
    print(line)
    tot += np.nansum(d(8, 1)[ks] + nxt[ks]) / frames
print(f"sum {tot / 1e3:.3f} ms per frame-step")
