"""Small run of both frame kernels for compute-sanitizer (memcheck / racecheck / synccheck), tiny preset:
batch-1 kernel persistent (flag-carrying hand-off) and stage-wise, batched tcgen05 kernel with 3 streams
(batched prefill + frame-steps), the sampling path (temperature 0.9) and the vocoder hand-off chunk."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import qtts_pkg
m = qtts_pkg.load()
preset = "tiny"
path = os.path.join(os.environ.get("QTTS_GGUF_CACHE", "/tmp"), f"synth-{preset}-1234.gguf")
if not os.path.exists(path):
    m.synth.write_gguf(path, preset, 1234)
cfg = m.synth.PRESETS[preset]
g = np.load(os.path.join(ROOT, "tests", "golden", "ref_tiny_seed1234.npz"))
T = m.TTSTransformer(max_streams=3, max_ctx=96)
assert T.load_model(path), T.get_error()
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 4
for numerics in (0, 1):
    T.set_numerics(numerics, 4)
    a = T.generate(g["tokens"], g["speaker"], max_len=frames, temperature=0.0)
    assert a is not None, T.get_error()
    print(f"persistent numerics {numerics}: {len(a)} frames, equal to golden: {np.array_equal(a, g['codes'][:len(a)])}", flush=True)
T.set_persistent(False)
b = T.generate(g["tokens"], g["speaker"], max_len=2, temperature=0.0)
T.set_persistent(True)
print("stage-wise:", len(b), "frames", flush=True)
c = T.generate(g["tokens"], g["speaker"], max_len=frames, temperature=0.9, seed=3)
print("sampled:", len(c), "frames", flush=True)
for i in range(3):
    assert T.stream_prepare(i, m.synth.synth_tokens(cfg, n_text=8 + i, seed=i), m.synth.synth_speaker(cfg, seed=i), max_len=frames, temperature=0.9 if i else 0.0, seed=i)
codes, produced, active = T.step_frames(frames)
print("batched:", produced.tolist(), flush=True)
ch = T.codes_chunk(0, 0, int(produced[0]), 40)
assert ch is not None and np.array_equal(ch[0, :, :produced[0]], codes[0, :produced[0]].T)
T.unload_model()
print("SANITIZE_RUN_OK", flush=True)
