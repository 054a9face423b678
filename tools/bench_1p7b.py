"""BASELINE configs[4]: the "1.7B architecture" (synth preset 1.7b: Talker H 2048 / FFN 6144, code predictor of
0.6B size behind mtp_proj; dims assumed, SURVEY §8d "Config 5"), random-init, on one B200:
  * batch 1: 256 frames through generate(), device ms/frame and HBM roofline fraction;
  * batch B (default 128) with max_len frames (default 4096, the KV-cache-bound regime): device ms per
    frame-step per 40-frame chunk, reported against n_kv so the KV term of the roofline is visible.
usage: python tools/bench_1p7b.py [B] [max_len] [out.json]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import qtts_pkg
pkg = qtts_pkg.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
max_len = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
out_path = sys.argv[3] if len(sys.argv) > 3 else None
preset = os.environ.get("QTTS_PRESET", "1.7b")
import tempfile
path = os.path.join(os.environ.get("QTTS_GGUF_CACHE") or tempfile.gettempdir(), f"synth-{preset}-1234.gguf")
if not os.path.exists(path):
    t0 = time.time(); pkg.synth.write_gguf(path, preset, 1234); print(f"gguf written in {time.time() - t0:.0f}s", flush=True)
cfg = pkg.synth.PRESETS[preset]
peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
hbm = float(peaks.get("hbm_gbs", 6560.3))
res = {"preset": preset, "hbm_peak_gbs": hbm}

def weight_bytes(c):
    H, F, cH, cF = c.hidden_size, c.intermediate_size, c.cp_hidden_size, c.cp_intermediate_size
    qkv = (c.n_heads + 2 * c.n_kv_heads) * 128
    talker = c.n_layers * (qkv * H + H * c.n_heads * 128 + 3 * F * H) + c.codec_vocab_size * H
    cqkv = (c.cp_n_heads + 2 * c.cp_n_kv_heads) * 128
    cp = c.cp_layers * (cqkv * cH + cH * c.cp_n_heads * 128 + 3 * cF * cH)
    heads = 15 * c.cp_vocab_size * cH
    mtp = 16 * cH * H if c.has_mtp else 0          # mtp_proj runs once per code-predictor position
    return 2 * (talker + 15 * cp + heads + mtp)
wb = weight_bytes(cfg)
kv_tok = 2 * 2 * cfg.n_layers * cfg.n_kv_heads * 128
res["weight_bytes_per_frame_step"] = wb; res["kv_bytes_per_token"] = kv_tok

# ---- batch 1 ----
T = pkg.TTSTransformer(device=0, max_streams=1, max_ctx=512)
t0 = time.time(); assert T.load_model(path), T.get_error(); print(f"load {time.time() - t0:.1f}s", flush=True)
T.set_numerics(pkg.NUMERICS_FAST, 4); T.set_ignore_eos(True)
toks = pkg.synth.synth_tokens(cfg, n_text=208, seed=0); spk = pkg.synth.synth_speaker(cfg, seed=0)
for rep in range(3):
    out = T.generate(toks, spk, max_len=256, temperature=0.9, top_k=50, seed=rep)
    assert out is not None and len(out) == 256, T.get_error()
    ms = T.last_loop_ms()
byt = wb + kv_tok * (10 + 128)
res["batch1"] = {"frames": 256, "ms_per_frame": ms / 256, "frames_per_s": 256e3 / ms,
                 "achieved_gbs": byt / (ms / 256 * 1e-3) / 1e9, "roofline_frac": byt / (ms / 256 * 1e-3) / 1e9 / hbm}
print("batch1", json.dumps(res["batch1"]), flush=True)
T.unload_model(); del T

# ---- batch B, long context ----
T = pkg.TTSTransformer(device=0, max_streams=B, max_ctx=max_len + 64)
assert T.load_model(path), T.get_error()
T.set_numerics(pkg.NUMERICS_FAST, 4); T.set_ignore_eos(True)
t0 = time.time()
for i in range(B):
    toks = pkg.synth.synth_tokens(cfg, n_text=64, seed=i)
    assert T.stream_begin(i, toks, pkg.synth.synth_speaker(cfg, seed=i), max_len=max_len, temperature=0.9, top_k=50, seed=i), T.get_error()
res["stream_begin_ms_per_stream"] = (time.time() - t0) / B * 1e3
chunks = []; tot = 0; ms_tot = 0.0; n_kv = 10
while True:
    codes, produced, active = T.step_frames(40)
    n = int(produced.sum()); ms = T.last_loop_ms()
    if n:
        steps = n / B
        byt = wb + B * kv_tok * (n_kv + steps / 2 + 1)
        chunks.append({"n_kv": int(n_kv + steps / 2), "ms_per_frame_step": ms / steps, "frames_per_s": n / ms * 1e3,
                       "achieved_gbs": byt / (ms / steps * 1e-3) / 1e9, "roofline_frac": byt / (ms / steps * 1e-3) / 1e9 / hbm})
        n_kv += steps; tot += n; ms_tot += ms
    if not active.any():
        break
res["batched"] = {"streams": B, "max_len": max_len, "frames": tot, "device_ms": ms_tot, "frames_per_s": tot / ms_tot * 1e3,
                  "ms_per_frame_step": ms_tot / (tot / B),
                  "by_context": [chunks[i] for i in sorted(set([0, len(chunks) // 8, len(chunks) // 4, len(chunks) // 2, len(chunks) - 1]))]}
print("batched", json.dumps(res["batched"]), flush=True)
if out_path:
    json.dump(res, open(out_path, "w"), indent=1)
