"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libqtts_ref.so,
built by oracle/Makefile.ref from /root/reference).  Run in the CPU container:

    python tests/golden/make_golden.py [tiny tiny-mtp 0.6b]

Each fixture holds, for one seeded synthetic GGUF (qwen3-tts-jetson_b200/synth.py, regenerated
from the seed at test time), the reference's outputs on the hot path:
  tokens, speaker                      inputs (encode_for_tts frame, speaker.bin payload)
  prefill, trailing, pad               build_prefill_graph   (src/tts_transformer.cpp:1093-1231)
  pre_hidden, pre_logits               forward_prefill       (:1829-1920)
  step_embd, step_logits, step_hidden  3 teacher-forced forward_step calls (:1952-2028)
  cp_cb0, cp_codes, cp_logits          predict_codes_autoregressive graphs (:2153-2340), greedy
  codes                                generate(), greedy, max_len frames (:2342-2574)
  threads                              ggml CPU threads used (4 = the reference's default)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import qtts_pkg  # noqa: E402
import ref  # noqa: E402

SEED = 1234
FRAMES = {"tiny": 48, "tiny-mtp": 48, "0.6b": 32, "1.7b": 8}
TEXT = "Hello, world!"


def make(preset: str, gguf_dir: str = "/tmp"):
    m = qtts_pkg.load()
    synth = m.synth
    cfg = synth.PRESETS[preset]
    path = os.path.join(gguf_dir, f"synth-{preset}-{SEED}.gguf")
    if not os.path.exists(path):
        synth.write_gguf(path, preset, SEED)
    toks = synth.synth_tokens(cfg, TEXT)
    spk = synth.synth_speaker(cfg)
    R = ref.Reference(path)          # ggml default: 4 threads
    pre, trail, pad = R.build_prefill(toks, spk)
    R.init_kv(512)
    pre_hidden, pre_logits = R.forward_prefill(pre)
    rng = np.random.default_rng(0)
    n_past = pre.shape[0]
    embs, sl, sh = [], [], []
    for step in range(3):
        e = (0.05 * rng.standard_normal(cfg.hidden_size)).astype(np.float32)
        lg, hd = R.forward_step(e, n_past + step)
        embs.append(e); sl.append(lg); sh.append(hd)
    cb0 = int(pre_logits[:2048].argmax())
    cp_codes, cp_logits = R.predict_codes_logits(pre_hidden, cb0)
    codes, _ = R.generate(toks, spk, max_len=FRAMES[preset], temperature=0.0)
    out = os.path.join(ROOT, "tests", "golden", f"ref_{preset}_seed{SEED}.npz")
    np.savez_compressed(
        out, tokens=toks, speaker=spk, prefill=pre, trailing=trail, pad=pad,
        pre_hidden=pre_hidden, pre_logits=pre_logits,
        step_embd=np.stack(embs), step_logits=np.stack(sl), step_hidden=np.stack(sh),
        cp_cb0=np.int32(cb0), cp_codes=cp_codes, cp_logits=cp_logits, codes=codes,
        threads=np.int32(4), seed=np.int32(SEED), max_len=np.int32(FRAMES[preset]))
    print(f"{out}: {codes.shape[0]} frames, {os.path.getsize(out) / 1e3:.0f} kB")
    R.close()


# ----------------------------------------------------------------------------------------------
# round 2: per-frame fixtures.  ref_<preset>_seed1234_frames.npz holds what the reference's
# generate() computes INSIDE its frame loop, dumped by replaying the loop through the reference's
# own methods (oracle/ref_harness.cpp:ref_generate_trace; the replay is accepted only if its codes
# equal the real generate()'s):
#   g_*   greedy "Hello, world!" run of the base fixture: raw Talker logits, hidden and next-input
#         embedding of the first G_FRAMES frames, code-predictor logits of frames 1..2
#   e_*   greedy "Hi" (10 tokens: EOS ramp from frame 40, src/tts_transformer.cpp:2439-2445) run
#         until the reference stops: all codes, the stop frame, raw + processed logits around the ramp
#   s_*   sampled run (temperature 0.9, top-k 50) with the reference's rng_ re-seeded to S_SEED:
#         codes and the uniforms each sampled code consumed (codebook 0: uniform_real_distribution<float>;
#         codes 1..15: the generate_canonical<double,53> draw of std::discrete_distribution)
# ----------------------------------------------------------------------------------------------
G_FRAMES = {"tiny": 12, "tiny-mtp": 12, "0.6b": 12, "1.7b": 6}
S_FRAMES = {"tiny": 24, "tiny-mtp": 24, "0.6b": 24, "1.7b": 8}
S_SEED = 3


def make_frames(preset: str, gguf_dir: str = "/tmp"):
    m = qtts_pkg.load()
    synth = m.synth
    cfg = synth.PRESETS[preset]
    path = os.path.join(gguf_dir, f"synth-{preset}-{SEED}.gguf")
    if not os.path.exists(path):
        synth.write_gguf(path, preset, SEED)
    toks = synth.synth_tokens(cfg, TEXT)
    spk = synth.synth_speaker(cfg)
    R = ref.Reference(path)
    out = {}
    # greedy base run
    n = FRAMES[preset]
    codes, _ = R.generate(toks, spk, max_len=n, temperature=0.0)
    tr = R.generate_trace(toks, spk, max_len=n, temperature=0.0)
    assert np.array_equal(tr["codes"], codes), "trace replay differs from generate()"
    g = G_FRAMES[preset]
    out.update(g_codes=codes, g_raw=tr["raw"][:g], g_hidden=tr["hidden"][:g], g_step_embd=tr["step_embd"][:g])
    cpl = []
    for f in (1, 2):
        c15, lg = R.predict_codes_logits(tr["hidden"][f], int(codes[f, 0]))
        assert np.array_equal(c15, codes[f, 1:])
        cpl.append(lg)
    out["g_cp_logits"] = np.stack(cpl)
    # run to EOS
    toks_e = synth.synth_tokens(cfg, "Hi")
    codes_e, _ = R.generate(toks_e, None, max_len=400, temperature=0.0)
    te = R.generate_trace(toks_e, None, max_len=400, temperature=0.0)
    assert np.array_equal(te["codes"], codes_e) and te["eos_frame"] == len(codes_e), "EOS trace differs from generate()"
    expected = max(20, 4 * len(toks_e))
    fr = sorted({expected - 1, expected, expected + 1, (expected + te["eos_frame"]) // 2, te["eos_frame"] - 1, te["eos_frame"]})
    fr = [f for f in fr if 0 <= f < te["raw"].shape[0]]
    out.update(e_tokens=toks_e, e_codes=codes_e, e_eos_frame=np.int32(te["eos_frame"]), e_frames=np.asarray(fr, np.int32),
               e_raw=te["raw"][fr], e_proc=te["proc"][fr])
    # sampled run, seeded
    ns = S_FRAMES[preset]
    R.seed_rng(S_SEED)
    codes_s, _ = R.generate(toks, spk, max_len=ns, temperature=0.9, top_k=50)
    R.seed_rng(S_SEED)
    ts = R.generate_trace(toks, spk, max_len=ns, temperature=0.9, top_k=50)
    assert np.array_equal(ts["codes"], codes_s), "sampled trace differs from generate()"
    u = np.zeros((ts["raw"].shape[0], 16), np.float64)
    u[:, 0] = ts["u_cb0"]
    u[:ts["n_frames"], 1:] = ts["u_cp"]
    out.update(s_codes=codes_s, s_u=u, s_eos_frame=np.int32(ts["eos_frame"]), s_seed=np.int32(S_SEED),
               s_proc=ts["proc"][:4], s_temperature=np.float32(0.9), s_top_k=np.int32(50))
    dst = os.path.join(ROOT, "tests", "golden", f"ref_{preset}_seed{SEED}_frames.npz")
    np.savez_compressed(dst, **out)
    print(f"{dst}: greedy {len(codes)} frames, EOS run stops at {te['eos_frame']}, sampled {len(codes_s)} frames "
          f"(eos {ts['eos_frame']}), {os.path.getsize(dst) / 1e3:.0f} kB")
    R.close()


if __name__ == "__main__":
    args = sys.argv[1:]
    if args and args[0] == "frames":
        for p in (args[1:] or ["tiny", "tiny-mtp", "0.6b", "1.7b"]):
            make_frames(p)
    else:
        for p in (args or ["tiny", "tiny-mtp"]):
            make(p)
