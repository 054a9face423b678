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


if __name__ == "__main__":
    for p in (sys.argv[1:] or ["tiny", "tiny-mtp"]):
        make(p)
