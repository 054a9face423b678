"""Development probe run on the GPU box: engine vs the compiled reference, verbose numbers."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import qtts_pkg  # noqa: E402
import ref  # noqa: E402

m = qtts_pkg.load()
synth = m.synth


def nmse(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))


def check(preset, n_frames, text="Hello, world!", speaker=True, n_text=None):
    path = f"/tmp/qtts-{preset}.gguf"
    if not os.path.exists(path):
        t0 = time.time(); synth.write_gguf(path, preset); print(f"[{preset}] wrote gguf in {time.time()-t0:.1f}s")
    cfg = synth.PRESETS[preset]
    toks = synth.synth_tokens(cfg, text) if n_text is None else synth.synth_tokens(cfg, n_text=n_text)
    spk = synth.synth_speaker(cfg) if speaker else None
    R = ref.Reference(path)
    t0 = time.time()
    T = m.TTSTransformer(max_ctx=1024)
    assert T.load_model(path), T.get_error()
    print(f"[{preset}] engine load {time.time()-t0:.1f}s stats={T.weight_stats()}")
    # --- prefill embeddings
    pre_r, trail_r, pad_r = R.build_prefill(toks, spk)
    out = T.build_prefill(toks, spk)
    assert out is not None, T.get_error()
    pre_g, trail_g, pad_g = out
    print(f"[{preset}] prefill rows {pre_r.shape} vs {pre_g.shape}; max|d| prefill {np.abs(pre_r-pre_g).max():.3e} "
          f"trailing {np.abs(trail_r-trail_g).max():.3e} pad {np.abs(pad_r-pad_g).max():.3e}")
    # --- forward_prefill
    R.init_kv(512)
    hid_r, log_r = R.forward_prefill(pre_r)
    T.clear_kv_cache()
    hid_g, log_g = T.forward_prefill(pre_r)
    hid_g = hid_g[-1]
    print(f"[{preset}] forward_prefill nmse logits {nmse(log_g, log_r):.3e} hidden {nmse(hid_g, hid_r):.3e} "
          f"argmax {log_r.argmax()} {log_g.argmax()}")
    # --- forward_step (teacher-forced on reference inputs), both numerics
    rng = np.random.default_rng(0)
    n_past = pre_r.shape[0]
    for step in range(3):
        emb = (0.05 * rng.standard_normal(cfg.hidden_size)).astype(np.float32)
        lr, hr = R.forward_step(emb, n_past + step)
        lg, hg = T.forward_step(emb, n_past + step)
        print(f"[{preset}] forward_step {step} nmse logits {nmse(lg, lr):.3e} hidden {nmse(hg, hr):.3e} max|d| {np.abs(lg-lr).max():.3e}")
    # --- code predictor logits
    cb0 = int(log_r[:2048].argmax())
    codes_r, cl_r = R.predict_codes_logits(hid_r, cb0)
    codes_g, cl_g = T.predict_codes_autoregressive(hid_r, cb0, 0.0, 0, forced=codes_r, want_logits=True)
    print(f"[{preset}] code-pred teacher-forced: codes equal {np.array_equal(codes_r, codes_g)} "
          f"nmse per step max {max(nmse(cl_g[i], cl_r[i]) for i in range(15)):.3e}")
    # --- generate greedy
    for mode, name in [(m.NUMERICS_REFERENCE, "reference-numerics"), (m.NUMERICS_FAST, "fast-numerics")]:
        T.set_numerics(mode, 4)
        for persistent in ([True, False] if preset.startswith("tiny") else [True]):
            T.set_persistent(persistent)
            t0 = time.time()
            g = T.generate(toks, spk, max_len=n_frames, temperature=0.0)
            dt = time.time() - t0
            assert g is not None, T.get_error()
            if mode == m.NUMERICS_REFERENCE and persistent:
                t1 = time.time(); r, ms = R.generate(toks, spk, max_len=n_frames, temperature=0.0); rt = time.time() - t1
                print(f"[{preset}] reference generate {r.shape[0]} frames in {rt:.2f}s")
                ref_codes = r
            n = min(len(g), len(ref_codes))
            neq = (g[:n] != ref_codes[:n]).any(axis=1)
            first_bad = int(np.argmax(neq)) if neq.any() else -1
            print(f"[{preset}] {name} persistent={persistent}: {len(g)} frames (ref {len(ref_codes)}), wall {dt:.3f}s, "
                  f"loop {T.last_loop_ms():.2f} ms => {T.last_loop_ms()/max(len(g),1):.3f} ms/frame, "
                  f"first mismatching frame {first_bad}")
    T.set_persistent(True)
    # sampling smoke
    T.set_numerics(m.NUMERICS_FAST, 4)
    g = T.generate(toks, spk, max_len=n_frames, temperature=0.9, top_k=50, seed=1)
    print(f"[{preset}] sampled: {len(g)} frames, loop {T.last_loop_ms():.2f} ms; cb0 head {g[:8,0].tolist()}")
    T.unload_model(); R.close()


if __name__ == "__main__":
    which = sys.argv[1:] or ["tiny", "tiny-mtp", "0.6b"]
    for p in which:
        check(p, 48 if p.startswith("tiny") else 32)
