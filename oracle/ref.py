"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libqtts_ref.so (the UNMODIFIED reference
`qwen3_tts::TTSTransformer` on the ggml CPU backend; built by oracle/Makefile.ref).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (qwen3-tts-jetson_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libqtts_ref.so")

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def available() -> bool:
    return os.path.exists(LIB_PATH)


_lib = None


def _load():
    global _lib
    if _lib is not None:
        return _lib
    lib = C.CDLL(LIB_PATH)
    lib.ref_create.restype = C.c_void_p
    lib.ref_destroy.argtypes = [C.c_void_p]
    lib.ref_error.restype = C.c_char_p
    lib.ref_error.argtypes = [C.c_void_p]
    lib.ref_load.argtypes = [C.c_void_p, C.c_char_p]
    lib.ref_set_threads.argtypes = [C.c_void_p, C.c_int]
    lib.ref_get_config.argtypes = [C.c_void_p, _i32p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.ref_build_prefill.argtypes = [C.c_void_p, _i32p, C.c_int, C.c_void_p, C.c_int, _f32p, _f32p,
                                      _f32p, C.POINTER(C.c_int)]
    lib.ref_init_kv.argtypes = [C.c_void_p, C.c_int]
    lib.ref_forward_prefill.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _f32p, _f32p]
    lib.ref_forward_step.argtypes = [C.c_void_p, _f32p, C.c_int, _f32p, _f32p]
    lib.ref_predict_codes.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_float, C.c_int, _i32p]
    lib.ref_predict_codes_logits.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_void_p, _i32p, _f32p]
    lib.ref_generate.argtypes = [C.c_void_p, _i32p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                 C.c_float, C.c_int, _i32p, C.POINTER(C.c_double)]
    lib.ref_seed_rng.argtypes = [C.c_void_p, C.c_uint32]
    lib.ref_sample_cb0.argtypes = [_f32p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, C.POINTER(C.c_int)]
    lib.ref_generate_trace.argtypes = [C.c_void_p, _i32p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float,
                                       C.c_int] + [C.c_void_p] * 7 + [C.POINTER(C.c_int)]
    lib.ref_nth_check.argtypes = [_f32p, C.c_int, C.c_int, _i32p, _i32p]
    _lib = lib
    return lib


def nth_check(vals, k: int):
    """(std::nth_element's arrangement, csrc/qtts_nth.cuh's arrangement) of the ids 0..n-1 by descending value."""
    v = np.ascontiguousarray(vals, dtype=np.float32)
    a = np.zeros(v.shape[0], np.int32); b = np.zeros(v.shape[0], np.int32)
    _load().ref_nth_check(v, v.shape[0], int(k), a, b)
    return a, b


def sample_cb0(processed_logits, eos_id: int, temperature: float, top_k: int, u: float):
    """src/tts_transformer.cpp:2451-2494 with the reference's own libstdc++ calls (std::nth_element
    threshold included) on post-processed codebook-0 logits.  Returns (token, candidates kept)."""
    lg = np.ascontiguousarray(processed_logits, dtype=np.float32)
    kept = C.c_int()
    tok = _load().ref_sample_cb0(lg, lg.shape[0], int(eos_id), float(temperature), int(top_k), float(u), C.byref(kept))
    return int(tok), int(kept.value)


class Reference:
    """The reference TTSTransformer, CPU backend.  threads<=0 keeps ggml's default of 4."""

    def __init__(self, gguf_path: str, threads: int = 0):
        self.lib = _load()
        self.h = self.lib.ref_create()
        if not self.lib.ref_load(self.h, gguf_path.encode()):
            raise RuntimeError("reference load_model failed: " + self.error())
        self.lib.ref_set_threads(self.h, threads)
        cfg = np.zeros(16, dtype=np.int32)
        eps, theta = C.c_float(), C.c_float()
        self.lib.ref_get_config(self.h, cfg, C.byref(eps), C.byref(theta))
        keys = ["hidden", "n_layers", "n_heads", "n_kv_heads", "ffn", "head_dim", "codec_vocab",
                "n_codebooks", "cp_layers", "cp_vocab", "cp_hidden", "cp_ffn", "cp_heads",
                "cp_kv_heads", "cp_head_dim", "text_embd_dim"]
        self.cfg = dict(zip(keys, cfg.tolist()))
        self.cfg["eps"], self.cfg["theta"] = eps.value, theta.value

    def error(self) -> str:
        return self.lib.ref_error(self.h).decode()

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _opt(a):
        return None if a is None else a.ctypes.data_as(C.c_void_p)

    def build_prefill(self, tokens, speaker=None, language_id=2050):
        H = self.cfg["hidden"]
        tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        spk = None if speaker is None else np.ascontiguousarray(speaker, dtype=np.float32)
        pre = np.zeros((16, H), dtype=np.float32)
        trail = np.zeros((len(tokens) + 1, H), dtype=np.float32)
        pad = np.zeros(H, dtype=np.float32)
        tl = C.c_int()
        n = self.lib.ref_build_prefill(self.h, tokens, len(tokens), self._opt(spk), language_id,
                                       pre, trail, pad, C.byref(tl))
        if n < 0:
            raise RuntimeError(self.error())
        return pre[:n].copy(), trail[:tl.value].copy(), pad

    def init_kv(self, n_ctx: int):
        if not self.lib.ref_init_kv(self.h, n_ctx):
            raise RuntimeError(self.error())

    def forward_prefill(self, embd, n_past=0):
        embd = np.ascontiguousarray(embd, dtype=np.float32)
        hid = np.zeros(self.cfg["hidden"], dtype=np.float32)
        logits = np.zeros(self.cfg["codec_vocab"], dtype=np.float32)
        if not self.lib.ref_forward_prefill(self.h, embd, embd.shape[0], n_past, hid, logits):
            raise RuntimeError(self.error())
        return hid, logits

    def forward_step(self, step_embd, n_past):
        step_embd = np.ascontiguousarray(step_embd, dtype=np.float32)
        hid = np.zeros(self.cfg["hidden"], dtype=np.float32)
        logits = np.zeros(self.cfg["codec_vocab"], dtype=np.float32)
        if not self.lib.ref_forward_step(self.h, step_embd, n_past, logits, hid):
            raise RuntimeError(self.error())
        return logits, hid

    def predict_codes(self, hidden, cb0, temperature=0.0, top_k=50):
        out = np.zeros(15, dtype=np.int32)
        hidden = np.ascontiguousarray(hidden, dtype=np.float32)
        if not self.lib.ref_predict_codes(self.h, hidden, int(cb0), temperature, top_k, out):
            raise RuntimeError(self.error())
        return out

    def predict_codes_logits(self, hidden, cb0, forced=None):
        out = np.zeros(15, dtype=np.int32)
        logits = np.zeros((15, self.cfg["cp_vocab"]), dtype=np.float32)
        hidden = np.ascontiguousarray(hidden, dtype=np.float32)
        f = None if forced is None else np.ascontiguousarray(forced, dtype=np.int32)
        if not self.lib.ref_predict_codes_logits(self.h, hidden, int(cb0), self._opt(f), out, logits):
            raise RuntimeError(self.error())
        return out, logits

    def seed_rng(self, seed: int):
        """Re-seed the reference's sampler (it seeds from std::random_device otherwise)."""
        self.lib.ref_seed_rng(self.h, int(seed) & 0xffffffff)

    def generate_trace(self, tokens, speaker=None, max_len=64, language_id=2050, rep_pen=1.05,
                       temperature=0.0, top_k=50):
        """generate() replayed through the reference's own methods with per-frame dumps
        (oracle/ref_harness.cpp:ref_generate_trace).  Returns a dict; rows = frames (+1 if EOS ended it)."""
        tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        spk = None if speaker is None else np.ascontiguousarray(speaker, dtype=np.float32)
        H, V, rows = self.cfg["hidden"], self.cfg["codec_vocab"], max_len + 1
        raw = np.zeros((rows, V), np.float32); proc = np.zeros((rows, V), np.float32)
        hid = np.zeros((rows, H), np.float32); emb = np.zeros((rows, H), np.float32)
        u0 = np.zeros(rows, np.float32); ucp = np.zeros((rows, 15), np.float64)
        codes = np.zeros((rows, 16), np.int32)
        eos = C.c_int()
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        n = self.lib.ref_generate_trace(self.h, tokens, len(tokens), self._opt(spk), max_len, language_id, rep_pen,
                                        temperature, top_k, p(raw), p(proc), p(hid), p(emb), p(u0), p(ucp), p(codes),
                                        C.byref(eos))
        if n < 0:
            raise RuntimeError(self.error() or "ref_generate_trace failed")
        r = n + (1 if eos.value >= 0 else 0)
        return {"n_frames": n, "eos_frame": eos.value, "raw": raw[:r].copy(), "proc": proc[:r].copy(),
                "hidden": hid[:n].copy(), "step_embd": emb[:max(n - 1, 0)].copy(), "u_cb0": u0[:r].copy(),
                "u_cp": ucp[:n].copy(), "codes": codes[:n].copy()}

    def generate(self, tokens, speaker=None, max_len=4096, language_id=2050, rep_pen=1.05,
                 temperature=0.0, top_k=50):
        tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        spk = None if speaker is None else np.ascontiguousarray(speaker, dtype=np.float32)
        codes = np.zeros(max_len * 16, dtype=np.int32)
        ms = C.c_double()
        n = self.lib.ref_generate(self.h, tokens, len(tokens), self._opt(spk), max_len, language_id,
                                  rep_pen, temperature, top_k, codes, C.byref(ms))
        if n < 0:
            raise RuntimeError(self.error())
        return codes[: n * 16].reshape(n, 16).copy(), ms.value
