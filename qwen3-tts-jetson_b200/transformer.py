"""Host-side mirror of `qwen3_tts::TTSTransformer` (reference: src/tts_transformer.h:164-245) and of
`qwen3_tts::TRTCodePredictor`'s loops (src/trt_code_predictor.h:25-87) over the C ABI of
libqtts_b200.so (include/qtts_b200.h).  Same method names, argument meaning and error
behaviour (bool return + get_error()); numpy arrays stand in for std::vector.

The library is the product; this module only marshals pointers.  There is no CPU fallback:
if the CUDA library is missing or no B200 is present every call fails.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build
from . import gguf_io

NUMERICS_REFERENCE = 0
NUMERICS_FAST = 1
F32, F16, BF16 = 0, 1, 2
EPI_STORE, EPI_ADD, EPI_SWIGLU, EPI_BIAS = 0, 1, 2, 3

_CFG_INT_FIELDS = ["hidden_size", "n_layers", "n_heads", "n_kv_heads", "intermediate_size", "head_dim",
                   "codec_vocab_size", "n_codebooks", "cp_layers", "cp_vocab_size", "cp_hidden_size",
                   "cp_intermediate_size", "cp_n_heads", "cp_n_kv_heads", "cp_head_dim",
                   "text_vocab_size", "text_embd_dim"]
_CFG_TAIL_FIELDS = ["codec_pad_id", "codec_bos_id", "codec_eos_id", "codec_think_id", "codec_nothink_id",
                    "codec_think_bos_id", "codec_think_eos_id", "tts_bos_id", "tts_eos_id", "tts_pad_id"]


class QttsConfig(C.Structure):
    _fields_ = ([(n, C.c_int32) for n in _CFG_INT_FIELDS] +
                [("rms_norm_eps", C.c_float), ("rope_theta", C.c_float)] +
                [(n, C.c_int32) for n in _CFG_TAIL_FIELDS])


FRAME_CB = C.CFUNCTYPE(C.c_int, C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_void_p)

# every symbol include/qtts_b200.h declares
EXPORTS = ["qtts_create", "qtts_destroy", "qtts_last_error", "qtts_create_error", "qtts_upload_tensor",
           "qtts_finalize", "qtts_weight_stats", "qtts_set_numerics", "qtts_set_persistent", "qtts_set_ignore_eos",
           "qtts_clear_kv_cache", "qtts_build_prefill", "qtts_forward_prefill", "qtts_forward_step",
           "qtts_predict_codes", "qtts_generate", "qtts_stream_begin", "qtts_stream_prepare", "qtts_prefill_pending",
           "qtts_step_frames",
           "qtts_set_batched", "qtts_set_batched_weight_format", "qtts_stream_logits", "qtts_batched_profile",
           "qtts_debug_bgemm", "qtts_profile", "qtts_debug_barrier_us", "qtts_launch_count", "qtts_last_loop_ms", "qtts_stream_handle",
           "qtts_forward_text", "qtts_forward_codec", "qtts_stream_force", "qtts_sample_cb0", "qtts_set_debug", "qtts_resize_kv", "qtts_get_hidden_states", "qtts_cp_step", "qtts_trace", "qtts_codes_chunk", "qtts_load_gguf", "qtts_load_stats", "qtts_codes_chunk_noalloc"]

_lib = None


def load_library(path: str | None = None):
    """dlopen libqtts_b200.so (built in-tree by build.py) and declare the prototypes."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or os.environ.get("QTTS_LIB_PATH") or _build.LIB      # QTTS_LIB_PATH: A/B builds of the same sources (tools/ab_build.py)
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run __graft_entry__.build() (nvcc, sm_100a). "
                           "There is no CPU fallback for this path.")
    lib = C.CDLL(path)
    vp, i32, f32p, i32p = C.c_void_p, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int32)
    lib.qtts_create.argtypes = [C.POINTER(QttsConfig), C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    lib.qtts_destroy.argtypes = [vp]
    lib.qtts_destroy.restype = None
    lib.qtts_last_error.argtypes = [vp]
    lib.qtts_last_error.restype = C.c_char_p
    lib.qtts_create_error.restype = C.c_char_p
    lib.qtts_upload_tensor.argtypes = [vp, C.c_char_p, C.c_int, C.POINTER(C.c_int64), C.c_int, vp]
    lib.qtts_finalize.argtypes = [vp]
    lib.qtts_load_gguf.argtypes = [vp, C.c_char_p]
    lib.qtts_load_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    lib.qtts_weight_stats.argtypes = [vp, C.POINTER(C.c_int64)]
    lib.qtts_set_numerics.argtypes = [vp, C.c_int, C.c_int]
    lib.qtts_set_persistent.argtypes = [vp, C.c_int]
    lib.qtts_set_ignore_eos.argtypes = [vp, C.c_int]
    lib.qtts_clear_kv_cache.argtypes = [vp, C.c_int]
    lib.qtts_build_prefill.argtypes = [vp, C.c_int, vp, i32, vp, i32, i32p, i32p, vp, vp, vp]
    lib.qtts_forward_prefill.argtypes = [vp, C.c_int, vp, i32, i32, vp, vp]
    lib.qtts_forward_step.argtypes = [vp, C.c_int, vp, i32, vp, vp]
    lib.qtts_forward_text.argtypes = [vp, C.c_int, vp, i32, vp, i32, vp]
    lib.qtts_forward_codec.argtypes = [vp, C.c_int, i32, i32, vp]
    lib.qtts_stream_force.argtypes = [vp, C.c_int, vp, i32, vp, i32]
    lib.qtts_sample_cb0.argtypes = [vp, C.c_int, vp, i32, i32, vp, i32, C.c_float, C.c_float, i32, C.c_float, i32p, vp]
    lib.qtts_set_debug.argtypes = [vp, C.c_int]
    lib.qtts_resize_kv.argtypes = [vp, i32, i32]
    lib.qtts_get_hidden_states.argtypes = [vp, C.c_int, vp]
    lib.qtts_codes_chunk.argtypes = [vp, C.c_int, i32, i32, i32, vp, C.c_int]
    lib.qtts_codes_chunk_noalloc.argtypes = [vp, C.c_int, i32, i32, i32, vp]
    lib.qtts_trace.argtypes = [vp, C.c_int, C.c_int, vp, i32]
    lib.qtts_cp_step.argtypes = [vp, C.c_int, vp, i32, i32, i32, i32, i32, C.c_int, vp, i32]
    lib.qtts_predict_codes.argtypes = [vp, C.c_int, vp, i32, C.c_float, i32, vp, vp, vp, vp]
    lib.qtts_generate.argtypes = [vp, C.c_int, vp, i32, vp, i32, i32, C.c_float, C.c_float, i32, C.c_uint64,
                                  vp, i32p, FRAME_CB, i32, vp]
    lib.qtts_stream_begin.argtypes = [vp, C.c_int, vp, i32, vp, i32, i32, C.c_float, C.c_float, i32, C.c_uint64]
    lib.qtts_stream_prepare.argtypes = [vp, C.c_int, vp, i32, vp, i32, i32, C.c_float, C.c_float, i32, C.c_uint64]
    lib.qtts_prefill_pending.argtypes = [vp]
    lib.qtts_step_frames.argtypes = [vp, i32, vp, vp, vp]
    lib.qtts_profile.argtypes = [vp, C.c_int, vp]
    lib.qtts_set_batched.argtypes = [vp, C.c_int]
    lib.qtts_set_batched_weight_format.argtypes = [vp, C.c_int]
    lib.qtts_stream_logits.argtypes = [vp, C.c_int, vp, vp]
    lib.qtts_batched_profile.argtypes = [vp, C.c_int, vp]
    lib.qtts_debug_bgemm.argtypes = [C.c_int] * 8 + [vp] * 5 + [C.c_float, vp, vp]
    lib.qtts_debug_barrier_us.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.qtts_debug_barrier_us.restype = C.c_double
    lib.qtts_launch_count.argtypes = [vp]
    lib.qtts_launch_count.restype = C.c_int64
    lib.qtts_last_loop_ms.argtypes = [vp]
    lib.qtts_last_loop_ms.restype = C.c_double
    lib.qtts_stream_handle.argtypes = [vp]
    lib.qtts_stream_handle.restype = vp
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class TTSTransformer:
    """Drop-in for the reference class on one B200.  `max_streams` > 1 adds the concurrent
    utterance slots the reference lacks (each reference-level method takes `slot`)."""

    def __init__(self, device: int = 0, max_streams: int = 1, max_ctx: int = 4224):
        self._lib = load_library()
        self._h = C.c_void_p()
        self._err = ""
        self.device, self.max_streams, self.max_ctx = device, max_streams, max_ctx
        self.config: dict = {}
        self.batched_weight_format = F16     # B operand of the batched kernel's GEMMs (applied at load)

    # -- error handling (bool + get_error(), like the reference) --------------------------------
    def get_error(self) -> str:
        if self._h:
            return self._lib.qtts_last_error(self._h).decode()
        return self._err

    def get_config(self) -> dict:
        return self.config

    # -- load_model (src/tts_transformer.cpp:51-153) ----------------------------------------------
    def load_model(self, model_path: str) -> bool:
        """parse_config from the GGUF metadata here (host logic), tensors by the library's own loader
        (qtts_load_gguf: mapped file -> pinned staging ring -> HBM, quantised types rejected by name)."""
        self.unload_model()
        try:
            cfg, _ = gguf_io.read_gguf(model_path)
        except Exception as ex:  # unreadable file
            self._err = f"Failed to open GGUF file: {ex}"
            return False
        if not self._create(cfg):
            return False
        if not self._lib.qtts_load_gguf(self._h, model_path.encode()):
            self._err = self._lib.qtts_last_error(self._h).decode()
            self.unload_model()
            return False
        if not self._lib.qtts_finalize(self._h):
            self._err = self._lib.qtts_last_error(self._h).decode()
            self.unload_model()
            return False
        return True

    def load_stats(self) -> dict:
        ms, nbytes = C.c_double(), C.c_int64()
        self._lib.qtts_load_stats(self._h, C.byref(ms), C.byref(nbytes))
        return {"load_ms": ms.value, "bytes": nbytes.value, "gb_per_s": nbytes.value / 1e6 / ms.value if ms.value else 0.0}

    def _create(self, cfg: dict) -> bool:
        self.config = dict(cfg)
        c = QttsConfig()
        for n in _CFG_INT_FIELDS + _CFG_TAIL_FIELDS:
            setattr(c, n, int(cfg[n]))
        c.rms_norm_eps, c.rope_theta = cfg["rms_norm_eps"], cfg["rope_theta"]
        h = C.c_void_p()
        if not self._lib.qtts_create(C.byref(c), self.device, self.max_streams, self.max_ctx, C.byref(h)):
            self._err = self._lib.qtts_create_error().decode()
            return False
        self._h = h
        return bool(self._lib.qtts_set_batched_weight_format(self._h, int(self.batched_weight_format)))

    def load_tensors(self, cfg: dict, tensors) -> bool:
        """Tensor-by-tensor upload (qtts_upload_tensor) for callers that hold the tensors in memory.
        cfg: gguf_io.parse_config dict; tensors: iterable of (name, dtype_code, ggml_dims, array)."""
        if not self._create(cfg):
            return False
        for name, code, dims, arr in tensors:
            d = (C.c_int64 * len(dims))(*dims)
            if not self._lib.qtts_upload_tensor(self._h, name.encode(), code, d, len(dims), _ptr(arr)):
                return False
        return bool(self._lib.qtts_finalize(self._h))

    def unload_model(self):
        if self._h:
            self._lib.qtts_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.unload_model()
        except Exception:
            pass

    def is_loaded(self) -> bool:
        return bool(self._h)

    # -- engine switches -----------------------------------------------------------------------
    def set_numerics(self, mode: int, reference_threads: int = 4) -> bool:
        return bool(self._lib.qtts_set_numerics(self._h, mode, reference_threads))

    def set_persistent(self, on: bool) -> bool:
        return bool(self._lib.qtts_set_persistent(self._h, 1 if on else 0))

    def set_ignore_eos(self, on: bool) -> bool:
        """Benchmark switch (fixed-length utterances); not reference behaviour."""
        return bool(self._lib.qtts_set_ignore_eos(self._h, 1 if on else 0))

    def set_batched(self, on: bool) -> bool:
        """step_frames: batched tcgen05 kernel (default) or the batch-1 kernel stream after stream."""
        return bool(self._lib.qtts_set_batched(self._h, 1 if on else 0))

    def stream_logits(self, slot: int, want_cp: bool = True):
        """Parity hook: processed cb0 logits and (batched mode) code-predictor logits of the last frame."""
        cb0 = np.zeros(self.config["codec_vocab_size"], np.float32)
        cp = np.zeros((self.config["n_codebooks"] - 1, self.config["cp_vocab_size"]), np.float32) if want_cp else None
        if not self._lib.qtts_stream_logits(self._h, slot, _ptr(cb0), _ptr(cp)):
            return None
        return cb0, cp

    def batched_profile(self, on: bool):
        out = np.zeros(64, np.uint64)
        self._lib.qtts_batched_profile(self._h, 1 if on else 0, _ptr(out))
        o = out.reshape(16, 4)
        names = {0: "gemm_f32in", 8: "gemm_f16in", 12: "norm", 2: "attn_cp", 4: "attn_talker", 6: "sample"}
        return {n: {"count": int(o[i, 0]), "stage_cycles": int(o[i, 1]), "barrier_cycles": int(o[i, 2])}
                for i, n in names.items()}

    def weight_stats(self) -> dict:
        out = (C.c_int64 * 4)()
        self._lib.qtts_weight_stats(self._h, out)
        return {"n_weights": out[0], "n_not_bf16_exact": out[1], "weight_bytes": out[2], "kv_bytes": out[3]}

    def profile(self, on):
        """Read-and-clear the stage cycle counters (see qtts_profile), then switch profiling."""
        out = np.zeros(64 + 4 * 256, np.uint64)
        self._lib.qtts_profile(self._h, int(on), _ptr(out))
        names = ["gemv", "attn", "sample_cb0", "sample_cp", "frame_end", "copy"]
        d = {n: out[i * 8:(i + 1) * 8].tolist() for i, n in enumerate(names)}
        d["per_cta"] = out[64:].reshape(256, 4)
        return d

    def trace(self, on: bool, cta: int = 0, max_stages: int = 8192):
        """Read-and-clear the stage timeline (qtts_trace): uint32 [stages, 8]; then switch recording."""
        out = np.zeros((max_stages, 8), np.uint32)
        self._lib.qtts_trace(self._h, 1 if on else 0, cta, _ptr(out), max_stages)
        return out

    def launch_count(self) -> int:
        return int(self._lib.qtts_launch_count(self._h))

    def last_loop_ms(self) -> float:
        return float(self._lib.qtts_last_loop_ms(self._h))

    # -- KV cache (src/tts_transformer.cpp:794-900) -------------------------------------------------
    def init_kv_cache(self, n_ctx: int, max_text: int = 0) -> bool:
        """Re-size the per-stream KV slabs (weights stay resident); streams in flight are dropped."""
        if not self._lib.qtts_resize_kv(self._h, n_ctx, max_text):
            return False
        self.max_ctx = n_ctx
        return True

    def set_debug(self, flags: int) -> bool:
        return bool(self._lib.qtts_set_debug(self._h, flags))

    def clear_kv_cache(self, slot: int = 0) -> bool:
        return bool(self._lib.qtts_clear_kv_cache(self._h, slot))

    # -- build_prefill_graph (src/tts_transformer.cpp:1093-1231) ------------------------------------
    def build_prefill(self, text_tokens, speaker_embd=None, language_id: int = 2050, slot: int = 0):
        toks = np.ascontiguousarray(text_tokens, dtype=np.int32)
        spk = None if speaker_embd is None else np.ascontiguousarray(speaker_embd, dtype=np.float32)
        H = self.config["hidden_size"]
        pre = np.zeros((16, H), np.float32)
        trail = np.zeros((max(len(toks), 1) + 1, H), np.float32)
        pad = np.zeros(H, np.float32)
        pl, tl = C.c_int32(), C.c_int32()
        ok = self._lib.qtts_build_prefill(self._h, slot, _ptr(toks), len(toks), _ptr(spk), language_id,
                                          C.byref(pl), C.byref(tl), _ptr(pre), _ptr(trail), _ptr(pad))
        if not ok:
            return None
        return pre[:pl.value].copy(), trail[:tl.value].copy(), pad

    # -- forward_prefill / forward_step (src/tts_transformer.cpp:1829-2028) -------------------------
    def forward_prefill(self, prefill_embd, n_past: int = 0, slot: int = 0, all_rows: bool = True):
        """Returns (hidden [n_tokens, H] like the reference's `output`, logits [V] of the last token);
        all_rows=False skips the per-row hidden states (hidden = [1, H], the last row)."""
        e = np.ascontiguousarray(prefill_embd, dtype=np.float32).reshape(-1, self.config["hidden_size"])
        n = e.shape[0]
        logits = np.zeros(self.config["codec_vocab_size"], np.float32)
        if all_rows:
            hid = np.zeros((max(n, 1), self.config["hidden_size"]), np.float32)
            ok = self._lib.qtts_forward_prefill(self._h, slot, _ptr(e), n, n_past, _ptr(hid), _ptr(logits))
            return (hid, logits) if ok else None
        ok = self._lib.qtts_forward_prefill(self._h, slot, _ptr(e), n, n_past, None, _ptr(logits))
        if not ok:
            return None
        return self.get_hidden_states(slot)[None, :], logits

    def get_hidden_states(self, slot: int = 0):
        """get_hidden_states (src/tts_transformer.cpp:2042-2048): the normalised hidden of the last forward call."""
        hid = np.zeros(self.config["hidden_size"], np.float32)
        if not self._lib.qtts_get_hidden_states(self._h, slot, _ptr(hid)):
            return None
        return hid

    def forward_text(self, text_tokens, speaker_embd=None, n_past: int = 0, slot: int = 0):
        toks = np.ascontiguousarray(text_tokens, dtype=np.int32)
        spk = None if speaker_embd is None else np.ascontiguousarray(speaker_embd, dtype=np.float32)
        hid = np.zeros((max(len(toks), 1), self.config["hidden_size"]), np.float32)
        if not self._lib.qtts_forward_text(self._h, slot, _ptr(toks), len(toks), _ptr(spk), n_past, _ptr(hid)):
            return None
        return hid

    def forward_codec(self, codec_token: int, n_past: int, slot: int = 0):
        logits = np.zeros(self.config["codec_vocab_size"], np.float32)
        if not self._lib.qtts_forward_codec(self._h, slot, int(codec_token), n_past, _ptr(logits)):
            return None
        return logits

    def codes_chunk(self, slot: int, first_frame: int, n_frames: int, fixed_frames: int = 40, device_ptr: int | None = None):
        """Vocoder hand-off: int64 [1, n_codebooks, fixed_frames] chunk of the stream's device-resident code ring,
        zero-padded (src/trt_vocoder.cpp:98-170).  device_ptr: write to that device address instead (returns True)."""
        nb = self.config["n_codebooks"]
        if device_ptr is not None:
            return bool(self._lib.qtts_codes_chunk(self._h, slot, first_frame, n_frames, fixed_frames, C.c_void_p(device_ptr), 1))
        out = np.zeros((1, nb, fixed_frames), np.int64)
        if not self._lib.qtts_codes_chunk(self._h, slot, first_frame, n_frames, fixed_frames, _ptr(out), 0):
            return None
        return out

    def codes_chunk_noalloc(self, slot: int, first_frame: int, n_frames: int, fixed_frames: int = 40, out=None):
        """qtts_codes_chunk_noalloc: the same chunk through the library's static device scratch (no device allocation;
        n_codebooks * fixed_frames <= NOALLOC_CHUNK_WORDS).  `out`: caller's int64 [1, n_codebooks, fixed_frames] array."""
        nb = self.config["n_codebooks"]
        if out is None:
            out = np.zeros((1, nb, fixed_frames), np.int64)
        assert out.dtype == np.int64 and out.flags.c_contiguous and out.size >= nb * max(fixed_frames, 0)
        if not self._lib.qtts_codes_chunk_noalloc(self._h, slot, first_frame, n_frames, fixed_frames, _ptr(out)):
            return None
        return out

    # TRTCodePredictor::step / step_with_token (src/trt_code_predictor.h:53-66): one code-predictor position
    def cp_step(self, position_id: int, lm_head_idx: int, input_vec=None, token_id: int = 0, embedding_idx: int = 0,
                needs_proj: bool = True, slot: int = 0):
        n = self.config["cp_vocab_size"] if lm_head_idx >= 0 else self.config["cp_hidden_size"]
        out = np.zeros(n, np.float32)
        v = None if input_vec is None else np.ascontiguousarray(input_vec, dtype=np.float32)
        if not self._lib.qtts_cp_step(self._h, slot, _ptr(v), 0 if v is None else v.shape[0], int(token_id), int(embedding_idx),
                                      position_id, lm_head_idx, 1 if needs_proj else 0, _ptr(out), n):
            return None
        return out

    # -- parity hooks ------------------------------------------------------------------------------------
    def stream_force(self, slot: int, forced_codes=None, uniforms=None) -> bool:
        """Teacher forcing / host-supplied uniforms for a begun or prepared stream (qtts_stream_force)."""
        f = None if forced_codes is None else np.ascontiguousarray(forced_codes, dtype=np.int32)
        u = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float32)
        return bool(self._lib.qtts_stream_force(self._h, slot, _ptr(f), 0 if f is None else f.shape[0],
                                                _ptr(u), 0 if u is None else u.shape[0]))

    def sample_cb0(self, raw_logits, frame: int, n_text_tokens: int, seen_tokens=(), repetition_penalty: float = 1.05,
                   temperature: float = 0.0, top_k: int = 50, u: float = 0.5, slot: int = 0):
        """Codebook-0 post-processing + sampling of one frame (src/tts_transformer.cpp:2417-2499) on raw logits.
        Returns (token, processed logits)."""
        lg = np.ascontiguousarray(raw_logits, dtype=np.float32)
        seen = np.ascontiguousarray(sorted(int(t) for t in seen_tokens), dtype=np.int32)
        tok = C.c_int32()
        proc = np.zeros(self.config["codec_vocab_size"], np.float32)
        if not self._lib.qtts_sample_cb0(self._h, slot, _ptr(lg), frame, n_text_tokens, _ptr(seen) if len(seen) else None,
                                         len(seen), repetition_penalty, temperature, top_k, u, C.byref(tok), _ptr(proc)):
            return None
        return int(tok.value), proc

    def forward_step(self, step_embd, n_past: int, slot: int = 0):
        e = np.ascontiguousarray(step_embd, dtype=np.float32)
        hid = np.zeros(self.config["hidden_size"], np.float32)
        logits = np.zeros(self.config["codec_vocab_size"], np.float32)
        if not self._lib.qtts_forward_step(self._h, slot, _ptr(e), n_past, _ptr(logits), _ptr(hid)):
            return None
        return logits, hid

    # -- predict_codes_autoregressive (src/tts_transformer.cpp:2153-2340) ---------------------------
    def predict_codes_autoregressive(self, hidden, codebook_0_token: int, temperature: float = 0.9,
                                     top_k: int = 50, uniforms=None, forced=None, want_logits: bool = False,
                                     slot: int = 0):
        hid = np.ascontiguousarray(hidden, dtype=np.float32)
        nb = self.config["n_codebooks"] - 1
        out = np.zeros(nb, np.int32)
        logits = np.zeros((nb, self.config["cp_vocab_size"]), np.float32) if want_logits else None
        u = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float32)
        f = None if forced is None else np.ascontiguousarray(forced, dtype=np.int32)
        ok = self._lib.qtts_predict_codes(self._h, slot, _ptr(hid), int(codebook_0_token), temperature, top_k,
                                          _ptr(u), _ptr(f), _ptr(out), _ptr(logits))
        if not ok:
            return None
        return (out, logits) if want_logits else out

    # TRTCodePredictor::run_greedy_loop / run_sampling_loop (src/trt_code_predictor.cpp:484-726)
    def run_greedy_loop(self, hidden, cb0_token: int, slot: int = 0):
        return self.predict_codes_autoregressive(hidden, cb0_token, 0.0, 0, slot=slot)

    def run_sampling_loop(self, hidden, cb0_token: int, temperature: float, top_k: int, rand_vals, slot: int = 0):
        return self.predict_codes_autoregressive(hidden, cb0_token, temperature, top_k, uniforms=rand_vals, slot=slot)

    # -- generate (src/tts_transformer.cpp:2342-2574) -----------------------------------------------
    def generate(self, text_tokens, speaker_embd=None, max_len: int = 4096, language_id: int = 2050,
                 repetition_penalty: float = 1.05, temperature: float = 0.9, top_k: int = 50,
                 on_frames=None, callback_interval: int = 40, seed: int = 0, slot: int = 0):
        """Returns int32 [n_frames, n_codebooks] or None (see get_error())."""
        toks = np.ascontiguousarray(text_tokens, dtype=np.int32)
        spk = None if speaker_embd is None else np.ascontiguousarray(speaker_embd, dtype=np.float32)
        nb = self.config["n_codebooks"]
        codes = np.zeros(max(max_len, 1) * nb, np.int32)
        nf = C.c_int32()

        def _cb(ptr, n_frames, n_books, _user):
            arr = np.ctypeslib.as_array(ptr, shape=(n_frames * n_books,)).reshape(n_frames, n_books)
            return 1 if on_frames(arr.copy(), n_frames, n_books) else 0

        cb = FRAME_CB(_cb) if on_frames else C.cast(None, FRAME_CB)
        ok = self._lib.qtts_generate(self._h, slot, _ptr(toks), len(toks), _ptr(spk), max_len, language_id,
                                     repetition_penalty, temperature, top_k, seed, _ptr(codes), C.byref(nf),
                                     cb, callback_interval, None)
        if not ok:
            return None
        return codes[: nf.value * nb].reshape(nf.value, nb).copy()

    # -- concurrent utterances ----------------------------------------------------------------------
    def stream_begin(self, slot: int, text_tokens, speaker_embd=None, max_len: int = 4096, language_id: int = 2050,
                     repetition_penalty: float = 1.05, temperature: float = 0.9, top_k: int = 50, seed: int = 0) -> bool:
        toks = np.ascontiguousarray(text_tokens, dtype=np.int32)
        spk = None if speaker_embd is None else np.ascontiguousarray(speaker_embd, dtype=np.float32)
        return bool(self._lib.qtts_stream_begin(self._h, slot, _ptr(toks), len(toks), _ptr(spk), max_len, language_id,
                                                repetition_penalty, temperature, top_k, seed))

    def stream_prepare(self, slot: int, text_tokens, speaker_embd=None, max_len: int = 4096, language_id: int = 2050,
                       repetition_penalty: float = 1.05, temperature: float = 0.9, top_k: int = 50, seed: int = 0) -> bool:
        """Phase 1 of stream_begin (embeddings + state); the Talker prefill runs in prefill_pending()."""
        toks = np.ascontiguousarray(text_tokens, dtype=np.int32)
        spk = None if speaker_embd is None else np.ascontiguousarray(speaker_embd, dtype=np.float32)
        return bool(self._lib.qtts_stream_prepare(self._h, slot, _ptr(toks), len(toks), _ptr(spk), max_len, language_id,
                                                  repetition_penalty, temperature, top_k, seed))

    def prefill_pending(self) -> bool:
        """Batched Talker prefill of every prepared stream (forward_prefill for many utterances at once)."""
        return bool(self._lib.qtts_prefill_pending(self._h))

    def step_frames(self, n_frames: int):
        nb = self.config["n_codebooks"]
        codes = np.zeros((self.max_streams, n_frames, nb), np.int32)
        produced = np.zeros(self.max_streams, np.int32)
        active = np.zeros(self.max_streams, np.int32)
        if not self._lib.qtts_step_frames(self._h, n_frames, _ptr(codes), _ptr(produced), _ptr(active)):
            return None
        return codes, produced, active


def debug_bgemm(W, X, n_tok: int = 1, epi: int = EPI_STORE, in_f16: bool = False, norm_w=None, res=None, bias=None,
                eps: float = 1e-6, weight_fmt: int = F16, device: int = 0, want_ssq: bool = False):
    """One projection stage of the batched tcgen05 kernel on caller-supplied operands (qtts_debug_bgemm)."""
    lib = load_library()
    W = np.ascontiguousarray(W, np.float32)
    X = np.ascontiguousarray(X, np.float32)
    rows, K = W.shape
    rows_act = X.shape[0]
    assert X.shape[1] == K and rows_act % n_tok == 0
    outw = rows // 2 if epi == EPI_SWIGLU else rows
    out = np.zeros((rows_act, outw), np.float32)
    ssq = np.zeros((rows_act, rows // 16), np.float32) if want_ssq else None
    f = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    nw, rs, bs = f(norm_w), f(res), f(bias)
    ok = lib.qtts_debug_bgemm(device, weight_fmt, rows_act // n_tok, n_tok, rows, K, epi, 1 if in_f16 else 0,
                              _ptr(W), _ptr(X), _ptr(nw), _ptr(rs), _ptr(bs), eps, _ptr(out), _ptr(ssq))
    if not ok:
        raise RuntimeError("qtts_debug_bgemm: " + lib.qtts_create_error().decode())
    return (out, ssq) if want_ssq else out
