"""GGUF -> (config, tensors) for the hot path.

Mirrors `TTSTransformer::parse_config` (reference: src/tts_transformer.cpp:288-442): every
hyper-parameter is looked up under the same list of alternative keys, with the same defaults.
The tensor set is the one `create_tensors` maps (src/tts_transformer.cpp:444-731).
"""
from __future__ import annotations

import numpy as np

# (config field, [keys in lookup order], default or name of the talker field to fall back to)
_U32 = [
    ("text_vocab_size", ["qwen3-tts.text.vocab_size", "qwen3-tts.text_vocab_size"], 151936),
    ("text_embd_dim", ["qwen3-tts.text.embedding_dim", "qwen3-tts.text_hidden_size"], 2048),
    ("hidden_size", ["qwen3-tts.talker.embedding_length", "qwen3-tts.embedding_length"], 1024),
    ("n_layers", ["qwen3-tts.talker.block_count", "qwen3-tts.block_count"], 28),
    ("n_heads", ["qwen3-tts.talker.attention.head_count", "qwen3-tts.attention.head_count"], 16),
    ("n_kv_heads", ["qwen3-tts.talker.attention.head_count_kv", "qwen3-tts.attention.head_count_kv"], 8),
    ("intermediate_size", ["qwen3-tts.talker.feed_forward_length", "qwen3-tts.feed_forward_length"], 3072),
    ("head_dim", ["qwen3-tts.talker.attention.key_length", "qwen3-tts.attention.key_length"], 128),
    ("codec_vocab_size", ["qwen3-tts.talker.codec_vocab_size", "qwen3-tts.vocab_size"], 3072),
    ("n_codebooks", ["qwen3-tts.talker.num_codebooks", "qwen3-tts.num_code_groups"], 16),
    ("cp_layers", ["qwen3-tts.code_pred.layer_count", "qwen3-tts.code_predictor.layer_count"], 5),
    ("cp_vocab_size", ["qwen3-tts.code_pred.vocab_size", "qwen3-tts.code_predictor.vocab_size"], 2048),
    ("cp_hidden_size", ["qwen3-tts.code_predictor.embedding_length"], "hidden_size"),
    ("cp_intermediate_size", ["qwen3-tts.code_predictor.feed_forward_length"], "intermediate_size"),
    ("cp_n_heads", ["qwen3-tts.code_predictor.attention.head_count"], "n_heads"),
    ("cp_n_kv_heads", ["qwen3-tts.code_predictor.attention.head_count_kv"], "n_kv_heads"),
    ("cp_head_dim", ["qwen3-tts.code_predictor.attention.key_length"], "head_dim"),
    ("codec_pad_id", ["qwen3-tts.codec.pad_id"], 2148),
    ("codec_bos_id", ["qwen3-tts.codec.bos_id"], 2149),
    ("codec_eos_id", ["qwen3-tts.codec.eos_id", "qwen3-tts.codec.eos_token_id"], 2150),
    ("tts_bos_id", ["qwen3-tts.tts_bos_token_id", "qwen3-tts.tts.bos_token_id", "qwen3-tts.tts.bos_id"], 151672),
    ("tts_eos_id", ["qwen3-tts.tts_eos_token_id", "qwen3-tts.tts.eos_token_id", "qwen3-tts.tts.eos_id"], 151673),
    ("tts_pad_id", ["qwen3-tts.tts_pad_token_id", "qwen3-tts.tts.pad_token_id", "qwen3-tts.tts.pad_id"], 151671),
    ("codec_think_id", ["qwen3-tts.codec.think_id", "qwen3-tts.codec_think_id"], 2154),
    ("codec_nothink_id", ["qwen3-tts.codec.nothink_id", "qwen3-tts.codec_nothink_id"], 2155),
    ("codec_think_bos_id", ["qwen3-tts.codec.think_bos_id", "qwen3-tts.codec_think_bos_id"], 2156),
    ("codec_think_eos_id", ["qwen3-tts.codec.think_eos_id", "qwen3-tts.codec_think_eos_id"], 2157),
    ("english_language_id", ["qwen3-tts.language.english_id", "qwen3-tts.codec.language.english_id",
                             "qwen3-tts.language_id"], 2050),
]
_F32 = [
    ("rms_norm_eps", ["qwen3-tts.talker.attention.layer_norm_rms_epsilon",
                      "qwen3-tts.attention.layer_norm_rms_epsilon"], 1e-6),
    ("rope_theta", ["qwen3-tts.talker.rope.freq_base", "qwen3-tts.rope.freq_base"], 1000000.0),
]


def parse_config(fields) -> dict:
    """fields: mapping key -> python scalar (already decoded GGUF KV metadata)."""
    cfg = {}
    for name, keys, default in _U32:
        val = None
        for k in keys:
            if k in fields:
                val = int(fields[k])
                break
        if val is None:
            val = cfg[default] if isinstance(default, str) else default
        cfg[name] = val
    for name, keys, default in _F32:
        val = default
        for k in keys:
            if k in fields:
                val = float(fields[k])
                break
        cfg[name] = float(np.float32(val))
    return cfg


def read_gguf(path: str):
    """Returns (config dict, iterator factory over (name, dtype_code, ggml_dims, ndarray))."""
    import gguf  # offline package in the image

    rd = gguf.GGUFReader(path)
    fields = {}
    for key, f in rd.fields.items():
        if len(f.types) == 1 and f.types[0] in (gguf.GGUFValueType.UINT32, gguf.GGUFValueType.INT32,
                                                gguf.GGUFValueType.UINT64, gguf.GGUFValueType.INT64,
                                                gguf.GGUFValueType.FLOAT32, gguf.GGUFValueType.FLOAT64,
                                                gguf.GGUFValueType.UINT16, gguf.GGUFValueType.UINT8):
            fields[key] = f.parts[f.data[0]][0]
    cfg = parse_config(fields)

    def tensors():
        for t in rd.tensors:
            if not (t.name.startswith("talker.") or t.name.startswith("code_pred.")):
                continue   # spk_enc.* / tok_* skipped like src/tts_transformer.cpp:473
            if t.tensor_type == gguf.GGMLQuantizationType.F32:
                code = 0
            elif t.tensor_type == gguf.GGMLQuantizationType.F16:
                code = 1
            else:
                raise ValueError(f"tensor {t.name}: GGUF type {t.tensor_type.name} is not supported "
                                 "(only F32/F16; quantised types are rejected)")
            dims = [int(d) for d in t.shape]   # ggml order: dims[0] innermost
            yield t.name, code, dims, np.ascontiguousarray(t.data)

    return cfg, tensors
