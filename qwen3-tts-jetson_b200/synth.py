"""Seeded random-init Qwen3-TTS GGUF files (synthetic weights of the real architecture).

The GGUF layout (tensor names, dtypes, metadata keys) follows the reference converter
`scripts/convert_tts_to_gguf.py:42-109,289-297,432-496` and the keys read by
`src/tts_transformer.cpp:288-442`: 2-D tensors are F16 `[out, in]` row-major, 1-D tensors F32.

Every matrix value is drawn N(0, std), rounded to a BF16-representable number and flushed to
zero below 2^-14, so the SAME real number is stored in the F16 file the reference reads and in
the BF16 copy the B200 engine streams.

The tokenizer is outside the hot path, but the reference's pipeline (`Qwen3TTS::load_models`,
`src/qwen3_tts.cpp:107-160`) loads it from the same GGUF, so the file carries a synthetic byte-level
vocabulary (`tokenizer.ggml.tokens`, no merges): id b < 256 is the GPT-2 byte-to-unicode character of
byte b, so the reference's `TextTokenizer::encode` (`src/text_tokenizer.cpp:234-291`) maps ASCII text to
one id per byte — exactly what `synth_tokens()` builds without a tokenizer — and `assistant`, `Ċ`,
`<|im_start|>`, `<|im_end|>` sit at the ids `encode_for_tts` (`:293-330`) expects.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, asdict

import numpy as np


@dataclass
class SynthConfig:
    name: str = "0.6b"
    text_vocab_size: int = 151936
    text_embd_dim: int = 2048
    hidden_size: int = 1024
    n_layers: int = 28
    n_heads: int = 16
    n_kv_heads: int = 8
    intermediate_size: int = 3072
    head_dim: int = 128
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1000000.0
    codec_vocab_size: int = 3072
    n_codebooks: int = 16
    cp_layers: int = 5
    cp_vocab_size: int = 2048
    cp_hidden_size: int = 1024
    cp_intermediate_size: int = 3072
    cp_n_heads: int = 16
    cp_n_kv_heads: int = 8
    cp_head_dim: int = 128
    tts_bos_id: int = 151672
    tts_eos_id: int = 151673
    tts_pad_id: int = 151671
    im_start_id: int = 151644
    im_end_id: int = 151645
    assistant_id: int = 77091
    newline_id: int = 198

    @property
    def has_mtp(self) -> bool:
        return self.cp_hidden_size != self.hidden_size


PRESETS = {
    # BASELINE.json configs[0..3]
    "0.6b": SynthConfig(),
    # BASELINE.json configs[4]; dims assumed (SURVEY.md §8d "Config 5"): Qwen3-1.7B-like talker,
    # 0.6B-size code predictor behind mtp_proj.
    "1.7b": SynthConfig(name="1.7b", hidden_size=2048, intermediate_size=6144),
    # small shapes for fast parity tests (same code paths: GQA 2:1, head_dim 128)
    "tiny": SynthConfig(name="tiny", text_vocab_size=4096, text_embd_dim=256, hidden_size=512,
                        n_layers=3, n_heads=4, n_kv_heads=2, intermediate_size=1024,
                        cp_layers=2, cp_hidden_size=512, cp_intermediate_size=1024,
                        cp_n_heads=4, cp_n_kv_heads=2,
                        tts_bos_id=4001, tts_eos_id=4002, tts_pad_id=4000,
                        im_start_id=4010, im_end_id=4011, assistant_id=4012, newline_id=10),
    # tiny with a code predictor narrower than the talker (exercises mtp_proj, the 1.7B path)
    "tiny-mtp": SynthConfig(name="tiny-mtp", text_vocab_size=4096, text_embd_dim=256,
                            hidden_size=512, n_layers=3, n_heads=4, n_kv_heads=2,
                            intermediate_size=1024, cp_layers=2, cp_hidden_size=256,
                            cp_intermediate_size=512, cp_n_heads=2, cp_n_kv_heads=1,
                            tts_bos_id=4001, tts_eos_id=4002, tts_pad_id=4000,
                            im_start_id=4010, im_end_id=4011, assistant_id=4012, newline_id=10),
}


def byte_to_unicode() -> list:
    """GPT-2's byte -> printable unicode character table (the reference's BYTE_TO_UNICODE, src/text_tokenizer.cpp:12-46)."""
    bs = list(range(33, 127)) + list(range(161, 173)) + list(range(174, 256))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    m = dict(zip(bs, cs))
    return [chr(m[b]) for b in range(256)]


def synth_vocab(cfg: "SynthConfig") -> list:
    """text_vocab_size token strings: byte characters at ids 0..255, the chat-frame specials at their ids, unique
    fillers elsewhere (no merges: BPE leaves every word as single characters)."""
    b2u = byte_to_unicode()
    toks = [f"<|filler_{i}|>" for i in range(cfg.text_vocab_size)]
    for b in range(256):
        toks[b] = b2u[b]
    toks[cfg.assistant_id] = "assistant"
    toks[cfg.newline_id] = b2u[10]              # "Ċ": the newline token is looked up by content (:157-160)
    toks[cfg.im_start_id] = "<|im_start|>"
    toks[cfg.im_end_id] = "<|im_end|>"
    return toks


def bf16_round(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even f32 -> bf16 -> f32, then flush |x| < 2^-14 to 0 (F16-normal range)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    u = (u + r) & np.uint32(0xFFFF0000)
    y = u.view(np.float32)
    y[np.abs(y) < 2.0 ** -14] = 0.0
    return y


def _matrix(rng: np.random.Generator, rows: int, cols: int, std: float) -> np.ndarray:
    w = rng.standard_normal((rows, cols), dtype=np.float32)
    w *= np.float32(std)
    return bf16_round(w).astype(np.float16)


def tensor_specs(cfg: SynthConfig):
    """Yield (name, kind, shape) in file order. kind: 'mat' (F16 2-D), 'norm' (F32 ones+noise), 'bias'."""
    H, E = cfg.hidden_size, cfg.text_embd_dim
    yield "talker.text_embd.weight", "mat", (cfg.text_vocab_size, E)
    yield "talker.text_proj.fc1.weight", "mat", (E, E)
    yield "talker.text_proj.fc1.bias", "bias", (E,)
    yield "talker.text_proj.fc2.weight", "mat", (H, E)
    yield "talker.text_proj.fc2.bias", "bias", (H,)
    yield "talker.codec_embd.weight", "mat", (cfg.codec_vocab_size, H)
    yield "talker.codec_head.weight", "mat", (cfg.codec_vocab_size, H)
    yield "talker.output_norm.weight", "norm", (H,)

    def block(prefix, h, nh, nkv, hd, ffn):
        yield f"{prefix}.attn_norm.weight", "norm", (h,)
        yield f"{prefix}.attn_q.weight", "mat", (nh * hd, h)
        yield f"{prefix}.attn_k.weight", "mat", (nkv * hd, h)
        yield f"{prefix}.attn_v.weight", "mat", (nkv * hd, h)
        yield f"{prefix}.attn_output.weight", "mat", (h, nh * hd)
        yield f"{prefix}.attn_q_norm.weight", "norm", (hd,)
        yield f"{prefix}.attn_k_norm.weight", "norm", (hd,)
        yield f"{prefix}.ffn_norm.weight", "norm", (h,)
        yield f"{prefix}.ffn_gate.weight", "mat", (ffn, h)
        yield f"{prefix}.ffn_up.weight", "mat", (ffn, h)
        yield f"{prefix}.ffn_down.weight", "mat", (h, ffn)

    for i in range(cfg.n_layers):
        yield from block(f"talker.blk.{i}", H, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim,
                         cfg.intermediate_size)
    for i in range(cfg.cp_layers):
        yield from block(f"code_pred.blk.{i}", cfg.cp_hidden_size, cfg.cp_n_heads,
                         cfg.cp_n_kv_heads, cfg.cp_head_dim, cfg.cp_intermediate_size)
    yield "code_pred.output_norm.weight", "norm", (cfg.cp_hidden_size,)
    if cfg.has_mtp:
        yield "code_pred.mtp_proj.weight", "mat", (cfg.cp_hidden_size, H)
        yield "code_pred.mtp_proj.bias", "bias", (cfg.cp_hidden_size,)
    for i in range(cfg.n_codebooks - 1):
        yield f"code_pred.codec_embd.{i}.weight", "mat", (cfg.cp_vocab_size, H)
    for i in range(cfg.n_codebooks - 1):
        yield f"code_pred.lm_head.{i}.weight", "mat", (cfg.cp_vocab_size, cfg.cp_hidden_size)


def write_gguf(path: str, preset: str | SynthConfig = "0.6b", seed: int = 1234,
               std: float = 0.02) -> SynthConfig:
    """Write a seeded random-init GGUF; returns the config used."""
    import gguf  # offline package in this image; only needed to WRITE synthetic files

    cfg = PRESETS[preset] if isinstance(preset, str) else preset
    arch = "qwen3-tts"
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    w = gguf.GGUFWriter(path, arch)
    w.add_name(f"qwen3-tts-synth-{cfg.name}-seed{seed}")
    # keys exactly as scripts/convert_tts_to_gguf.py:459-494 writes them
    w.add_block_count(cfg.n_layers)
    w.add_embedding_length(cfg.hidden_size)
    w.add_feed_forward_length(cfg.intermediate_size)
    w.add_head_count(cfg.n_heads)
    w.add_head_count_kv(cfg.n_kv_heads)
    w.add_key_length(cfg.head_dim)
    w.add_value_length(cfg.head_dim)
    w.add_rope_freq_base(cfg.rope_theta)
    w.add_layer_norm_rms_eps(cfg.rms_norm_eps)
    w.add_vocab_size(cfg.codec_vocab_size)
    w.add_uint32(f"{arch}.text_vocab_size", cfg.text_vocab_size)
    w.add_uint32(f"{arch}.text_hidden_size", cfg.text_embd_dim)
    w.add_uint32(f"{arch}.num_code_groups", cfg.n_codebooks)
    w.add_uint32(f"{arch}.code_predictor.layer_count", cfg.cp_layers)
    w.add_uint32(f"{arch}.code_predictor.vocab_size", cfg.cp_vocab_size)
    w.add_uint32(f"{arch}.code_predictor.embedding_length", cfg.cp_hidden_size)
    w.add_uint32(f"{arch}.code_predictor.feed_forward_length", cfg.cp_intermediate_size)
    w.add_uint32(f"{arch}.code_predictor.attention.head_count", cfg.cp_n_heads)
    w.add_uint32(f"{arch}.code_predictor.attention.head_count_kv", cfg.cp_n_kv_heads)
    w.add_uint32(f"{arch}.code_predictor.attention.key_length", cfg.cp_head_dim)
    w.add_uint32(f"{arch}.codec.pad_id", 2148)
    w.add_uint32(f"{arch}.codec.bos_id", 2149)
    w.add_uint32(f"{arch}.codec.eos_id", 2150)
    w.add_uint32(f"{arch}.tts_bos_token_id", cfg.tts_bos_id)
    w.add_uint32(f"{arch}.tts_eos_token_id", cfg.tts_eos_id)
    w.add_uint32(f"{arch}.tts_pad_token_id", cfg.tts_pad_id)
    # synthetic byte-level vocabulary for the reference's TextTokenizer (src/text_tokenizer.cpp:80-165)
    w.add_array("tokenizer.ggml.tokens", synth_vocab(cfg))
    w.add_uint32("tokenizer.ggml.bos_token_id", cfg.im_start_id)
    w.add_uint32("tokenizer.ggml.eos_token_id", cfg.im_end_id)

    rng = np.random.Generator(np.random.PCG64(seed))
    for name, kind, shape in tensor_specs(cfg):
        if kind == "mat":
            t = _matrix(rng, shape[0], shape[1], std)
        elif kind == "norm":
            # ones + small noise so a wrong norm vector is detectable; BF16-exact values
            t = bf16_round(1.0 + 0.05 * rng.standard_normal(shape, dtype=np.float32))
        else:
            t = bf16_round(std * rng.standard_normal(shape, dtype=np.float32))
        w.add_tensor(name, t)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return cfg


def synth_tokens(cfg: SynthConfig, text: str | None = None, n_text: int | None = None,
                 seed: int = 0) -> np.ndarray:
    """`encode_for_tts` frame (src/text_tokenizer.cpp:293-330) around synthetic text ids.

    text -> one id per UTF-8 byte; otherwise n_text seeded ids in [32, 1000).
    """
    if text is not None:
        body = [b for b in text.encode("utf-8")]
    else:
        r = np.random.Generator(np.random.PCG64(seed))
        body = r.integers(32, 1000, size=int(n_text)).tolist()
    ids = ([cfg.im_start_id, cfg.assistant_id, cfg.newline_id] + body +
           [cfg.im_end_id, cfg.newline_id, cfg.im_start_id, cfg.assistant_id, cfg.newline_id])
    return np.asarray(ids, dtype=np.int32)


def synth_speaker(cfg: SynthConfig, seed: int = 7) -> np.ndarray:
    """speaker.bin payload: raw float32[hidden] (src/main.cpp:38-59); seeded N(0,1)*0.1."""
    r = np.random.Generator(np.random.PCG64(seed))
    return (0.1 * r.standard_normal(cfg.hidden_size)).astype(np.float32)


def config_dict(cfg: SynthConfig) -> dict:
    return asdict(cfg)


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--preset", default="0.6b", choices=sorted(PRESETS))
    ap.add_argument("--seed", type=int, default=1234)
    a = ap.parse_args()
    c = write_gguf(a.out, a.preset, a.seed)
    print(f"wrote {a.out} ({os.path.getsize(a.out) / 1e6:.1f} MB) preset={c.name}")
