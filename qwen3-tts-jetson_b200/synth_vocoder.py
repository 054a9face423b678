"""Seeded random-init vocoder GGUF (`qwen3-tts-tokenizer-f16.gguf`) for the reference's GGML audio decoder.

The vocoder is OUTSIDE the optimised path (north_star: "only streamed"); this file exists so that the reference's
own `qwen3-tts-cli` — and the same CLI built on the B200 engine (oracle/Makefile.dropin) — runs end to end on
synthetic weights (BASELINE configs[0] is the literal CLI run).  Nothing here is used by
the engine.

Tensor names are the ones `AudioTokenizerDecoder::load_model` maps (src/audio_tokenizer_decoder.cpp:75-330; written by
scripts/convert_tokenizer_to_gguf.py:40-110); shapes follow from how `build_graph` uses each tensor
(src/audio_tokenizer_decoder.cpp:405-800) with the decoder's fixed configuration (src/audio_tokenizer_decoder.h:14-28:
16 codebooks x 2048 x 256, hidden 512, latent 1024, 8 pre-transformer layers, decoder 1536 -> 96 channels, upsample
8 x 5 x 4 x 3 after two 2x ConvNeXt blocks = 1920 samples per frame).  numpy shapes below are the REVERSE of ggml's ne
order.  Convolution kernels are F16 (ggml_conv_1d's im2col path), matrices F16, vectors F32.
"""
from __future__ import annotations

import os

import numpy as np

CODEBOOKS, CB_SIZE, CB_DIM = 16, 2048, 256
HIDDEN, LATENT, FFN, N_TFM = 512, 1024, 1024, 8
DEC_DIM, RATES = 1536, (8, 5, 4, 3)
PW_EXPAND = 4096


def tensor_specs():
    """Yield (name, kind, numpy shape); kind: 'w' F16 N(0, std) | 'v' F32 N(0, small) | 'one' F32 ones | 'usage' F32 ones."""
    p = "tok_dec."
    yield p + "vq_first.0.codebook", "w", (CB_SIZE, CB_DIM)
    yield p + "vq_first.0.usage", "usage", (CB_SIZE,)
    yield p + "vq_first.output_proj.weight", "w", (HIDDEN, CB_DIM, 1)
    for i in range(CODEBOOKS - 1):
        yield p + f"vq_rest.{i}.codebook", "w", (CB_SIZE, CB_DIM)
        yield p + f"vq_rest.{i}.usage", "usage", (CB_SIZE,)
    yield p + "vq_rest.output_proj.weight", "w", (HIDDEN, CB_DIM, 1)
    yield p + "pre_conv.weight", "w", (LATENT, HIDDEN, 3)
    yield p + "pre_conv.bias", "v", (LATENT,)
    yield p + "pre_tfm.input_proj.weight", "w", (HIDDEN, LATENT)
    yield p + "pre_tfm.input_proj.bias", "v", (HIDDEN,)
    for i in range(N_TFM):
        b = p + f"pre_tfm.blk.{i}."
        yield b + "attn_norm.weight", "one", (HIDDEN,)
        yield b + "attn_q.weight", "w", (LATENT, HIDDEN)
        yield b + "attn_k.weight", "w", (LATENT, HIDDEN)
        yield b + "attn_v.weight", "w", (LATENT, HIDDEN)
        yield b + "attn_output.weight", "w", (HIDDEN, LATENT)
        yield b + "attn_scale", "one", (HIDDEN,)
        yield b + "ffn_norm.weight", "one", (HIDDEN,)
        yield b + "ffn_gate.weight", "w", (FFN, HIDDEN)
        yield b + "ffn_up.weight", "w", (FFN, HIDDEN)
        yield b + "ffn_down.weight", "w", (HIDDEN, FFN)
        yield b + "ffn_scale", "one", (HIDDEN,)
    yield p + "pre_tfm.norm.weight", "one", (HIDDEN,)
    yield p + "pre_tfm.output_proj.weight", "w", (LATENT, HIDDEN)
    yield p + "pre_tfm.output_proj.bias", "v", (LATENT,)
    for i in range(2):
        b = p + f"upsample.{i}."
        yield b + "conv.weight", "w", (LATENT, LATENT, 2)           # conv_transpose_1d, stride 2: ne [K, OC, IC]
        yield b + "conv.bias", "v", (LATENT,)
        yield b + "dwconv.weight", "w", (LATENT, 1, 7)
        yield b + "dwconv.bias", "v", (LATENT,)
        yield b + "norm.weight", "one", (LATENT,)
        yield b + "norm.bias", "v", (LATENT,)
        yield b + "pwconv1.weight", "w", (PW_EXPAND, LATENT)
        yield b + "pwconv1.bias", "v", (PW_EXPAND,)
        yield b + "pwconv2.weight", "w", (LATENT, PW_EXPAND)
        yield b + "pwconv2.bias", "v", (LATENT,)
        yield b + "gamma", "v", (LATENT,)
    yield p + "dec.0.conv.weight", "w", (DEC_DIM, LATENT, 7)
    yield p + "dec.0.conv.bias", "v", (DEC_DIM,)
    c_in = DEC_DIM
    for i, rate in enumerate(RATES, start=1):
        c_out = c_in // 2
        b = p + f"dec.{i}."
        yield b + "snake.alpha", "v", (c_in,)
        yield b + "snake.beta", "v", (c_in,)
        yield b + "conv_t.weight", "w", (c_in, c_out, 2 * rate)     # ne [K = 2 * stride, OC, IC]
        yield b + "conv_t.bias", "v", (c_out,)
        for r in (2, 3, 4):
            rb = b + f"res.{r}."
            yield rb + "act1.alpha", "v", (c_out,)
            yield rb + "act1.beta", "v", (c_out,)
            yield rb + "conv1.weight", "w", (c_out, c_out, 7)
            yield rb + "conv1.bias", "v", (c_out,)
            yield rb + "act2.alpha", "v", (c_out,)
            yield rb + "act2.beta", "v", (c_out,)
            yield rb + "conv2.weight", "w", (c_out, c_out, 1)
            yield rb + "conv2.bias", "v", (c_out,)
        c_in = c_out
    yield p + "dec.5.snake.alpha", "v", (c_in,)
    yield p + "dec.5.snake.beta", "v", (c_in,)
    yield p + "dec.6.conv.weight", "w", (1, c_in, 7)
    yield p + "dec.6.conv.bias", "v", (1,)


def write_vocoder_gguf(path: str, seed: int = 1234, std: float = 0.02) -> int:
    """Writes the file; returns the number of tensors."""
    import gguf
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    w = gguf.GGUFWriter(path, "qwen3-tts")
    w.add_name(f"qwen3-tts-synth-vocoder-seed{seed}")
    w.add_uint32("qwen3-tts.tokenizer.sample_rate", 24000)
    w.add_uint32("qwen3-tts.tokenizer.num_codebooks", CODEBOOKS)
    w.add_uint32("qwen3-tts.tokenizer.codebook_size", CB_SIZE)
    rng = np.random.Generator(np.random.PCG64(seed))
    n = 0
    for name, kind, shape in tensor_specs():
        if kind == "w":
            fan_in = int(np.prod(shape[1:]))
            t = (rng.standard_normal(shape, dtype=np.float32) * np.float32(max(std, 1.0 / np.sqrt(fan_in)))).astype(np.float16)
        elif kind == "v":
            t = (0.01 * rng.standard_normal(shape, dtype=np.float32)).astype(np.float32)
        else:
            t = np.ones(shape, np.float32)
        w.add_tensor(name, t)
        n += 1
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return n


def write_model_dir(model_dir: str, preset: str = "0.6b", seed: int = 1234) -> str:
    """The directory layout `Qwen3TTS::load_models` expects (hard-coded names, src/qwen3_tts.cpp:117-118)."""
    from . import synth
    os.makedirs(model_dir, exist_ok=True)
    tts = os.path.join(model_dir, "qwen3-tts-0.6b-f16.gguf")
    voc = os.path.join(model_dir, "qwen3-tts-tokenizer-f16.gguf")
    if not os.path.exists(tts):
        synth.write_gguf(tts, preset, seed)
    if not os.path.exists(voc):
        write_vocoder_gguf(voc, seed)
    return model_dir


if __name__ == "__main__":
    import sys
    out = sys.argv[1] if len(sys.argv) > 1 else "qwen3-tts-tokenizer-f16.gguf"
    print(f"wrote {out}: {write_vocoder_gguf(out)} tensors, {os.path.getsize(out) / 1e6:.1f} MB")
