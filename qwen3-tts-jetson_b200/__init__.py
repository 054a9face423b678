"""qwen3-tts-jetson_b200 — B200-native (sm_100a) Qwen3-TTS codec-frame engine.

Only the hot path of syb0rg/qwen3-tts-jetson lives here: csrc/ (CUDA kernels + C ABI) and the
host-side mirror of the reference's TTSTransformer interface (transformer.py).  The directory
name is not a Python identifier; import it through `qtts_pkg.load()` at the repo root.
"""
from .transformer import (TTSTransformer, load_library, debug_bgemm, EXPORTS, NUMERICS_FAST,  # noqa: F401
                          NUMERICS_REFERENCE, F32, F16, BF16, EPI_STORE, EPI_ADD, EPI_SWIGLU, EPI_BIAS)
from . import build, gguf_io, serve, sharding, synth, synth_vocoder  # noqa: F401
