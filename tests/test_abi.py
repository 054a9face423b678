"""CPU tests of the drop-in boundary and the host logic: the C-ABI library loads, exports every
symbol include/qtts_b200.h declares, fails loudly without a GPU (no CPU fallback), and the GGUF
config parser follows the reference's key fall-back lists (src/tts_transformer.cpp:288-442)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_cuda

HEADER = os.path.join(ROOT, "include", "qtts_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(qtts_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_symbols_exported(pkg):
    lib = pkg.load_library()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"libqtts_b200.so does not export {s}"
    assert sorted(pkg.EXPORTS) == syms, "transformer.EXPORTS out of sync with include/qtts_b200.h"


def test_library_is_sm100a_only(pkg):
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", pkg.build.LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


@pytest.mark.skipif(has_cuda(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(pkg, make_gguf):
    path, _ = make_gguf("tiny", 1234)
    T = pkg.TTSTransformer()
    assert T.load_model(path) is False
    assert "no CUDA device" in T.get_error() and "no CPU fallback" in T.get_error()
    assert not T.is_loaded()


def test_load_model_bad_path(pkg):
    T = pkg.TTSTransformer()
    assert T.load_model("/nonexistent/model.gguf") is False
    assert "Failed to open GGUF file" in T.get_error()


def test_parse_config_key_fallbacks(pkg):
    pc = pkg.gguf_io.parse_config
    d = pc({})
    assert (d["hidden_size"], d["n_layers"], d["codec_vocab_size"], d["cp_layers"]) == (1024, 28, 3072, 5)
    assert d["cp_hidden_size"] == d["hidden_size"] and d["codec_eos_id"] == 2150
    # first spelling wins over the second; second is what the converter writes
    d = pc({"qwen3-tts.talker.embedding_length": 2048, "qwen3-tts.embedding_length": 512,
            "qwen3-tts.block_count": 3, "qwen3-tts.code_predictor.embedding_length": 1024})
    assert d["hidden_size"] == 2048 and d["n_layers"] == 3 and d["cp_hidden_size"] == 1024
    assert d["cp_intermediate_size"] == d["intermediate_size"]


def test_synth_gguf_roundtrip(pkg, make_gguf):
    path, cfg = make_gguf("tiny-mtp", 1234)
    conf, tensors = pkg.gguf_io.read_gguf(path)
    assert conf["hidden_size"] == cfg.hidden_size and conf["cp_hidden_size"] == cfg.cp_hidden_size
    names = {}
    for name, code, dims, arr in tensors():
        names[name] = (code, dims)
        if code == 1:   # F16 matrices hold BF16-representable values (synth.bf16_round)
            f = arr.astype(np.float32).view(np.uint32)
            assert (f & 0xFFFF).max() == 0
    assert names["code_pred.mtp_proj.weight"] == (1, [cfg.hidden_size, cfg.cp_hidden_size])
    assert names["talker.blk.0.attn_norm.weight"][0] == 0
    want = {n for n, _, _ in pkg.synth.tensor_specs(cfg)}
    assert want == set(names)


def test_synth_tokens_frame(pkg):
    cfg = pkg.synth.PRESETS["0.6b"]
    t = pkg.synth.synth_tokens(cfg, "Hello, world!")
    # encode_for_tts frame: src/text_tokenizer.cpp:293-330 (tests/test_tokenizer.cpp:14 head/tail)
    assert t[:3].tolist() == [151644, 77091, 198] and t[-5:].tolist() == [151645, 198, 151644, 77091, 198]
    assert len(t) == 21


def test_quantised_gguf_is_rejected_by_name(pkg, tmp_path):
    """N2: scripts/convert_tts_to_gguf.py can emit Q8_0 / Q4_0 / Q4_K; the loaders reject such a tensor with its name
    and type instead of reading garbage (Python metadata reader, C++ GGUF reader used by the library's loader)."""
    import gguf
    import subprocess
    path = str(tmp_path / "q8.gguf")
    w = gguf.GGUFWriter(path, "qwen3-tts")
    w.add_block_count(1)
    w.add_tensor("talker.output_norm.weight", np.ones(64, np.float32))
    q = np.zeros((8, 34 * 2), np.uint8)                    # 8 rows x 64 values of Q8_0: 2 blocks of 34 bytes per row
    w.add_tensor("talker.codec_head.weight", q, raw_dtype=gguf.GGMLQuantizationType.Q8_0)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    cfg, tensors = pkg.gguf_io.read_gguf(path)
    with pytest.raises(ValueError, match="talker.codec_head.weight.*Q8_0"):
        list(tensors())
    exe = pkg.build.build_adapter_check()
    r = subprocess.run([exe, "--tensors", path], capture_output=True, text=True)
    assert r.returncode == 1 and "talker.codec_head.weight" in r.stderr and "Q8_0" in r.stderr and "rejected" in r.stderr
