"""C++ host adapter (cpp/qtts_transformer.hpp): the class surface of the reference's TTSTransformer
(src/tts_transformer.h:164-245) and TRTCodePredictor loops (src/trt_code_predictor.h:61-79) over the
C ABI.  CPU: it compiles, links, parses GGUF like the Python host and fails loudly without a GPU.
GPU: generate(), the frame callback and the code-predictor loops return what the ctypes path returns."""
import json
import subprocess

import numpy as np
import pytest

from conftest import has_cuda


@pytest.fixture(scope="module")
def exe(pkg):
    pkg.build.build()
    return pkg.build.build_adapter_check()


def test_adapter_parses_gguf_like_python_host(pkg, exe, make_gguf):
    for preset in ("tiny", "tiny-mtp"):
        path, _ = make_gguf(preset, 1234)
        r = subprocess.run([exe, "--config", path], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        got = json.loads(r.stdout)
        cfg, tensors = pkg.gguf_io.read_gguf(path)
        for k, v in got.items():
            if k in cfg:
                assert v == pytest.approx(cfg[k], rel=1e-6), k
        ts = list(tensors())
        assert got["n_tensors"] >= len(ts)
        name, _, _, arr = ts[0]
        if got["first_tensor"] == name:
            assert got["first_tensor_bytesum"] == int(arr.view(np.uint8).astype(np.uint64).sum())


@pytest.mark.skipif(has_cuda(), reason="checks the no-GPU failure mode")
def test_adapter_fails_loudly_without_gpu(exe, make_gguf, tmp_path):
    path, _ = make_gguf("tiny", 1234)
    tok = tmp_path / "tok.i32"
    np.arange(10, 22, dtype=np.int32).tofile(tok)
    r = subprocess.run([exe, path, str(tok), "8", "0", "0", "1"], capture_output=True, text=True)
    assert r.returncode == 3
    assert "no CPU fallback" in json.loads(r.stdout)["load_error"]


@pytest.mark.gpu
@pytest.mark.parametrize("preset,temperature", [("tiny", 0.0), ("tiny-mtp", 0.0), ("tiny", 0.9)])
def test_adapter_matches_ctypes_path(pkg, exe, make_gguf, tmp_path, preset, temperature):
    path, _ = make_gguf(preset, 1234)
    tokens = np.arange(10, 22, dtype=np.int32)
    tok = tmp_path / "tok.i32"
    tokens.tofile(tok)
    max_len, top_k, seed = 10, 50, 77
    r = subprocess.run([exe, path, str(tok), str(max_len), str(temperature), str(top_k), str(seed)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    got = json.loads(r.stdout)

    T = pkg.TTSTransformer(device=0, max_streams=1, max_ctx=4096)
    assert T.load_model(path), T.get_error()
    want = T.generate(tokens, None, max_len=max_len, temperature=temperature, top_k=top_k, seed=seed)
    assert want is not None, T.get_error()
    nb = got["n_codebooks"]
    codes = np.array(got["codes"], np.int32).reshape(-1, nb)
    assert np.array_equal(codes, want)
    assert got["n_frames"] == len(want)
    # callback_interval 4: every full group of 4 frames plus the remainder (src/tts_transformer.cpp:2540-2560)
    assert got["cb_frames"] == len(want) and got["cb_calls"] == -(-len(want) // 4)
    assert got["bad_width_rejected"] is True
    assert got["greedy15"] == got["predict15"]
    # TRTCodePredictor::step / step_with_token position by position (teacher-forced on the loop's codes): the same ids
    # except where the single-token attention of positions 0/1 (F16 cache) meets a near-tie of the merged F32 pass
    same = sum(int(a == b) for a, b in zip(got["stepwise15"], got["greedy15"]))
    assert len(got["stepwise15"]) == 15 and same >= 13, (got["stepwise15"], got["greedy15"])
    assert got["upload_validation"] is True
    # a request beyond the context cap fails without unloading the model (round-1 advisor finding)
    assert got["too_long_fails"] is True and got["still_loaded"] is True
    second = np.array(got["codes_second_call"], np.int32).reshape(-1, nb)
    if temperature > 0.0:
        assert not np.array_equal(second, codes), "a second generate() replayed the first call's uniforms"
    else:
        assert np.array_equal(second, codes)
    if temperature == 0.0:
        # first frame by hand (prefill -> greedy cb0 -> code predictor loop) == first frame of generate()
        assert got["cb0"] == int(want[0, 0])
        assert got["greedy15"] == [int(v) for v in want[0, 1:]]
    T.unload_model()
