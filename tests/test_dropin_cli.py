"""Compiled proof of the drop-in boundary (round-1 verdict, missing #3): the reference's OWN pipeline and CLI sources
(src/qwen3_tts.cpp, src/main.cpp, compiled where they lie by oracle/Makefile.dropin with cpp/dropin/qtts_dropin.h
force-included) linked against qtts_b200::TTSTransformer, next to the untouched reference CLI.  Both run on a synthetic
model directory (TTS GGUF with the synthetic tokenizer vocabulary + random-init vocoder GGUF, synth_vocoder.py).

CPU: the binaries exist, print the reference's usage text, the B200 one carries no reference TTSTransformer code and
fails loudly without a GPU through the reference's own error path.  GPU: same flags, same text -> the same WAV as
the reference CLI on the CPU (greedy, frames inside the exact-id prefix of the tiny fixture), one-shot and --serve."""
import os
import subprocess

import numpy as np
import pytest

from conftest import has_cuda

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "qwen3-tts-cli")
B200_CLI = os.path.join(ROOT, "oracle", "_ref", "qwen3-tts-cli-b200")


@pytest.fixture(scope="module")
def clis():
    if not (os.path.exists(REF_CLI) and os.path.exists(B200_CLI)):
        if not os.path.isdir("/root/reference/src"):
            pytest.skip("oracle/_ref CLIs not built and no reference tree to build them from")
        subprocess.run(["make", "-f", "oracle/Makefile.ref", "REF=/root/reference", "-j8"], cwd=ROOT, check=True, stdout=subprocess.DEVNULL)
        subprocess.run(["make", "-f", "oracle/Makefile.dropin", "REF=/root/reference", "-j8"], cwd=ROOT, check=True, stdout=subprocess.DEVNULL)
    return REF_CLI, B200_CLI


@pytest.fixture(scope="module")
def model_dir(pkg, tmp_path_factory):
    import importlib
    sv = importlib.import_module(pkg.__name__ + ".synth_vocoder")
    d = os.environ.get("QTTS_GGUF_CACHE") or str(tmp_path_factory.mktemp("mdl"))
    return sv.write_model_dir(os.path.join(d, "mdl_tiny"), "tiny", 1234)


def run(cli, model_dir, out, text="Hello, world!", extra=(), env=None, stdin=None):
    cmd = [cli, "-m", model_dir, "-o", out, "--temperature", "0", "--max-tokens", "8", *extra]
    if text is not None:
        cmd += ["-t", text]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, input=stdin)


def test_dropin_binaries_built_and_bound(clis):
    ref_cli, b200_cli = clis
    for cli in clis:
        r = subprocess.run([cli, "--help"], capture_output=True, text=True)
        assert "--max-tokens" in r.stdout + r.stderr and "--serve" in r.stdout + r.stderr
    nm = subprocess.run(["nm", "-C", b200_cli], capture_output=True, text=True).stdout
    assert "qwen3_tts::TTSTransformer::" not in nm, "the drop-in CLI still carries the reference's TTSTransformer"
    assert "qwen3_tts::Qwen3TTS::synthesize" in nm and "qtts_generate" in nm
    assert "qwen3_tts::TTSTransformer::generate" in subprocess.run(["nm", "-C", ref_cli], capture_output=True, text=True).stdout


def test_reference_cli_end_to_end_on_cpu(clis, model_dir, tmp_path):
    """BASELINE configs[0] in miniature: the untouched reference CLI, synthetic TTS + vocoder GGUFs, greedy, WAV out."""
    out = str(tmp_path / "ref.wav")
    r = run(clis[0], model_dir, out)
    assert r.returncode == 0 and os.path.getsize(out) > 44 + 2 * 1000, r.stderr[-800:]
    assert "Output saved to" in r.stdout + r.stderr


@pytest.mark.skipif(has_cuda(), reason="checks the no-GPU failure mode")
def test_dropin_cli_fails_loudly_without_gpu(clis, model_dir, tmp_path):
    r = run(clis[1], model_dir, str(tmp_path / "x.wav"))
    assert "no CUDA device: the B200 engine has no CPU fallback" in r.stdout + r.stderr
    assert not os.path.exists(tmp_path / "x.wav")


@pytest.mark.gpu
def test_dropin_cli_writes_the_reference_wav(clis, model_dir, tmp_path):
    a, b = str(tmp_path / "ref.wav"), str(tmp_path / "b200.wav")
    ra = run(clis[0], model_dir, a)
    rb = run(clis[1], model_dir, b)
    assert ra.returncode == 0, ra.stderr[-800:]
    assert rb.returncode == 0 and os.path.exists(b), (rb.stdout + rb.stderr)[-1200:]
    wa, wb = np.fromfile(a, np.uint8), np.fromfile(b, np.uint8)
    assert wa.size == wb.size and np.array_equal(wa, wb), "same text, same flags: the two CLIs must write the same audio"
    # server mode (src/main.cpp:108-163): the reference's line protocol over the B200 engine
    c = str(tmp_path / "serve.wav")
    rs = run(clis[1], model_dir, str(tmp_path / "unused.wav"), text=None, extra=("--serve",), stdin=f"Hello, world!\t{c}\nquit\n")
    assert rs.returncode == 0 and "OK\t" in rs.stdout, (rs.stdout + rs.stderr)[-1200:]
    assert np.array_equal(np.fromfile(c, np.uint8), wa)
