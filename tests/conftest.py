import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import qtts_pkg  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def pkg():
    return qtts_pkg.load()


@pytest.fixture(scope="session")
def gguf_dir(tmp_path_factory):
    d = os.environ.get("QTTS_GGUF_CACHE") or str(tmp_path_factory.mktemp("gguf"))
    os.makedirs(d, exist_ok=True)
    return d


@pytest.fixture(scope="session")
def make_gguf(pkg, gguf_dir):
    """Returns f(preset, seed) -> (path, SynthConfig); files are cached per session."""
    cache = {}

    def f(preset="tiny", seed=1234):
        key = (preset, seed)
        if key not in cache:
            path = os.path.join(gguf_dir, f"synth-{preset}-{seed}.gguf")
            cfg = pkg.synth.PRESETS[preset]
            if not os.path.exists(path):
                pkg.synth.write_gguf(path, preset, seed)
            cache[key] = (path, cfg)
        return cache[key]

    return f


def has_cuda() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def ref_mod():
    import ref
    if not ref.available():
        pytest.skip("oracle/_ref/libqtts_ref.so not built (make -f oracle/Makefile.ref)")
    return ref


def nmse(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))
