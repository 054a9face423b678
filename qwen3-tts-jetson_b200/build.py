"""Builds libqtts_b200.so (sm_100a only) in-tree with nvcc.  Used by __graft_entry__.build()."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libqtts_b200.so")
NVCC_FLAGS = (os.environ.get("QTTS_NVCC_EXTRA", "").split()) + ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")))


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    hdr = os.path.join(HERE, "..", "include", "qtts_b200.h")
    return any(os.path.getmtime(s) > t for s in sources() + [hdr])


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB, os.path.join(CSRC, "libqtts_b200.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


CPP = os.path.join(HERE, "..", "cpp")
ADAPTER_CHECK = os.path.join(CPP, "adapter_check")


def build_adapter_check(force: bool = False) -> str:
    """g++ build of the C++ host adapter's driver (cpp/adapter_check.cpp) against the C-ABI library."""
    srcs = [os.path.join(CPP, f) for f in ("adapter_check.cpp", "qtts_transformer.hpp", "qtts_gguf.hpp")]
    if not force and os.path.exists(ADAPTER_CHECK) and \
            all(os.path.getmtime(s) <= os.path.getmtime(ADAPTER_CHECK) for s in srcs + [LIB]):
        return ADAPTER_CHECK
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", srcs[0], "-o", ADAPTER_CHECK,
           "-L" + HERE, "-lqtts_b200", "-Wl,-rpath,$ORIGIN/../" + os.path.basename(HERE)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return ADAPTER_CHECK


if __name__ == "__main__":
    print(build(force=True, verbose=True))
