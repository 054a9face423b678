"""A/B builds of libqtts_b200.so: same sources, extra -D switches, written to build_ab/ (git-ignored, travels to the
GPU box).  usage: python tools/ab_build.py name=-DFOO=0,-DBAR=1 ...   then QTTS_LIB_PATH=build_ab/libqtts_<name>.so"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import qtts_pkg
b = qtts_pkg.load().build
out = os.path.join(ROOT, "build_ab"); os.makedirs(out, exist_ok=True)
procs = []
for a in sys.argv[1:]:
    name, _, flags = a.partition("=")
    lib = os.path.join(out, f"libqtts_{name}.so")
    cmd = ["nvcc"] + [f for f in flags.split(",") if f] + b.NVCC_FLAGS + ["-o", lib, os.path.join(b.CSRC, "libqtts_b200.cu")]
    procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
for name, p in procs:
    o, _ = p.communicate()
    print(name, "rc", p.returncode, o[-400:] if p.returncode else "")
