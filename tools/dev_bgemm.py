"""Bring-up helper: one batched-GEMM stage case per process (a CUDA fault poisons the context)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import qtts_pkg
pkg = qtts_pkg.load()
fmt, n_streams, n_tok, rows, K, epi, in_f16, norm = sys.argv[1], *map(int, sys.argv[2:9])
r = np.random.default_rng(1)
W = pkg.synth.bf16_round((0.05 * r.standard_normal((rows, K))).astype(np.float32))
X = r.standard_normal((n_streams * n_tok, K)).astype(np.float32)
if in_f16:
    X = X.astype(np.float16).astype(np.float32)
nw = (1.0 + 0.1 * r.standard_normal(K)).astype(np.float32) if norm else None
res = r.standard_normal((n_streams * n_tok, rows)).astype(np.float32) if epi == 1 else None
bias = r.standard_normal(rows).astype(np.float32) if epi == 3 else None
try:
    out = pkg.debug_bgemm(W, X, n_tok=n_tok, epi=epi, in_f16=bool(in_f16), norm_w=nw, res=res, bias=bias,
                          weight_fmt=pkg.BF16 if fmt == "bf16" else pkg.F16)
except RuntimeError as ex:
    print("CASE", sys.argv[1:], "dbg", os.environ.get("QTTS_BDBG"), "ERROR", ex)
    sys.exit(0)
Xn = X
if norm and not in_f16:
    ss = (X ** 2).astype(np.float64).sum(1)
    sc = (1.0 / np.sqrt((ss / K).astype(np.float32) + np.float32(1e-6))).astype(np.float32)
    Xn = (X * sc[:, None]) * nw[None, :]
A = Xn.astype(np.float16).astype(np.float64)
want = A @ W.astype(np.float64).T
if epi == 1: want += res
if epi == 3: want += bias[None, :]
if epi == 2:
    g, u = want[:, 0::2], want[:, 1::2]
    want = g / (1 + np.exp(-g)) * u
err = np.abs(out - want)
print("CASE", sys.argv[1:], "dbg", os.environ.get("QTTS_BDBG"), "max_err", float(err.max()), "max_ref", float(np.abs(want).max()),
      "bad_rows", int((err.max(1) > 1e-2).sum()), "bad_cols", int((err.max(0) > 1e-2).sum()))
if err.max() > 1e-2:
    br = np.where(err.max(1) > 1e-2)[0]; bc = np.where(err.max(0) > 1e-2)[0]
    print("  bad rows", br[:20], "bad cols", bc[:40])
    print("  out[0,:8]", out[0, :8], "want", want[0, :8])
