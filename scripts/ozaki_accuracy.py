"""How many int8 slices would an integer-tensor-core emulation of the FP64 block products need?  

NumPy model of the Ozaki splitting for the products the N=64 stepping kernel forms: A = a parity block of the first- or
second-derivative operator (rows scaled to a common exponent), B = a parity block of a converged-looking state (columns
scaled), both cut into signed 7-bit slices; slice pairs (i, j) with i + j < s are multiplied exactly (int32 accumulation is
exact for K = 32) and recombined in FP64.  Prints the relative error of the product against the FP64 product for s = 4..10.
CPU only; nothing here is on the product path.
"""
import sys

import numpy as np

sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
from ldc_b200 import operators as ops


def slices(M, axis, n):
    """Scale each row (axis=1) / column (axis=0) by a power of two so that |x| < 1, then n signed 7-bit digits."""
    amax = np.max(np.abs(M), axis=axis, keepdims=True)
    e = np.ceil(np.log2(np.where(amax > 0, amax, 1.0))) + 1
    r = M / 2.0**e
    out = []
    for _ in range(n):
        r = r * 128.0
        q = np.trunc(r)
        out.append(q)
        r = r - q
    return out, e


def emulate(A, B, s):
    As, ea = slices(A, 1, s)
    Bs, eb = slices(B, 0, s)
    C = np.zeros((A.shape[0], B.shape[1]))
    for g in range(s - 1, -1, -1):          # small terms first
        acc = np.zeros_like(C)
        for i in range(g + 1):
            acc += As[i] @ Bs[g - i]        # exact in int32 on the device (|entries| <= 127, K = 32)
        C += acc * 2.0 ** (-7 * (g + 2))
    return C * 2.0**ea * 2.0**eb


N = 64
lv = ops.build_level(N, N)
m, h = N - 1, N // 2
x, y = np.meshgrid(lv.x[1:-1], lv.y[1:-1], indexing="ij")
U = np.sin(np.pi * x) ** 2 * np.sin(2 * np.pi * y) + 1e-3 * np.cos(7 * x * y)      # smooth field + small-scale content
D, DD = lv.Dx[1:-1, 1:-1], lv.Dxx[1:-1, 1:-1]
J = np.eye(m)[::-1]
for name, Op in (("D", D), ("DD", DD)):
    # parity blocks as the kernels use them: rows < h of (Op +- Op J), columns < h of (U +- J U)
    A = (Op + Op @ J)[:h, :h]
    Bm = 0.5 * (U - J @ U)[:h, :h]
    ref = A @ Bm
    print(name, "block product 32^3:", " ".join(f"s={s}: {np.linalg.norm(emulate(A, Bm, s) - ref) / np.linalg.norm(ref):.1e}" for s in range(4, 11)))
