"""Experiment: how far do converged fields move when the 1-D operators are replaced by their exactly
centro-(anti)symmetric parts?  (Pre-study for the even-odd DMMA kernel: it halves the contraction flops but
only sees the symmetric part; the reference's matrices are symmetric to rounding only.)"""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200
from ldc_b200 import _lib, operators as ops

G = np.load("tests/golden/solves.npz")


def sym(A, sign):
    """centro-symmetric (sign=+1) or centro-antisymmetric (sign=-1) part; for n x m (interpolation) too."""
    return 0.5 * (A + sign * A[::-1, ::-1])


def set_sym_ops(s):
    lv = s.level
    Dx, Dy = sym(lv.Dx, -1), sym(lv.Dy, -1)
    Dxx, Dyy = sym(lv.Dxx, +1), sym(lv.Dyy, +1)
    Ix, Iy = sym(lv.Ix, +1), sym(lv.Iy, +1)
    keep = [np.ascontiguousarray(a) for a in (Dx, Dxx, Dy, Dyy, Ix, Iy)]
    _lib.check(s.lib.ldc_set_operators(s._h, *[_lib.ptr(a) for a in keep], lv.dx_min, lv.dy_min), "set_operators")
    return max(np.abs(lv.Dx - Dx).max() / np.abs(lv.Dx).max(), np.abs(lv.Dxx - Dxx).max() / np.abs(lv.Dxx).max())


def fsg(N, Re, cs, symmetrise, max_iter):
    orders = ops.hierarchy_orders(N, 2)
    prev, total, asym = None, 0, 0.0
    for n in orders:
        lvl = ldc_b200.BatchedSpectralLDC(n, batch=1)
        if symmetrise:
            asym = max(asym, set_sym_ops(lvl))
        lvl.set_cases(Re, corner_smoothing=cs)
        if prev is None:
            lvl.init_state()
        else:
            prev.prolongate_to(lvl, "fft", bc_quirk=True)
            prev.close()
        r = lvl.solve(1e-6, max_iter, warmup=0, freeze_pressure=False, nan_exit=True)
        total += int(r.iterations[0])
        prev = lvl
    u, v, p = prev.get_state()
    prev.close()
    return u[0], v[0], p[0], total, asym


rel = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
out = {}
for tag, N, Re, cs, mi in (("fsg_Re100_N32", 32, 100.0, 0.15, 10**7), ("fsg_Re1000_N30_cs015", 30, 1000.0, 0.15, 500000),
                           ("fsg_Re400_N64", 64, 400.0, 0.15, 10**7)):
    for symm in (False, True):
        u, v, p, it, asym = fsg(N, Re, cs, symm, mi)
        out[f"{tag}_{'sym' if symm else 'asis'}"] = dict(iters=it, ref_iters=int(G[f"{tag}_meta"][0]), rel_u=rel(u, G[f"{tag}_u"]),
                                                        rel_v=rel(v, G[f"{tag}_v"]), rel_p=rel(p, G[f"{tag}_p"]), op_asym=asym)
        print(tag, symm, out[f"{tag}_{'sym' if symm else 'asis'}"], flush=True)
json.dump(out, open("gpurun_out/exp_symmetrised.json", "w"), indent=1)
