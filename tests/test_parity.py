"""Even-odd (parity) algebra of the `eo` kernel (tests/parity_model.py) against the oracle residual and the
reference-minted residual fixture: the formulation is pinned on the CPU before the CUDA kernel is trusted."""
import numpy as np
import pytest

from ldc_b200 import operators as ops
import parity_model as par
from oracle import ldc_oracle as orc

rel = lambda a, b: np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("m", [1, 2, 7, 14, 29, 63])
def test_split_merge_roundtrip_and_centre(m):
    rng = np.random.default_rng(m)
    F = rng.standard_normal((3, m, m))
    blocks = par.split_field(F)
    assert blocks.shape == (3, 2, 2, par.half(m), par.half(m))
    assert np.allclose(par.merge_field(blocks, m), F, atol=1e-15)
    if m % 2:  # the self-mirrored line has no odd part
        c = par.half(m) - 1
        assert not blocks[:, 1, :, c, :].any() and not blocks[:, :, 1, :, c].any()
    f = rng.standard_normal(m)
    S, A = par.split_vector(f)
    assert np.allclose(par.inverse(m) @ np.concatenate([S, A]), f, atol=1e-15)


@pytest.mark.parametrize("N", [12, 15, 16, 30, 64])
def test_operator_asymmetry_is_rounding_only(N):
    lv = ops.build_level(N, N)
    G = (lv.Dx @ lv.Ix)[1:-1, :]
    for M, symm in ((lv.Dx[1:-1, 1:-1], False), (lv.Dxx[1:-1, 1:-1], True), (G, False)):
        assert par.operator_blocks(M, symm)[2] < 1e-12


@pytest.mark.parametrize("N", [12, 15, 16, 30, 64])
def test_parity_residual_matches_oracle(N):
    rng = np.random.default_rng(N)
    n, m = N + 1, N - 1
    u, v = rng.standard_normal((2, n, n))
    p = rng.standard_normal((m, m))
    Ru, Rv, Rp = orc.residual(orc.build_level(N), u, v, p, 400.0, 5.0)
    ru, rv, rp = par.residual_interior_parity(ops.build_level(N, N), u, v, p, 400.0, 5.0)
    assert rel(ru, Ru[1:-1, 1:-1]) < 1e-12 and rel(rv, Rv[1:-1, 1:-1]) < 1e-12 and rel(rp, Rp) < 1e-12


@pytest.mark.parametrize("N", [16, 30])
def test_parity_residual_matches_reference_fixture(golden, N):
    g = golden.residual
    ru, rv, rp = par.residual_interior_parity(ops.build_level(N, N), g[f"u_{N}"], g[f"v_{N}"], g[f"p_{N}"],
                                              float(g["Re"]), 5.0)
    assert rel(ru, g[f"Ru_mg_{N}"][1:-1, 1:-1]) < 1e-12
    assert rel(rv, g[f"Rv_mg_{N}"][1:-1, 1:-1]) < 1e-12
    assert rel(rp, g[f"Rp_mg_{N}"]) < 1e-12
