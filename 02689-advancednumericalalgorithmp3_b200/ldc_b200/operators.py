"""Host-side setup of the spectral operators the CUDA kernels consume (NumPy, O(n^2)-O(n^3) once per N).

Everything here is per polynomial order, not per step: collocation nodes, first/second differentiation
matrices, the inner->full pressure interpolation, quadrature weights, lid profiles and the dense 1-D
transfer maps between grid levels.  Reference sites are cited per function (paths relative to
src/solvers/ of the reference).
"""
from __future__ import annotations

from dataclasses import dataclass
from functools import lru_cache

import numpy as np
from numpy.polynomial import chebyshev as _cheb
from numpy.polynomial import legendre as _leg

BASES = ("chebyshev", "legendre")


# ---------------------------------------------------------------------------------------------------
# nodes, differentiation, quadrature
# ---------------------------------------------------------------------------------------------------
def reference_nodes(n: int, basis: str) -> np.ndarray:
    """Gauss-Lobatto nodes on [-1, 1], ascending (spectral/basis/spectral.py:18-39; polynomial.py:164-195)."""
    if basis == "chebyshev":
        return -np.cos(np.pi * np.arange(n) / (n - 1))
    if basis == "legendre":
        inner = _leg.Legendre.basis(n - 1).deriv().roots()
        return np.sort(np.concatenate(([-1.0], inner, [1.0])))
    raise ValueError(f"Unknown basis_type: {basis}. Use 'legendre' or 'chebyshev'")


def map_nodes(xi: np.ndarray, a: float, b: float) -> np.ndarray:
    """Affine map [-1,1] -> [a,b] (spectral/basis/spectral.py:498-502)."""
    if (a, b) == (-1.0, 1.0):
        return xi
    return 0.5 * (b - a) * (xi + 1.0) + a


def diff_matrix_reference_interval(xi: np.ndarray, basis: str) -> np.ndarray:
    """First-derivative collocation matrix on [-1,1].

    Chebyshev: D_ij = (c_i/c_j) (-1)^(i+j) / (xi_i - xi_j), diagonal by the negative-row-sum trick
    (spectral/basis/spectral.py:42-90).  Legendre: D = V_x V^-1 with the Legendre Vandermonde pair
    (spectral/basis/spectral.py:93-130)."""
    n = xi.size
    if basis == "chebyshev":
        idx = np.arange(n)
        c = np.ones(n)
        c[0] = c[-1] = 2.0
        sign = np.where((idx[:, None] + idx[None, :]) % 2 == 0, 1.0, -1.0)
        with np.errstate(divide="ignore", invalid="ignore"):
            D = (c[:, None] / c[None, :]) * sign / (xi[:, None] - xi[None, :])
        D[idx, idx] = 0.0
        for i in range(n):  # row sums in the reference's order (np.sum over a row with a zero diagonal)
            D[i, i] = -np.sum(D[i, :])
        return D
    if basis == "legendre":
        V = _leg.legvander(xi, n - 1)
        Vx = np.zeros_like(V)
        for k in range(n):
            coef = np.zeros(n)
            coef[k] = 1.0
            Vx[:, k] = _leg.legval(xi, _leg.legder(coef))
        return Vx @ np.linalg.solve(V, np.eye(n))
    raise ValueError(f"Unknown basis_type: {basis}. Use 'legendre' or 'chebyshev'")


def interpolation_inner_to_full(nodes_full: np.ndarray) -> np.ndarray:
    """(n x m) matrix taking values on the m = n-2 inner nodes to all n nodes through the degree m-1
    interpolating polynomial (spectral/sg.py:212-248; multigrid/fsg.py:38-73).  Its interior rows are the
    identity up to rounding; rows 0 and n-1 extrapolate."""
    inner = nodes_full[1:-1]
    a, b = nodes_full[0], nodes_full[-1]
    m = inner.size
    t_in = 2 * (inner - a) / (b - a) - 1
    t_fu = 2 * (nodes_full - a) / (b - a) - 1
    return _cheb.chebvander(t_fu, m - 1) @ np.linalg.solve(_cheb.chebvander(t_in, m - 1), np.eye(m))


def quadrature_weights_reference_interval(n: int, basis: str) -> np.ndarray:
    """Clenshaw-Curtis (spectral/basis/spectral.py:411-470) or LGL (polynomial.py:198-243) weights on [-1,1]."""
    N = n - 1
    if basis == "chebyshev":
        if N == 0:
            return np.array([2.0])
        if N == 1:
            return np.array([1.0, 1.0])
        w = np.zeros(n)
        ks = np.arange(N // 2 + 1)
        bk = np.where(ks == 0, 1.0, 2.0 / (1.0 - 4.0 * ks * ks))
        if N % 2 == 0:
            bk[-1] *= 0.5
        for j in range(n):
            acc = 0.0
            for k in range(N // 2 + 1):  # sequential accumulation, as the reference sums it
                acc += bk[k] * np.cos(2 * np.pi * k * j / N)
            w[j] = 2.0 * acc / N
        w[0] *= 0.5
        w[N] *= 0.5
        return w
    if basis == "legendre":
        if N == 0:
            return np.array([2.0])
        xi = reference_nodes(n, basis)
        coef = np.zeros(n)
        coef[N] = 1.0
        PN = _leg.legval(xi, coef)
        return 2.0 / (N * (N + 1) * PN**2)
    raise ValueError(f"Unknown basis_type: {basis}. Use 'legendre' or 'chebyshev'")


# ---------------------------------------------------------------------------------------------------
# lid profiles (the swept hyper-parameter)
# ---------------------------------------------------------------------------------------------------
CORNER_METHODS = ("smoothing", "saad", "polynomial")


def lid_velocity_profile(x: np.ndarray, method: str, smoothing_width: float, lid_velocity: float, Lx: float):
    """(u_lid, v_lid) along the north boundary (spectral/operators/corner.py:80-112, 148-169, 192-223).

    smoothing: cosine ramps of width smoothing_width*Lx at both ends (strict inequalities, width <= 0 means
    no smoothing); saad / polynomial: 16 xi^2 (1 - xi)^2.  v_lid is zero."""
    x = np.asarray(x, dtype=np.float64).ravel()
    key = method.lower()
    v = np.zeros_like(x)
    if key == "smoothing":
        u = np.full_like(x, lid_velocity)
        if smoothing_width > 0:
            s = smoothing_width * Lx
            lo = x < s
            u[lo] = 0.5 * (1 - np.cos(np.pi * x[lo] / s)) * lid_velocity
            hi = x > (Lx - s)
            u[hi] = 0.5 * (1 - np.cos(np.pi * (Lx - x[hi]) / s)) * lid_velocity
        return u, v
    if key in ("saad", "polynomial"):
        t = x / Lx
        return 16.0 * t**2 * (1.0 - t) ** 2 * lid_velocity, v
    raise ValueError(f"Unknown corner treatment method: {method}. Use 'smoothing', 'polynomial', or 'saad'.")


# ---------------------------------------------------------------------------------------------------
# grid level
# ---------------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class LevelOperators:
    """Everything the kernels need for one polynomial order (nx, ny) on [0,Lx] x [0,Ly]."""
    nx: int
    ny: int
    Lx: float
    Ly: float
    basis: str
    x: np.ndarray     # (nx+1,)
    y: np.ndarray     # (ny+1,)
    Dx: np.ndarray    # (nx+1, nx+1)
    Dy: np.ndarray
    Dxx: np.ndarray   # Dx @ Dx, the reference's second-derivative operator (spectral/sg.py:192-193)
    Dyy: np.ndarray
    Ix: np.ndarray    # (nx+1, nx-1)
    Iy: np.ndarray
    dx_min: float
    dy_min: float
    W: np.ndarray     # (nx+1, ny+1) tensor-product quadrature weights (spectral/sg.py:479-493)
    wx: np.ndarray = None   # 1-D quadrature weights
    wy: np.ndarray = None

    @property
    def shape_full(self):
        return (self.nx + 1, self.ny + 1)

    @property
    def shape_inner(self):
        return (self.nx - 1, self.ny - 1)


def _one_direction(n_order: int, L: float, basis: str):
    n = n_order + 1
    xi = reference_nodes(n, basis)
    nodes = map_nodes(xi, 0.0, L)
    D = (2.0 / (L - 0.0)) * diff_matrix_reference_interval(xi, basis)
    DD = D @ D
    I = interpolation_inner_to_full(nodes)
    w = quadrature_weights_reference_interval(n, basis) * (L - 0.0) / 2
    return nodes, D, DD, I, w


@lru_cache(maxsize=64)
def build_level(nx: int, ny: int, Lx: float = 1.0, Ly: float = 1.0, basis: str = "chebyshev") -> LevelOperators:
    """Grids + operators (spectral/sg.py:103-119, 181-210; multigrid/fsg.py:357-481).  The reference's dense
    n^2 x n^2 Kronecker operators are deliberately not built: nothing on the stepping path reads them."""
    basis = basis.lower()
    if basis not in BASES:
        raise ValueError(f"Unknown basis_type: {basis}. Use 'legendre' or 'chebyshev'")
    x, Dx, Dxx, Ix, wx = _one_direction(nx, Lx, basis)
    if ny == nx and Ly == Lx:
        y, Dy, Dyy, Iy, wy = x, Dx, Dxx, Ix, wx
    else:
        y, Dy, Dyy, Iy, wy = _one_direction(ny, Ly, basis)
    return LevelOperators(nx=nx, ny=ny, Lx=Lx, Ly=Ly, basis=basis, x=x, y=y, Dx=Dx, Dy=Dy, Dxx=Dxx, Dyy=Dyy, Ix=Ix,
                          Iy=Iy, dx_min=float(np.min(np.diff(x))), dy_min=float(np.min(np.diff(y))),
                          W=np.outer(wx, wy), wx=wx, wy=wy)


def hierarchy_orders(n_fine: int, n_levels: int, coarsest_n: int = 12):
    """Polynomial orders coarse -> fine: N, N//2, ... up to n_levels, stopping before an order < coarsest_n
    (multigrid/fsg.py:489-543)."""
    orders = []
    n = n_fine
    for _ in range(n_levels):
        orders.append(n)
        if n // 2 < coarsest_n:
            break
        n //= 2
    return orders[::-1]


# ---------------------------------------------------------------------------------------------------
# level-transfer maps as dense matrices
# ---------------------------------------------------------------------------------------------------
def _dct1_matrix(n: int) -> np.ndarray:
    """SciPy's un-normalised DCT-I as a matrix: y_k = x_0 + (-1)^k x_{n-1} + 2 sum_j x_j cos(pi k j / (n-1))."""
    N = n - 1
    k = np.arange(n)
    M = 2.0 * np.cos(np.pi * np.outer(k, k) / N)
    M[:, 0] = 1.0
    M[:, -1] = np.where(k % 2 == 0, 1.0, -1.0)
    return M


def _cheb_coefficient_map(n: int) -> np.ndarray:
    """values -> "Chebyshev coefficients" exactly as FFTProlongation computes them: halve both endpoint
    samples, DCT-I / N, halve the first and last coefficient (operators/transfer_operators.py:243-252).
    SciPy's DCT-I already weights the endpoints by one half relative to the interior, so this map is NOT
    the interpolating one; it is reproduced on purpose."""
    N = n - 1
    A = _dct1_matrix(n)
    A[:, 0] *= 0.5
    A[:, -1] *= 0.5
    A /= N
    A[0, :] *= 0.5
    A[-1, :] *= 0.5
    return A


def prolongation_matrix(n_coarse: int, n_fine: int, method: str = "fft") -> np.ndarray:
    """(n_fine x n_coarse) 1-D coarse->fine map.  'fft' = FFTProlongation (operators/transfer_operators.py:209-255
    with the cosine-sum evaluation of 31-50); 'polynomial' = chebfit/chebval interpolation (333-376)."""
    if n_coarse > n_fine:
        raise ValueError(f"Prolongation requires n_coarse ({n_coarse}) <= n_fine ({n_fine})")
    if n_coarse == n_fine:
        return np.eye(n_fine)
    key = method.lower()
    if key == "fft":
        theta = np.pi * np.arange(n_fine) / (n_fine - 1)
        return np.cos(np.outer(theta, np.arange(n_coarse))) @ _cheb_coefficient_map(n_coarse)
    if key == "polynomial":
        xc = np.cos(np.pi * np.arange(n_coarse) / (n_coarse - 1))
        xf = np.cos(np.pi * np.arange(n_fine) / (n_fine - 1))
        # chebfit at full degree solves the Vandermonde system by least squares, as the reference does
        cols = [_cheb.chebval(xf, _cheb.chebfit(xc, e, deg=n_coarse - 1)) for e in np.eye(n_coarse)]
        return np.column_stack(cols)
    raise ValueError(f"Unknown prolongation method: {method}. Use 'fft' or 'polynomial'.")


def restriction_matrix(n_fine: int, n_coarse: int, method: str = "fft") -> np.ndarray:
    """(n_coarse x n_fine) 1-D fine->coarse map.  'fft' = DCT truncation (operators/transfer_operators.py:269-325),
    'injection' = every other node / nearest node (388-428).  The reference's FSG driver never restricts."""
    if n_fine < n_coarse:
        raise ValueError(f"Restriction requires n_fine ({n_fine}) >= n_coarse ({n_coarse})")
    if n_fine == n_coarse:
        return np.eye(n_coarse)
    key = method.lower()
    if key == "fft":
        theta = np.pi * np.arange(n_coarse) / (n_coarse - 1)
        keep = min(n_coarse, n_fine)
        return np.cos(np.outer(theta, np.arange(keep))) @ _cheb_coefficient_map(n_fine)[:keep, :]
    if key == "injection":
        R = np.zeros((n_coarse, n_fine))
        if n_fine == 2 * n_coarse - 1:
            R[np.arange(n_coarse), 2 * np.arange(n_coarse)] = 1.0
        else:
            xf = np.cos(np.pi * np.arange(n_fine) / (n_fine - 1))
            xc = np.cos(np.pi * np.arange(n_coarse) / (n_coarse - 1))
            for i, xv in enumerate(xc):
                R[i, int(np.argmin(np.abs(xf - xv)))] = 1.0
        return R
    raise ValueError(f"Unknown restriction method: {method}. Use 'fft' or 'injection'.")


def extrapolate_pressure_to_full(p_inner: np.ndarray) -> np.ndarray:
    """Output pressure: inner (m x m) values + linear extrapolation to the four edges, corners = mean of the
    two neighbours (spectral/sg.py:144-179).  Accepts (..., mx, my)."""
    p = np.asarray(p_inner)
    out = np.zeros(p.shape[:-2] + (p.shape[-2] + 2, p.shape[-1] + 2))
    out[..., 1:-1, 1:-1] = p
    out[..., 0, 1:-1] = 2 * out[..., 1, 1:-1] - out[..., 2, 1:-1]
    out[..., -1, 1:-1] = 2 * out[..., -2, 1:-1] - out[..., -3, 1:-1]
    out[..., 1:-1, 0] = 2 * out[..., 1:-1, 1] - out[..., 1:-1, 2]
    out[..., 1:-1, -1] = 2 * out[..., 1:-1, -2] - out[..., 1:-1, -3]
    out[..., 0, 0] = 0.5 * (out[..., 0, 1] + out[..., 1, 0])
    out[..., 0, -1] = 0.5 * (out[..., 0, -2] + out[..., 1, -1])
    out[..., -1, 0] = 0.5 * (out[..., -1, 1] + out[..., -2, 0])
    out[..., -1, -1] = 0.5 * (out[..., -1, -2] + out[..., -2, -1])
    return out
