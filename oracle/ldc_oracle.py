"""CPU oracle for the spectral lid-driven-cavity pseudo-time stepper.

TEST INFRASTRUCTURE ONLY.  This module is a NumPy restatement of the
reference's Chebyshev-collocation residual + RK4 pseudo-time stepping
(`solver=spectral/sg` and `solver=spectral/fsg`).  Only `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py` may import it; the product package never does.

Parity status: PINNED against the reference itself.  The reference ships no
tests or golden vectors, so `tests/golden/make_golden.py`
imports the reference from /root/reference (in the authoring container) and
writes operator matrices, one residual evaluation, K-step trajectories and
full-solve summaries into `tests/golden/*.npz`; `tests/test_oracle.py` checks
this file against those fixtures.

Each function cites the reference file:line it follows (paths relative to
/root/reference/src/solvers).  Layout: fields are (n, n) C-order arrays with
x the slow axis (meshgrid indexing="ij"), n = N + 1; pressure lives on the
(m, m) inner grid, m = N - 1.  d/dx = Dx @ U, d/dy = U @ Dy.T.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

RK_ALPHA = (0.25, 1.0 / 3.0, 0.5, 1.0)  # spectral/sg.py:430, multigrid/fsg.py:973


# --------------------------------------------------------------------------
# 1-D building blocks
# --------------------------------------------------------------------------
def cgl_nodes(num_points: int, a: float = -1.0, b: float = 1.0) -> np.ndarray:
    """Chebyshev-Gauss-Lobatto nodes, ascending.  spectral/basis/spectral.py:18-39, 484-502."""
    N = num_points - 1
    xi = -np.cos(np.pi * np.arange(num_points) / N)
    if (a, b) == (-1.0, 1.0):
        return xi
    return 0.5 * (b - a) * (xi + 1.0) + a


def cheb_diff_reference_interval(xi: np.ndarray) -> np.ndarray:
    """D on [-1,1] with negative-row-sum diagonal.  spectral/basis/spectral.py:42-90."""
    n = xi.size
    N = n - 1
    if N == 0:
        return np.zeros((1, 1))
    c = np.ones(n)
    c[0] = c[N] = 2.0
    D = np.zeros((n, n))
    for i in range(n):
        for j in range(n):
            if i != j:
                D[i, j] = (c[i] / c[j]) * (-1.0) ** (i + j) / (xi[i] - xi[j])
    for i in range(n):
        D[i, i] = -np.sum(D[i, :])
    return D


def cheb_diff_matrix(num_points: int, a: float, b: float) -> np.ndarray:
    """Scaled to [a,b].  spectral/basis/spectral.py:504-522."""
    return (2.0 / (b - a)) * cheb_diff_reference_interval(cgl_nodes(num_points))


def interp_inner_to_full(nodes_inner: np.ndarray, nodes_full: np.ndarray) -> np.ndarray:
    """(n x m) polynomial interpolation matrix.  spectral/sg.py:212-248, multigrid/fsg.py:38-73."""
    from numpy.polynomial.chebyshev import chebvander

    m = nodes_inner.size
    a, b = nodes_full[0], nodes_full[-1]
    xi_in = 2 * (nodes_inner - a) / (b - a) - 1
    xi_fu = 2 * (nodes_full - a) / (b - a) - 1
    V_in = chebvander(xi_in, m - 1)
    V_fu = chebvander(xi_fu, m - 1)
    return V_fu @ np.linalg.solve(V_in, np.eye(m))


def clenshaw_curtis(num_points: int) -> np.ndarray:
    """Weights on [-1,1].  spectral/basis/spectral.py:411-470."""
    N = num_points - 1
    if N == 0:
        return np.array([2.0])
    if N == 1:
        return np.array([1.0, 1.0])
    w = np.zeros(num_points)
    for j in range(num_points):
        s = 0.0
        for k in range(N // 2 + 1):
            bk = 1.0 if k == 0 else 2.0 / (1 - 4 * k * k)
            if k == N // 2 and N % 2 == 0:
                bk *= 0.5
            s += bk * np.cos(2 * np.pi * k * j / N)
        w[j] = 2.0 * s / N
    w[0] *= 0.5
    w[N] *= 0.5
    return w


def lid_profile(x: np.ndarray, method: str, width: float, U: float, Lx: float):
    """(u_lid, v_lid) on the top boundary.  spectral/operators/corner.py:80-112, 148-169, 192-223."""
    x = np.asarray(x, dtype=float).ravel()
    m = method.lower()
    v = np.zeros_like(x)
    if m == "smoothing":
        u = np.full_like(x, U, dtype=float)
        if width > 0:
            s = width * Lx
            left = x < s
            if np.any(left):
                u[left] = 0.5 * (1 - np.cos(np.pi * x[left] / s)) * U
            right = x > (Lx - s)
            if np.any(right):
                u[right] = 0.5 * (1 - np.cos(np.pi * (Lx - x[right]) / s)) * U
        return u, v
    if m in ("saad", "polynomial"):
        xi = x / Lx
        return 16.0 * xi**2 * (1.0 - xi) ** 2 * U, v
    raise ValueError(f"Unknown corner treatment method: {method}. Use 'smoothing', 'polynomial', or 'saad'.")


# --------------------------------------------------------------------------
# Transfer operators (1-D maps, applied as P_x @ U @ P_y.T)
# --------------------------------------------------------------------------
def prolong_1d_fft(u_c: np.ndarray, n_f: int) -> np.ndarray:
    """DCT-I coefficients + cosine-sum evaluation, endpoint weighting kept exactly as the
    reference has it.  operators/transfer_operators.py:209-255, 31-50."""
    from scipy.fft import dct

    n_c = u_c.size
    if n_c == n_f:
        return u_c.copy()
    if n_c > n_f:
        raise ValueError(f"Prolongation requires n_coarse ({n_c}) <= n_fine ({n_f})")
    N_c = n_c - 1
    w = u_c.astype(float).copy()
    w[0] /= 2
    w[-1] /= 2
    coef = dct(w, type=1) / N_c
    coef[0] /= 2
    coef[-1] /= 2
    N_f = n_f - 1
    theta = np.pi * np.arange(n_f) / N_f
    k = np.arange(n_c)
    return np.cos(np.outer(theta, k)) @ coef


def prolong_1d_poly(u_c: np.ndarray, n_f: int) -> np.ndarray:
    """chebfit / chebval interpolant.  operators/transfer_operators.py:333-376."""
    from numpy.polynomial.chebyshev import chebfit, chebval

    n_c = u_c.size
    if n_c == n_f:
        return u_c.copy()
    xc = np.cos(np.pi * np.arange(n_c) / (n_c - 1))
    xf = np.cos(np.pi * np.arange(n_f) / (n_f - 1))
    return chebval(xf, chebfit(xc, u_c, deg=n_c - 1))


def restrict_1d_fft(u_f: np.ndarray, n_c: int) -> np.ndarray:
    """DCT truncation (uncalled by the FSG driver).  operators/transfer_operators.py:269-325."""
    from scipy.fft import dct

    n_f = u_f.size
    if n_f == n_c:
        return u_f.copy()
    if n_f < n_c:
        raise ValueError(f"Restriction requires n_fine ({n_f}) >= n_coarse ({n_c})")
    N_f = n_f - 1
    w = u_f.astype(float).copy()
    w[0] /= 2
    w[-1] /= 2
    coef = dct(w, type=1) / N_f
    coef[0] /= 2
    coef[-1] /= 2
    keep = min(n_c, coef.size)
    theta = np.pi * np.arange(n_c) / (n_c - 1)
    return np.cos(np.outer(theta, np.arange(keep))) @ coef[:keep]


def restrict_1d_injection(u_f: np.ndarray, n_c: int) -> np.ndarray:
    """Every other node.  operators/transfer_operators.py:388-428."""
    n_f = u_f.size
    if n_f == n_c:
        return u_f.copy()
    if n_f == 2 * n_c - 1 or (n_f % 2 == 1 and n_c == (n_f + 1) // 2):
        return u_f[::2].copy()
    xf = np.cos(np.pi * np.arange(n_f) / (n_f - 1))
    xc = np.cos(np.pi * np.arange(n_c) / (n_c - 1))
    return np.array([u_f[np.argmin(np.abs(xf - x))] for x in xc])


def transfer_matrix(fn, n_from: int, n_to: int) -> np.ndarray:
    """Dense (n_to x n_from) matrix of a linear 1-D transfer map."""
    M = np.zeros((n_to, n_from))
    eye = np.eye(n_from)
    for j in range(n_from):
        M[:, j] = fn(eye[j], n_to)
    return M


def transfer_2d(fn, U: np.ndarray, shape_to) -> np.ndarray:
    """Columns first, then rows.  operators/transfer_operators.py:93-130."""
    nx_c, ny_c = U.shape
    nx_f, ny_f = shape_to
    if (nx_c, ny_c) == (nx_f, ny_f):
        return U.copy()
    tmp = np.zeros((nx_f, ny_c))
    for j in range(ny_c):
        tmp[:, j] = fn(U[:, j], nx_f)
    out = np.zeros((nx_f, ny_f))
    for i in range(nx_f):
        out[i, :] = fn(tmp[i, :], ny_f)
    return out


PROLONG_1D = {"fft": prolong_1d_fft, "polynomial": prolong_1d_poly}
RESTRICT_1D = {"fft": restrict_1d_fft, "injection": restrict_1d_injection}


# --------------------------------------------------------------------------
# One grid level
# --------------------------------------------------------------------------
@dataclass
class Level:
    N: int
    Lx: float
    Ly: float
    x: np.ndarray
    y: np.ndarray
    Dx: np.ndarray
    Dy: np.ndarray
    Dxx: np.ndarray
    Dyy: np.ndarray
    Ix: np.ndarray
    Iy: np.ndarray
    dx_min: float
    dy_min: float
    W: np.ndarray
    extra: dict = field(default_factory=dict)

    @property
    def n(self):
        return self.N + 1

    @property
    def m(self):
        return self.N - 1


def build_level(N: int, Lx: float = 1.0, Ly: float = 1.0) -> Level:
    """Grids + operators for polynomial order N (Chebyshev basis).
    spectral/sg.py:103-119, 181-210, 479-493; multigrid/fsg.py:357-481."""
    n = N + 1
    x = cgl_nodes(n, 0.0, Lx)
    y = cgl_nodes(n, 0.0, Ly)
    Dx = cheb_diff_matrix(n, 0.0, Lx)
    Dy = cheb_diff_matrix(n, 0.0, Ly)
    Dxx = Dx @ Dx  # matrix square, not an analytic second-derivative matrix (Q6)
    Dyy = Dy @ Dy
    Ix = interp_inner_to_full(x[1:-1], x)
    Iy = interp_inner_to_full(y[1:-1], y)
    wx = clenshaw_curtis(n) * (Lx - 0.0) / 2
    wy = clenshaw_curtis(n) * (Ly - 0.0) / 2
    return Level(N=N, Lx=Lx, Ly=Ly, x=x, y=y, Dx=Dx, Dy=Dy, Dxx=Dxx, Dyy=Dyy, Ix=Ix, Iy=Iy,
                 dx_min=float(np.min(np.diff(x))), dy_min=float(np.min(np.diff(y))),
                 W=np.outer(wx, wy))


def hierarchy_orders(n_fine: int, n_levels: int, coarsest_n: int = 12):
    """Coarse-to-fine polynomial orders.  multigrid/fsg.py:519-531."""
    orders = []
    n = n_fine
    for _ in range(n_levels):
        orders.append(n)
        nxt = n // 2
        if nxt < coarsest_n:
            break
        n = nxt
    return orders[::-1]


# --------------------------------------------------------------------------
# Residual, BCs, time step, RK4 step
# --------------------------------------------------------------------------
def pressure_gradient(lv: Level, p: np.ndarray):
    """p_full = Ix p Iy^T; (Dx p_full, p_full Dy^T).  spectral/sg.py:255-276, multigrid/fsg.py:181-216."""
    pf = lv.Ix @ p @ lv.Iy.T
    return lv.Dx @ pf, pf @ lv.Dy.T


def residual(lv: Level, u, v, p, Re: float, beta2: float):
    """R_u, R_v on all n x n nodes; R_p on the m x m inner grid.
    spectral/sg.py:278-346, multigrid/fsg.py:857-910 (+ njit kernels 81-178)."""
    du_dx = lv.Dx @ u
    du_dy = u @ lv.Dy.T
    dv_dx = lv.Dx @ v
    dv_dy = v @ lv.Dy.T
    lap_u = lv.Dxx @ u + u @ lv.Dyy.T
    lap_v = lv.Dxx @ v + v @ lv.Dyy.T
    dp_dx, dp_dy = pressure_gradient(lv, p)
    nu = 1.0 / Re
    Ru = -(u * du_dx + v * du_dy) - dp_dx + nu * lap_u
    Rv = -(u * dv_dx + v * dv_dy) - dp_dy + nu * lap_v
    Rp = -beta2 * (du_dx + dv_dy)[1:-1, 1:-1]
    return Ru, Rv, Rp


def enforce_bc(u, v, u_lid, v_lid):
    """W, E, S walls = 0, then the lid column last (corners take the lid value).
    spectral/sg.py:348-385, multigrid/fsg.py:912-932."""
    u[0, :] = 0.0
    v[0, :] = 0.0
    u[-1, :] = 0.0
    v[-1, :] = 0.0
    u[:, 0] = 0.0
    v[:, 0] = 0.0
    u[:, -1] = u_lid
    v[:, -1] = v_lid


def adaptive_dt(lv: Level, u, v, Re, beta2, CFL, lid_velocity):
    """spectral/sg.py:387-408, multigrid/fsg.py:934-948."""
    u_max = max(np.max(np.abs(u)), lid_velocity)
    v_max = max(np.max(np.abs(v)), 1e-10)
    nu = 1.0 / Re
    lam_x = (u_max + np.sqrt(u_max**2 + beta2)) / lv.dx_min + nu / lv.dx_min**2
    lam_y = (v_max + np.sqrt(v_max**2 + beta2)) / lv.dy_min + nu / lv.dy_min**2
    return CFL / (lam_x + lam_y)


@dataclass
class Case:
    """Per-cavity physical parameters."""
    Re: float = 100.0
    beta2: float = 5.0
    CFL: float = 1.5
    lid_velocity: float = 1.0
    corner_treatment: str = "smoothing"
    corner_smoothing: float = 0.15


def rk4_step(lv: Level, u, v, p, case: Case, u_lid, v_lid, freeze_pressure: bool):
    """One low-storage 4-stage RK pseudo-step, in place on (u, v, p).

    freeze_pressure=True  -> `SGSolver.step` semantics: every stage takes grad(p^n) (SURVEY Q1;
                             spectral/sg.py:270 reads self.p_2d, not the staged p).
    freeze_pressure=False -> `MultigridSmoother.step` semantics (multigrid/fsg.py:956-995).
    Returns (u_res, v_res, last-stage R_u, R_v, R_p)."""
    u0, v0, p0 = u.copy(), v.copy(), p.copy()
    dt = adaptive_dt(lv, u0, v0, case.Re, case.beta2, case.CFL, case.lid_velocity)
    ui, vi, pi = u0, v0, p0
    Ru = Rv = Rp = None
    for s, alpha in enumerate(RK_ALPHA):
        Ru, Rv, Rp = residual(lv, ui, vi, p0 if freeze_pressure else pi, case.Re, case.beta2)
        ui = u0 + alpha * dt * Ru
        vi = v0 + alpha * dt * Rv
        pi = p0 + alpha * dt * Rp
        enforce_bc(ui, vi, u_lid, v_lid)
    u[:] = ui
    v[:] = vi
    p[:] = pi
    u_res = np.linalg.norm(u - u0) / (np.linalg.norm(u0) + 1e-12)
    v_res = np.linalg.norm(v - v0) / (np.linalg.norm(v0) + 1e-12)
    return u_res, v_res, Ru, Rv, Rp


# --------------------------------------------------------------------------
# Diagnostics and output pressure
# --------------------------------------------------------------------------
def diagnostics(lv: Level, u, v):
    """(energy, enstrophy, palinstrophy) with Clenshaw-Curtis quadrature.  spectral/sg.py:495-550."""
    E = 0.5 * float(np.sum(lv.W * (u * u + v * v)))
    om = lv.Dx @ v - u @ lv.Dy.T
    Z = 0.5 * float(np.sum(lv.W * om * om))
    gx = lv.Dx @ om
    gy = om @ lv.Dy.T
    P = 0.5 * float(np.sum(lv.W * (gx**2 + gy**2)))
    return E, Z, P


def evaluate_at_points(lv: Level, u, v, xq, yq):
    """Bilinear evaluation of u, v at scattered points, as the reference does it (base.py:1122-1160): the fields are
    re-laid out as (y, x) and handed to SciPy's RegularGridInterpolator(method="linear", fill_value=nan)."""
    from scipy.interpolate import RegularGridInterpolator

    pts = np.column_stack([yq, xq])
    fu = RegularGridInterpolator((lv.y, lv.x), np.ascontiguousarray(u.T), method="linear", bounds_error=False, fill_value=np.nan)
    fv = RegularGridInterpolator((lv.y, lv.x), np.ascontiguousarray(v.T), method="linear", bounds_error=False, fill_value=np.nan)
    return fu(pts), fv(pts)


def validation_errors(lv: Level, u, v, xq, yq, ref_u, ref_v, Lx=1.0, Ly=1.0):
    """(u_L2_error, v_L2_error, n_valid) against a stored solution.  base.py:1015-1040."""
    cu, cv = evaluate_at_points(lv, u, v, xq, yq)
    margin = 1e-10
    interior = (xq > margin) & (xq < Lx - margin) & (yq > margin) & (yq < Ly - margin)
    valid = interior & ~(np.isnan(cu) | np.isnan(cv))
    eu = np.linalg.norm(cu[valid] - ref_u[valid]) / (np.linalg.norm(ref_u[valid]) + 1e-12)
    ev = np.linalg.norm(cv[valid] - ref_v[valid]) / (np.linalg.norm(ref_v[valid]) + 1e-12)
    return float(eu), float(ev), int(np.sum(valid))


def extrapolate_pressure(p: np.ndarray) -> np.ndarray:
    """Inner (m x m) -> full (n x n) by linear extrapolation; corners = mean of two neighbours.
    spectral/sg.py:144-179."""
    m = p.shape[0]
    f = np.zeros((m + 2, p.shape[1] + 2))
    f[1:-1, 1:-1] = p
    f[0, 1:-1] = 2 * f[1, 1:-1] - f[2, 1:-1]
    f[-1, 1:-1] = 2 * f[-2, 1:-1] - f[-3, 1:-1]
    f[1:-1, 0] = 2 * f[1:-1, 1] - f[1:-1, 2]
    f[1:-1, -1] = 2 * f[1:-1, -2] - f[1:-1, -3]
    f[0, 0] = 0.5 * (f[0, 1] + f[1, 0])
    f[0, -1] = 0.5 * (f[0, -2] + f[1, -1])
    f[-1, 0] = 0.5 * (f[-1, 1] + f[-2, 0])
    f[-1, -1] = 0.5 * (f[-1, -2] + f[-2, -1])
    return f


# --------------------------------------------------------------------------
# Drivers
# --------------------------------------------------------------------------
def initial_state(lv: Level, u_lid, v_lid):
    """Zeros + lid column.  spectral/sg.py:98, 250-253."""
    n, m = lv.n, lv.m
    u = np.zeros((n, n))
    v = np.zeros((n, n))
    p = np.zeros((m, m))
    u[:, -1] = u_lid
    v[:, -1] = v_lid
    return u, v, p


def solve_sg(N: int, case: Case, tol: float, max_iter: int, Lx=1.0, Ly=1.0, history=False):
    """`LidDrivenCavitySolver.solve` driving `SGSolver.step`: converged iff i >= 10 and
    rel < tol, iterations = i + 1.  base.py:243-313."""
    lv = build_level(N, Lx, Ly)
    u_lid, v_lid = lid_profile(lv.x, case.corner_treatment, case.corner_smoothing, case.lid_velocity, Lx)
    u, v, p = initial_state(lv, u_lid, v_lid)
    it, conv, rel = 0, False, float("inf")
    hist = []
    Ru = Rv = Rp = None
    for i in range(max_iter):
        it = i + 1
        ur, vr, Ru, Rv, Rp = rk4_step(lv, u, v, p, case, u_lid, v_lid, freeze_pressure=True)
        rel = max(ur, vr)
        if history:
            # base.py:258-276: residual norms of the LAST stage's residual, then E, Z, P of the new state
            hist.append((rel, np.linalg.norm(Ru), np.linalg.norm(Rv), np.linalg.norm(Rp)) + tuple(diagnostics(lv, u, v)))
        conv = (i >= 10) and (rel < tol)
        if conv:
            break
    out = dict(u=u, v=v, p=p, iterations=it, converged=conv, rel=rel, level=lv,
               res_norms=(np.linalg.norm(Ru), np.linalg.norm(Rv), np.linalg.norm(Rp)))
    if history:
        out["history"] = np.array(hist)
    return out


def prolongate_state(lc: Level, lf: Level, uc, vc, pc, method: str, lid_velocity: float):
    """Coarse -> fine initial guess including the reference's swapped-axes "BC re-enforcement"
    (SURVEY Q3): row 0 := 0, row -1 := lid_velocity (the EAST wall), col 0 := 0, col -1 := 0.
    multigrid/fsg.py:551-614."""
    fn = PROLONG_1D[method]
    uf = transfer_2d(fn, uc, (lf.n, lf.n))
    vf = transfer_2d(fn, vc, (lf.n, lf.n))
    uf[0, :] = 0.0
    vf[0, :] = 0.0
    uf[-1, :] = lid_velocity
    vf[-1, :] = 0.0
    uf[:, 0] = 0.0
    vf[:, 0] = 0.0
    uf[:, -1] = 0.0
    vf[:, -1] = 0.0
    pf = transfer_2d(fn, pc, (lf.m, lf.m))
    return uf, vf, pf


def solve_fsg(N: int, case: Case, tol: float, max_iter: int, n_levels=2, coarse_tolerance_factor=1.0,
              prolongation="fft", Lx=1.0, Ly=1.0):
    """Coarse-to-fine FSG driver.  multigrid/fsg.py:1053-1221, spectral/fsg.py:39-129.
    Convergence per level: max(u_res, v_res) < tol * ctf**levels_from_finest, no warm-up;
    non-finite -> abort every level (finest stays zero, Q8); max_iter applies per level."""
    orders = hierarchy_orders(N, n_levels)
    levels = [build_level(o, Lx, Ly) for o in orders]
    total = 0
    conv = False
    diverged = False
    per_level = []
    state = None
    for li, lv in enumerate(levels):
        u_lid, v_lid = lid_profile(lv.x, case.corner_treatment, case.corner_smoothing, case.lid_velocity, Lx)
        ltol = tol * (coarse_tolerance_factor ** (len(levels) - 1 - li))
        if li == 0:
            u = np.zeros((lv.n, lv.n))
            v = np.zeros((lv.n, lv.n))
            p = np.zeros((lv.m, lv.m))
        else:
            u, v, p = prolongate_state(levels[li - 1], lv, *state, prolongation, case.lid_velocity)
        u[:, -1] = u_lid  # smoother.initialize_lid, multigrid/fsg.py:950-954, 1161
        v[:, -1] = v_lid
        conv = False
        it_l = 0
        for _ in range(max_iter):
            ur, vr, *_ = rk4_step(lv, u, v, p, case, u_lid, v_lid, freeze_pressure=False)
            it_l += 1
            total += 1
            mr = max(ur, vr)
            if mr < ltol:
                conv = True
                break
            if not np.isfinite(mr):
                diverged = True
                break
        per_level.append(it_l)
        state = (u, v, p)
        if diverged:
            break
    fin = levels[-1]
    if diverged and len(per_level) < len(levels):
        u = np.zeros((fin.n, fin.n))
        v = np.zeros((fin.n, fin.n))
        p = np.zeros((fin.m, fin.m))
    else:
        u, v, p = state
    return dict(u=u, v=v, p=p, iterations=total, converged=bool(conv and not diverged), diverged=diverged,
                per_level=per_level, level=fin, orders=orders)
