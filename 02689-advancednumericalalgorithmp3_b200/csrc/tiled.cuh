// Tiled path: state lives in global memory (L2-resident for every realistic batch), one launch per
// RK stage.  Works for any (nx, ny); it is also the full-fidelity evaluation of the reference residual on
// ALL nodes (ldc_residual), the diagnostics and the level-transfer kernels.
#pragma once
#include "common.cuh"

namespace ldc {

constexpr int RED_THREADS = 256;

// Deterministic block-wide reductions (fixed tree, no atomics): result valid in every thread.
template <int NV, bool IS_MAX>
__device__ inline void block_reduce(double (&v)[NV], double* sh /* [NV * 32] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        double x = v[q];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double y = __shfl_xor_sync(0xffffffffu, x, o);
            x = IS_MAX ? fmax(x, y) : x + y;
        }
        if (lane == 0) sh[q * 32 + warp] = x;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        double x = sh[q * 32];
        for (int w = 1; w < nwarps; ++w) x = IS_MAX ? fmax(x, sh[q * 32 + w]) : x + sh[q * 32 + w];
        v[q] = x;
    }
    __syncthreads();
}

// Convergence bookkeeping after one RK4 step of one cavity.  Mirrors base.solve (base.py:250-286: warm-up
// of `warmup` iterations, iterations = i + 1) and solve_fsg (multigrid/fsg.py:1168-1202: convergence test
// first, then the non-finite exit).  max() follows Python's: the first argument wins unless the second is
// strictly greater.
__device__ inline void update_status(CavityStatus& s, double ures, double vres, double tol, int warmup,
                                     long long max_iter, int nan_exit) {
    s.iters += 1;
    s.u_res = ures;
    s.v_res = vres;
    const double r = (vres > ures) ? vres : ures;
    if ((s.iters - 1) >= warmup && r < tol) {
        s.converged = 1;
        s.active = 0;
    } else if (nan_exit && !isfinite(r)) {
        s.diverged = 1;
        s.active = 0;
    } else if (s.iters >= max_iter) {
        s.active = 0;
    }
}

// out[b] (ra x cx) = A (ra x ca, shared by the batch) * X[b] (ca x cx)
__global__ void k_left_mul(const double* __restrict__ A, const double* __restrict__ X, double* __restrict__ out,
                           int ra, int ca, int cx, const CavityStatus* __restrict__ mask) {
    const int b = blockIdx.z;
    if (mask && !mask[b].active) return;
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= ra || j >= cx) return;
    const double* Xb = X + (size_t)b * ca * cx;
    double acc = 0.0;
    for (int k = 0; k < ca; ++k) acc = fma(A[i * ca + k], Xb[k * cx + j], acc);
    out[(size_t)b * ra * cx + i * cx + j] = acc;
}

// out[b] (rx x ra) = X[b] (rx x ca) * A^T, A given TRANSPOSED as AT (ca x ra) so reads coalesce.
__global__ void k_right_mul(const double* __restrict__ X, const double* __restrict__ AT, double* __restrict__ out,
                            int rx, int ca, int ra, const CavityStatus* __restrict__ mask) {
    const int b = blockIdx.z;
    if (mask && !mask[b].active) return;
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= rx || j >= ra) return;
    const double* Xb = X + (size_t)b * rx * ca;
    double acc = 0.0;
    for (int k = 0; k < ca; ++k) acc = fma(Xb[i * ca + k], AT[k * ra + j], acc);
    out[(size_t)b * rx * ra + i * ra + j] = acc;
}

__global__ void k_transpose(const double* __restrict__ A, double* __restrict__ AT, int r, int c) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= r * c) return;
    const int i = idx / c, j = idx % c;
    AT[j * r + i] = A[idx];
}

// Residual on all nodes, the reference formula (spectral/sg.py:278-346; multigrid/fsg.py:81-178, 857-910):
//   R_u = -(u du/dx + v du/dy) - dp/dx + nu lap(u),  R_v likewise,  R_p = -beta^2 (du/dx + dv/dy) on the inner grid
// with d/dx = Dx @ U, d/dy = U @ Dy^T, lap = Dxx @ U + U @ Dyy^T, p_full = Ix p Iy^T.
__global__ void __launch_bounds__(256)
k_residual(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ pfull,
           const double* __restrict__ Dx, const double* __restrict__ Dxx, const double* __restrict__ DyT,
           const double* __restrict__ DyyT, const double* __restrict__ Re, const double* __restrict__ beta2,
           double* __restrict__ Ru, double* __restrict__ Rv, double* __restrict__ Rp, int n1x, int n1y,
           const CavityStatus* __restrict__ mask) {
    const int b = blockIdx.z;
    if (mask && !mask[b].active) return;
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= n1x || j >= n1y) return;
    const size_t off = (size_t)b * n1x * n1y;
    const double *ub = u + off, *vb = v + off, *pb = pfull + off;
    double dudx = 0, dvdx = 0, dxxu = 0, dxxv = 0, dpdx = 0;
    for (int k = 0; k < n1x; ++k) {
        const double d1 = Dx[i * n1x + k], d2 = Dxx[i * n1x + k];
        const double uu = ub[k * n1y + j], vv = vb[k * n1y + j], pp = pb[k * n1y + j];
        dudx = fma(d1, uu, dudx);
        dvdx = fma(d1, vv, dvdx);
        dpdx = fma(d1, pp, dpdx);
        dxxu = fma(d2, uu, dxxu);
        dxxv = fma(d2, vv, dxxv);
    }
    double dudy = 0, dvdy = 0, dyyu = 0, dyyv = 0, dpdy = 0;
    for (int k = 0; k < n1y; ++k) {
        const double d1 = DyT[k * n1y + j], d2 = DyyT[k * n1y + j];
        const double uu = ub[i * n1y + k], vv = vb[i * n1y + k], pp = pb[i * n1y + k];
        dudy = fma(uu, d1, dudy);
        dvdy = fma(vv, d1, dvdy);
        dpdy = fma(pp, d1, dpdy);
        dyyu = fma(uu, d2, dyyu);
        dyyv = fma(vv, d2, dyyv);
    }
    const double nu = 1.0 / Re[b];
    const double uc = ub[i * n1y + j], vc = vb[i * n1y + j];
    const double conv_u = uc * dudx + vc * dudy;
    const double conv_v = uc * dvdx + vc * dvdy;
    Ru[off + i * n1y + j] = -conv_u - dpdx + nu * (dxxu + dyyu);
    Rv[off + i * n1y + j] = -conv_v - dpdy + nu * (dxxv + dyyv);
    if (i >= 1 && i < n1x - 1 && j >= 1 && j < n1y - 1) {
        const int mx = n1x - 2, my = n1y - 2;
        Rp[(size_t)b * mx * my + (i - 1) * my + (j - 1)] = -beta2[b] * (dudx + dvdy);
    }
}

// Start of an RK4 step: keep phi^n, reduce max|u|, max|v| of phi^n and form dt (Q9: start-of-step state).
__global__ void __launch_bounds__(RED_THREADS)
k_step_begin(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ p,
             double* __restrict__ u0, double* __restrict__ v0, double* __restrict__ p0, double* __restrict__ dt,
             const double* __restrict__ Re, const double* __restrict__ beta2, const double* __restrict__ CFL,
             const double* __restrict__ lidvel, double dx_min, double dy_min, int nfull, int ninner,
             CavityStatus* __restrict__ status) {
    __shared__ double sh[2 * 32];
    const int b = blockIdx.x;
    if (!status[b].active) return;
    const size_t of = (size_t)b * nfull, oi = (size_t)b * ninner;
    double m[2] = {0.0, 0.0};
    for (int e = threadIdx.x; e < nfull; e += blockDim.x) {
        const double uu = u[of + e], vv = v[of + e];
        u0[of + e] = uu;
        v0[of + e] = vv;
        // NaN-propagating max, as np.max(np.abs(.)) propagates NaN
        m[0] = (fabs(uu) > m[0] || uu != uu) ? fabs(uu) : m[0];
        m[1] = (fabs(vv) > m[1] || vv != vv) ? fabs(vv) : m[1];
    }
    for (int e = threadIdx.x; e < ninner; e += blockDim.x) p0[oi + e] = p[oi + e];
    // fmax drops NaN; carry NaN through an explicit flag instead
    double nanflag[2] = {m[0] != m[0] ? 1.0 : 0.0, m[1] != m[1] ? 1.0 : 0.0};
    if (m[0] != m[0]) m[0] = 0.0;
    if (m[1] != m[1]) m[1] = 0.0;
    block_reduce<2, true>(m, sh);
    block_reduce<2, true>(nanflag, sh);
    if (threadIdx.x == 0) {
        const double um = nanflag[0] > 0.0 ? nan("") : m[0], vm = nanflag[1] > 0.0 ? nan("") : m[1];
        status[b].umax = um;
        status[b].vmax = vm;
        // Python max(np.max(abs(u)), lid): NaN first argument wins -> dt = NaN, as in the reference
        const double uarg = (lidvel[b] > um) ? lidvel[b] : um;
        const double varg = (1e-10 > vm) ? 1e-10 : vm;
        const double nu = 1.0 / Re[b];
        const double lam_x = (uarg + sqrt(uarg * uarg + beta2[b])) / dx_min + nu / (dx_min * dx_min);
        const double lam_y = (varg + sqrt(varg * varg + beta2[b])) / dy_min + nu / (dy_min * dy_min);
        dt[b] = CFL[b] / (lam_x + lam_y);
    }
}

// Stage update phi = phi^n + alpha dt R followed by the wall / lid boundary conditions
// (spectral/sg.py:348-385: W, E, S walls zero, lid column last so both top corners take the lid value).
__global__ void __launch_bounds__(256)
k_stage_update(double* __restrict__ u, double* __restrict__ v, double* __restrict__ p, const double* __restrict__ u0,
               const double* __restrict__ v0, const double* __restrict__ p0, const double* __restrict__ Ru,
               const double* __restrict__ Rv, const double* __restrict__ Rp, const double* __restrict__ dt,
               const double* __restrict__ ulid, const double* __restrict__ vlid, double alpha, int n1x, int n1y,
               int write_p, const CavityStatus* __restrict__ status) {
    const int b = blockIdx.z;
    if (!status[b].active) return;
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= n1x || j >= n1y) return;
    const size_t off = (size_t)b * n1x * n1y;
    const int e = i * n1y + j;
    const double adt = alpha * dt[b];
    double un = u0[off + e] + adt * Ru[off + e];
    double vn = v0[off + e] + adt * Rv[off + e];
    if (j == n1y - 1) {
        un = ulid[(size_t)b * n1x + i];
        vn = vlid[(size_t)b * n1x + i];
    } else if (i == 0 || i == n1x - 1 || j == 0) {
        un = 0.0;
        vn = 0.0;
    }
    u[off + e] = un;
    v[off + e] = vn;
    const int mx = n1x - 2, my = n1y - 2;
    if (write_p && i < mx && j < my) {
        const size_t oi = (size_t)b * mx * my + i * my + j;
        p[oi] = p0[oi] + adt * Rp[oi];
    }
}

// End of an RK4 step: u_res, v_res = ||phi^{n+1} - phi^n|| / (||phi^n|| + 1e-12) (base.py:250-258,
// multigrid/fsg.py:992-993), then the convergence bookkeeping.
__global__ void __launch_bounds__(RED_THREADS)
k_step_end(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ u0,
           const double* __restrict__ v0, int nfull, const double* __restrict__ tol, int warmup, long long max_iter,
           int nan_exit, CavityStatus* __restrict__ status) {
    __shared__ double sh[4 * 32];
    const int b = blockIdx.x;
    if (!status[b].active) return;
    const size_t of = (size_t)b * nfull;
    double s[4] = {0, 0, 0, 0};
    for (int e = threadIdx.x; e < nfull; e += blockDim.x) {
        const double a = u0[of + e], c = v0[of + e];
        const double du = u[of + e] - a, dv = v[of + e] - c;
        s[0] = fma(du, du, s[0]);
        s[1] = fma(a, a, s[1]);
        s[2] = fma(dv, dv, s[2]);
        s[3] = fma(c, c, s[3]);
    }
    block_reduce<4, false>(s, sh);
    if (threadIdx.x == 0) {
        const double ures = sqrt(s[0]) / (sqrt(s[1]) + 1e-12);
        const double vres = sqrt(s[2]) / (sqrt(s[3]) + 1e-12);
        CavityStatus st = status[b];
        update_status(st, ures, vres, tol ? tol[b] : -1.0, warmup, max_iter, nan_exit);
        status[b] = st;
    }
}

__global__ void k_reset_status(CavityStatus* status, int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    CavityStatus s;
    s.iters = 0;
    s.active = 1;
    s.converged = 0;
    s.diverged = 0;
    s.pad = 0;
    s.u_res = s.v_res = 0.0;
    s.umax = s.vmax = 0.0;
    status[b] = s;
}

__global__ void k_count_active(const CavityStatus* status, int B, int* n_active) {
    __shared__ int cnt;
    if (threadIdx.x == 0) cnt = 0;
    __syncthreads();
    int c = 0;
    for (int b = threadIdx.x; b < B; b += blockDim.x) c += status[b].active ? 1 : 0;
    atomicAdd(&cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) *n_active = cnt;
}

// ||R_u||, ||R_v||, ||R_p|| per cavity (spectral/sg.py:463-473).
__global__ void __launch_bounds__(RED_THREADS)
k_residual_norms(const double* __restrict__ Ru, const double* __restrict__ Rv, const double* __restrict__ Rp,
                 int nfull, int ninner, double* __restrict__ out) {
    __shared__ double sh[3 * 32];
    const int b = blockIdx.x;
    double s[3] = {0, 0, 0};
    for (int e = threadIdx.x; e < nfull; e += blockDim.x) {
        const double a = Ru[(size_t)b * nfull + e], c = Rv[(size_t)b * nfull + e];
        s[0] = fma(a, a, s[0]);
        s[1] = fma(c, c, s[1]);
    }
    for (int e = threadIdx.x; e < ninner; e += blockDim.x) {
        const double a = Rp[(size_t)b * ninner + e];
        s[2] = fma(a, a, s[2]);
    }
    block_reduce<3, false>(s, sh);
    if (threadIdx.x == 0) {
        out[b * 3 + 0] = sqrt(s[0]);
        out[b * 3 + 1] = sqrt(s[1]);
        out[b * 3 + 2] = sqrt(s[2]);
    }
}

// omega = Dx v - u Dy^T on all nodes (spectral/sg.py:516-523).
__global__ void __launch_bounds__(256)
k_vorticity(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ Dx,
            const double* __restrict__ DyT, double* __restrict__ om, int n1x, int n1y) {
    const int b = blockIdx.z;
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= n1x || j >= n1y) return;
    const size_t off = (size_t)b * n1x * n1y;
    double dvdx = 0, dudy = 0;
    for (int k = 0; k < n1x; ++k) dvdx = fma(Dx[i * n1x + k], v[off + k * n1y + j], dvdx);
    for (int k = 0; k < n1y; ++k) dudy = fma(u[off + i * n1y + k], DyT[k * n1y + j], dudy);
    om[off + i * n1y + j] = dvdx - dudy;
}

// E = 1/2 sum W (u^2 + v^2), Z = 1/2 sum W omega^2, P = 1/2 sum W |grad omega|^2 (spectral/sg.py:495-550).
__global__ void __launch_bounds__(RED_THREADS)
k_diagnostics(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ om,
              const double* __restrict__ Dx, const double* __restrict__ DyT, const double* __restrict__ W,
              double* __restrict__ out, int n1x, int n1y) {
    __shared__ double sh[3 * 32];
    const int b = blockIdx.x;
    const int nfull = n1x * n1y;
    const size_t off = (size_t)b * nfull;
    double s[3] = {0, 0, 0};
    for (int e = threadIdx.x; e < nfull; e += blockDim.x) {
        const int i = e / n1y, j = e % n1y;
        double gx = 0, gy = 0;
        for (int k = 0; k < n1x; ++k) gx = fma(Dx[i * n1x + k], om[off + k * n1y + j], gx);
        for (int k = 0; k < n1y; ++k) gy = fma(om[off + i * n1y + k], DyT[k * n1y + j], gy);
        const double w = W[e], uu = u[off + e], vv = v[off + e], oo = om[off + e];
        s[0] += w * (uu * uu + vv * vv);
        s[1] += w * oo * oo;
        s[2] += w * (gx * gx + gy * gy);
    }
    block_reduce<3, false>(s, sh);
    if (threadIdx.x == 0) {
        out[b * 3 + 0] = 0.5 * s[0];
        out[b * 3 + 1] = 0.5 * s[1];
        out[b * 3 + 2] = 0.5 * s[2];
    }
}

// ---- per-iteration history (base.solve, base.py:243-313) ----------------------------------------------------------
// One record per cavity and iteration: [rel_iter, ||R_u||, ||R_v||, ||R_p||, E, Z, P].  `it` is the 0-based index of the
// iteration that has just been taken; a cavity records iff it took that step (status.iters == it + 1).
constexpr int HIST_W = 7;

// rel_iter = max(u_res, v_res) (Python max: first wins unless the second is strictly greater) and the norms of the LAST
// stage's residual on all nodes, as _compute_algebraic_residuals reads them from arrays.R_* (spectral/sg.py:463-473).
__global__ void __launch_bounds__(RED_THREADS)
k_hist_norms(const double* __restrict__ Ru, const double* __restrict__ Rv, const double* __restrict__ Rp, int nfull,
             int ninner, const CavityStatus* __restrict__ status, long long it, long long slot, size_t cap,
             double* __restrict__ hist) {
    __shared__ double sh[3 * 32];
    const int b = blockIdx.x;
    if (status[b].iters != it + 1) return;
    double s[3] = {0, 0, 0};
    for (int e = threadIdx.x; e < nfull; e += blockDim.x) {
        const double a = Ru[(size_t)b * nfull + e], c = Rv[(size_t)b * nfull + e];
        s[0] = fma(a, a, s[0]);
        s[1] = fma(c, c, s[1]);
    }
    for (int e = threadIdx.x; e < ninner; e += blockDim.x) {
        const double a = Rp[(size_t)b * ninner + e];
        s[2] = fma(a, a, s[2]);
    }
    block_reduce<3, false>(s, sh);
    if (threadIdx.x == 0) {
        double* r = hist + ((size_t)b * cap + slot) * HIST_W;
        const double ur = status[b].u_res, vr = status[b].v_res;
        r[0] = (vr > ur) ? vr : ur;
        r[1] = sqrt(s[0]);
        r[2] = sqrt(s[1]);
        r[3] = sqrt(s[2]);
    }
}

// E, Z, P of the new state (spectral/sg.py:495-550); om = vorticity from k_vorticity.
__global__ void __launch_bounds__(RED_THREADS)
k_hist_diag(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ om,
            const double* __restrict__ Dx, const double* __restrict__ DyT, const double* __restrict__ W, int n1x, int n1y,
            const CavityStatus* __restrict__ status, long long it, long long slot, size_t cap, double* __restrict__ hist) {
    __shared__ double sh[3 * 32];
    const int b = blockIdx.x;
    if (status[b].iters != it + 1) return;
    const int nfull = n1x * n1y;
    const size_t off = (size_t)b * nfull;
    double s[3] = {0, 0, 0};
    for (int e = threadIdx.x; e < nfull; e += blockDim.x) {
        const int i = e / n1y, j = e % n1y;
        double gx = 0, gy = 0;
        for (int k = 0; k < n1x; ++k) gx = fma(Dx[i * n1x + k], om[off + k * n1y + j], gx);
        for (int k = 0; k < n1y; ++k) gy = fma(om[off + i * n1y + k], DyT[k * n1y + j], gy);
        const double w = W[e], uu = u[off + e], vv = v[off + e], oo = om[off + e];
        s[0] += w * (uu * uu + vv * vv);
        s[1] += w * oo * oo;
        s[2] += w * (gx * gx + gy * gy);
    }
    block_reduce<3, false>(s, sh);
    if (threadIdx.x == 0) {
        double* r = hist + ((size_t)b * cap + slot) * HIST_W;
        r[4] = 0.5 * s[0];
        r[5] = 0.5 * s[1];
        r[6] = 0.5 * s[2];
    }
}

// u = v = p = 0 then the lid column.
__global__ void k_init_state(double* __restrict__ u, double* __restrict__ v, double* __restrict__ p,
                             const double* __restrict__ ulid, const double* __restrict__ vlid, int n1x, int n1y) {
    const int b = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const int nfull = n1x * n1y, ninner = (n1x - 2) * (n1y - 2);
    if (e < nfull) {
        const int i = e / n1y, j = e % n1y;
        const bool lid = (j == n1y - 1);
        u[(size_t)b * nfull + e] = lid ? ulid[(size_t)b * n1x + i] : 0.0;
        v[(size_t)b * nfull + e] = lid ? vlid[(size_t)b * n1x + i] : 0.0;
    }
    if (e < ninner) p[(size_t)b * ninner + e] = 0.0;
}

// Boundary treatment after a level transfer.  quirk=1 is the reference's prolongate_solution
// (multigrid/fsg.py:589-599, axes swapped): row 0 := 0, row -1 := lid_velocity for u (0 for v),
// col 0 := 0, col -1 := 0, then initialize_lid (950-954) rewrites the lid column only.  quirk=0 is the plain
// wall + lid boundary condition.
__global__ void k_transfer_bc(double* __restrict__ u, double* __restrict__ v, const double* __restrict__ ulid,
                              const double* __restrict__ vlid, const double* __restrict__ lidvel, int n1x, int n1y,
                              int quirk) {
    const int b = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n1x * n1y) return;
    const int i = e / n1y, j = e % n1y;
    const size_t o = (size_t)b * n1x * n1y + e;
    if (j == n1y - 1) {
        u[o] = ulid[(size_t)b * n1x + i];
        v[o] = vlid[(size_t)b * n1x + i];
    } else if (j == 0) {
        u[o] = 0.0;
        v[o] = 0.0;
    } else if (i == n1x - 1) {
        u[o] = quirk ? lidvel[b] : 0.0;
        v[o] = 0.0;
    } else if (i == 0) {
        u[o] = 0.0;
        v[o] = 0.0;
    }
}

}  // namespace ldc
