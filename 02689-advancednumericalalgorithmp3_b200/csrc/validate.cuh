// L2 error of every cavity against a stored reference solution given at scattered points
// (compute_validation_errors, src/solvers/base.py:970-1054; _evaluate_at_points 1122-1160): bilinear evaluation of
// u, v on the collocation grid -- SciPy's RegularGridInterpolator(method="linear") -- then
//   err_f = ||f(x_q) - ref_f||_2 / (||ref_f||_2 + 1e-12)   over the masked points whose interpolated values are not NaN.
// HBM/L2-bound gather + reduction: 4 neighbours x 2 fields per point; reductions in a fixed order (deterministic).
#pragma once
#include "tiled.cuh"

namespace ldc {

constexpr int VAL_CHUNKS = 16;

// partial sums of one chunk of points of one cavity: [sum du^2, sum ref_u^2, sum dv^2, sum ref_v^2, n_valid]
__global__ void __launch_bounds__(RED_THREADS)
k_val_partial(const double* __restrict__ u, const double* __restrict__ v, int nfull, int n1y, long long nq,
              const int* __restrict__ cell, const double* __restrict__ tx, const double* __restrict__ ty,
              const int* __restrict__ mask, const double* __restrict__ ref_u, const double* __restrict__ ref_v,
              double* __restrict__ partial) {
    __shared__ double sh[5 * 32];
    const int b = blockIdx.y, c = blockIdx.x;
    const double *ub = u + (size_t)b * nfull, *vb = v + (size_t)b * nfull;
    const long long per = (nq + VAL_CHUNKS - 1) / VAL_CHUNKS;
    const long long q0 = c * per, q1 = (q0 + per < nq) ? q0 + per : nq;
    double s[5] = {0, 0, 0, 0, 0};
    for (long long q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
        if (!mask[q]) continue;
        const int o = cell[q];   // node (ix, iy) of the cell's lower corner: o = ix * n1y + iy
        const double ax = tx[q], ay = ty[q];
        // SciPy order: corners (y0,x0) (y0,x1) (y1,x0) (y1,x1), weight = w_y * w_x, accumulated left to right
        const double w00 = __dmul_rn(1.0 - ay, 1.0 - ax), w01 = __dmul_rn(1.0 - ay, ax);
        const double w10 = __dmul_rn(ay, 1.0 - ax), w11 = __dmul_rn(ay, ax);
        double fu = __dmul_rn(ub[o], w00);
        fu = __dadd_rn(fu, __dmul_rn(ub[o + n1y], w01));
        fu = __dadd_rn(fu, __dmul_rn(ub[o + 1], w10));
        fu = __dadd_rn(fu, __dmul_rn(ub[o + n1y + 1], w11));
        double fv = __dmul_rn(vb[o], w00);
        fv = __dadd_rn(fv, __dmul_rn(vb[o + n1y], w01));
        fv = __dadd_rn(fv, __dmul_rn(vb[o + 1], w10));
        fv = __dadd_rn(fv, __dmul_rn(vb[o + n1y + 1], w11));
        if (fu != fu || fv != fv) continue;   // ~(isnan(u) | isnan(v)), base.py:1028
        const double ru = ref_u[q], rv = ref_v[q];
        const double du = fu - ru, dv = fv - rv;
        s[0] = fma(du, du, s[0]);
        s[1] = fma(ru, ru, s[1]);
        s[2] = fma(dv, dv, s[2]);
        s[3] = fma(rv, rv, s[3]);
        s[4] += 1.0;
    }
    block_reduce<5, false>(s, sh);
    if (threadIdx.x == 0) {
        double* out = partial + ((size_t)b * VAL_CHUNKS + c) * 5;
#pragma unroll
        for (int k = 0; k < 5; ++k) out[k] = s[k];
    }
}

__global__ void k_val_final(const double* __restrict__ partial, int B, double* __restrict__ out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double s[5] = {0, 0, 0, 0, 0};
    for (int c = 0; c < VAL_CHUNKS; ++c)
#pragma unroll
        for (int k = 0; k < 5; ++k) s[k] += partial[((size_t)b * VAL_CHUNKS + c) * 5 + k];
    out[b * 3 + 0] = sqrt(s[0]) / (sqrt(s[1]) + 1e-12);
    out[b * 3 + 1] = sqrt(s[2]) / (sqrt(s[3]) + 1e-12);
    out[b * 3 + 2] = s[4];
}

}  // namespace ldc
