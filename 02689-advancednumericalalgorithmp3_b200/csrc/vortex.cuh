// Streamfunction and vortex metrics of a batch of cavities on the device.
//
// The reference solves lap(psi) = -omega, psi = 0 on the walls, with a sparse Kronecker Laplacian and SuperLU
// (spectral/sg.py:556-619; 4.65 s per call at N = 64) and then scans psi and omega for the primary vortex, the three
// corner vortices and the vorticity of largest magnitude (spectral/sg.py:621-743).  Restricted to the interior nodes the
// system is the Sylvester equation A psi + psi B^T = F, A = Dxx[1:-1,1:-1], B = Dyy[1:-1,1:-1], F = -omega_int.  With the
// eigen-decompositions A = Va La Va^-1, B = Vb Lb Vb^-1 (two per polynomial order, computed once on the host):
//     psi = Va ((Va^-1 F Vb^-T) / (la_i + lb_j)) Vb^T
// i.e. four batched GEMMs, followed by one step of iterative refinement against A, B themselves (A is not normal), and a
// fixed-order argmin / argmax scan per cavity.  Everything runs on the state already resident in the handle.
#pragma once
#include "tiled.cuh"

namespace ldc {

// F[b][i][j] = -om[b][i + 1][j + 1]
__global__ void k_vm_interior_neg(const double* __restrict__ om, double* __restrict__ F, int n1x, int n1y) {
    const int b = blockIdx.y, mx = n1x - 2, my = n1y - 2;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= mx * my) return;
    const int i = e / my, j = e % my;
    F[(size_t)b * mx * my + e] = -om[(size_t)b * n1x * n1y + (i + 1) * n1y + j + 1];
}

// G[b][i][j] *= inv[i][j]   (inv = 1 / (la_i + lb_j))
__global__ void k_vm_scale(double* __restrict__ G, const double* __restrict__ inv, int mm) {
    const int b = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < mm) G[(size_t)b * mm + e] *= inv[e];
}

// R = F - (AP + PB)
__global__ void k_vm_residual(const double* __restrict__ F, const double* __restrict__ AP, const double* __restrict__ PB,
                              double* __restrict__ R, int mm) {
    const int b = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < mm) {
        const size_t o = (size_t)b * mm + e;
        R[o] = F[o] - (AP[o] + PB[o]);
    }
}

// psi_full (zero on the boundary) = psi_int + corr
__global__ void k_vm_embed(const double* __restrict__ P0, const double* __restrict__ corr, double* __restrict__ psi, int n1x,
                           int n1y) {
    const int b = blockIdx.y, mx = n1x - 2, my = n1y - 2;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n1x * n1y) return;
    const int i = e / n1y, j = e % n1y;
    double val = 0.0;
    if (i > 0 && i < n1x - 1 && j > 0 && j < n1y - 1) {
        const size_t o = (size_t)b * mx * my + (i - 1) * my + (j - 1);
        val = P0[o] + corr[o];
    }
    psi[(size_t)b * n1x * n1y + e] = val;
}

constexpr int VM_OUT = 19;
constexpr int VM_THREADS = 256;

// (value, index) candidates: `better` decides, ties go to the lower flat index (NumPy's argmin / argmax: first occurrence)
struct VmBest {
    double val;
    int idx;
};
template <bool MIN>
__device__ __forceinline__ void vm_take(VmBest& a, double val, int idx) {
    // NaN never wins a comparison here; NumPy would return the first NaN -- diverged cavities are not post-processed
    const bool better = MIN ? (val < a.val) : (val > a.val);
    if (better || (val == a.val && idx < a.idx)) {
        a.val = val;
        a.idx = idx;
    }
}

// One block per cavity.  out[b] = psi_min, psi_min_x, psi_min_y, omega_center, omega_max, omega_max_x, omega_max_y, then
// (psi, omega, x, y) of the BR, BL, TL corner vortices (zeros when the quadrant's maximum of psi is not positive).
__global__ void __launch_bounds__(VM_THREADS) k_vm_scan(const double* __restrict__ psi, const double* __restrict__ om,
                                                        const double* __restrict__ x, const double* __restrict__ y, int n1x, int n1y,
                                                        double* __restrict__ out) {
    __shared__ double sval[5][VM_THREADS];
    __shared__ int sidx[5][VM_THREADS];
    const int b = blockIdx.x, nf = n1x * n1y, tid = threadIdx.x;
    const double* P = psi + (size_t)b * nf;
    const double* W = om + (size_t)b * nf;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    VmBest best[5] = {{inf, nf}, {-inf, nf}, {-inf, nf}, {-inf, nf}, {-inf, nf}};   // psi min | |omega| max | BR | BL | TL
    for (int e = tid; e < nf; e += VM_THREADS) {
        const int i = e / n1y, j = e % n1y;
        const double ps = P[e], xx = x[i], yy = y[j];
        vm_take<true>(best[0], ps, e);
        vm_take<false>(best[1], fabs(W[e]), e);
        if (xx > 0.5 && yy < 0.5) vm_take<false>(best[2], ps, e);
        if (xx < 0.5 && yy < 0.5) vm_take<false>(best[3], ps, e);
        if (xx < 0.5 && yy > 0.5) vm_take<false>(best[4], ps, e);
    }
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        sval[q][tid] = best[q].val;
        sidx[q][tid] = best[q].idx;
    }
    __syncthreads();
    for (int s = VM_THREADS / 2; s > 0; s >>= 1) {
        if (tid < s) {
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                VmBest a = {sval[q][tid], sidx[q][tid]};
                if (q == 0) vm_take<true>(a, sval[q][tid + s], sidx[q][tid + s]);
                else vm_take<false>(a, sval[q][tid + s], sidx[q][tid + s]);
                sval[q][tid] = a.val;
                sidx[q][tid] = a.idx;
            }
        }
        __syncthreads();
    }
    if (tid == 0) {
        double* o = out + (size_t)b * VM_OUT;
        const int k0 = sidx[0][0] < nf ? sidx[0][0] : 0, k1 = sidx[1][0] < nf ? sidx[1][0] : 0;
        o[0] = P[k0]; o[1] = x[k0 / n1y]; o[2] = y[k0 % n1y]; o[3] = W[k0];
        o[4] = W[k1]; o[5] = x[k1 / n1y]; o[6] = y[k1 % n1y];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int k = sidx[2 + q][0];
            const bool have = k < nf && P[k] > 0.0;
            o[7 + 4 * q + 0] = have ? P[k] : 0.0;
            o[7 + 4 * q + 1] = have ? W[k] : 0.0;
            o[7 + 4 * q + 2] = have ? x[k / n1y] : 0.0;
            o[7 + 4 * q + 3] = have ? y[k % n1y] : 0.0;
        }
    }
}

}  // namespace ldc
