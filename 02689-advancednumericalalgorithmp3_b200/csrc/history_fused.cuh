// Fused history stepper for small grids: base.solve's per-iteration records (base.py:243-313) without a launch per stage.
//
// ldc_solve_history needs, after EVERY pseudo-step, the norms of the last stage's residual on ALL nodes (boundary rows
// included), and energy / enstrophy / palinstrophy of the new state -- quantities the interior-only DMMA kernels never
// form.  The launch-per-stage path delivers them with 22 launches per iteration (117 us at N = 16).  Here one CTA per cavity
// keeps the operators, the full-grid state and every work array of the reference (datastructures.py:282-353: u, v, p,
// u_prev.., R_u.., p_full) in shared memory and runs whole chunks of iterations in one launch.  The arithmetic per node is
// the tiled path's (same loops, same fma order: tiled.cuh), plain DFMA: the grids that fit (N <= 36) are latency-bound, not
// flop-bound.  BASELINE config 1 (SGSolver Re=100 N=32, 59 649 iterations) is the case this exists for.
#pragma once
#include "tiled.cuh"

namespace ldc {

constexpr int HF_THREADS = 384;   // 12 warps: the blocked residual has ceil(n1x / 4) x n1y work items (297 at N = 32)
constexpr int HF_IB = 4;          // rows per thread in the residual loop

struct HistFusedParams {
    double *u, *v, *p;                                        // canonical state [B][n1x][n1y], [B][mx][my] (in / out)
    const double *Dx, *Dxx, *DyT, *DyyT, *Ix, *IyT;           // as the tiled path holds them
    const double *Re, *beta2, *CFL, *lidvel, *ulid, *vlid, *W, *tol;
    CavityStatus* status;
    double* hist;                                             // [B][cap][HIST_W]
    size_t cap;
    long long start_iter, end_iter, max_iter;                 // this launch takes iterations [start_iter, end_iter)
    int warmup, nan_exit, freeze;
    int n1x, n1y;
    double dx_min, dy_min;
};

__host__ __device__ inline size_t hist_fused_smem_doubles(int n1x, int n1y) {
    const size_t nf = (size_t)n1x * n1y, mx = n1x - 2, my = n1y - 2, ni = mx * my;
    return 2 * (size_t)n1x * n1x + 2 * (size_t)n1y * n1y + n1x * mx + my * n1y   // Dx Dxx DyT DyyT Ix IyT
           + nf                                                                   // W
           + 8 * nf                                                               // u v u0 v0 Ru Rv pfull omega
           + 3 * ni                                                               // p p0 Rp
           + n1x * my                                                             // ptmp = Ix p
           + 8 * 32;                                                              // reduction scratch
}

__global__ void __launch_bounds__(HF_THREADS, 1) k_history_fused(const HistFusedParams P) {
    extern __shared__ __align__(16) double smem[];
    const int b = blockIdx.x, tid = threadIdx.x, nthr = blockDim.x;
    const int n1x = P.n1x, n1y = P.n1y, mx = n1x - 2, my = n1y - 2;
    const int nf = n1x * n1y, ni = mx * my, nib = (n1x + HF_IB - 1) / HF_IB;
    double* sDx = smem;
    double* sDxx = sDx + n1x * n1x;
    double* sDyT = sDxx + n1x * n1x;
    double* sDyyT = sDyT + n1y * n1y;
    double* sIx = sDyyT + n1y * n1y;
    double* sIyT = sIx + n1x * mx;
    double* sW = sIyT + my * n1y;
    double* su = sW + nf;
    double* sv = su + nf;
    double* su0 = sv + nf;
    double* sv0 = su0 + nf;
    double* sRu = sv0 + nf;
    double* sRv = sRu + nf;
    double* spf = sRv + nf;
    double* som = spf + nf;
    double* sp = som + nf;
    double* sp0 = sp + ni;
    double* sRp = sp0 + ni;
    double* spt = sRp + ni;
    double* sh = spt + n1x * my;

    CavityStatus st = P.status[b];
    if (!st.active) return;
    for (int e = tid; e < n1x * n1x; e += nthr) { sDx[e] = P.Dx[e]; sDxx[e] = P.Dxx[e]; }
    for (int e = tid; e < n1y * n1y; e += nthr) { sDyT[e] = P.DyT[e]; sDyyT[e] = P.DyyT[e]; }
    for (int e = tid; e < n1x * mx; e += nthr) sIx[e] = P.Ix[e];
    for (int e = tid; e < my * n1y; e += nthr) sIyT[e] = P.IyT[e];
    for (int e = tid; e < nf; e += nthr) {
        sW[e] = P.W[e];
        su[e] = P.u[(size_t)b * nf + e];
        sv[e] = P.v[(size_t)b * nf + e];
    }
    for (int e = tid; e < ni; e += nthr) sp[e] = P.p[(size_t)b * ni + e];
    const double Re = P.Re[b], nu = 1.0 / Re, beta2 = P.beta2[b], CFL = P.CFL[b], lidvel = P.lidvel[b];
    const double tol = P.tol ? P.tol[b] : -1.0;
    const double* ulid = P.ulid + (size_t)b * n1x;
    const double* vlid = P.vlid + (size_t)b * n1x;
    __syncthreads();

    for (long long it = P.start_iter; it < P.end_iter && st.active; ++it) {
#include "history_fused_step.inc"
    }

    for (int e = tid; e < nf; e += nthr) {
        P.u[(size_t)b * nf + e] = su[e];
        P.v[(size_t)b * nf + e] = sv[e];
    }
    for (int e = tid; e < ni; e += nthr) P.p[(size_t)b * ni + e] = sp[e];
    if (tid == 0) P.status[b] = st;
}

}  // namespace ldc
