// Resident path: one CTA advances one cavity through many RK4 steps with the stage state (u, v, p) and the
// 1-D operators resident in shared memory.  Every contraction of the residual is an FP64 tensor-core tile
// (mma.sync m8n8k4 f64 = SASS DMMA.8x8x4; sm_100a lowers every f64 mma shape to it) fed from shared memory;
// the convective / viscous / continuity terms, the RK stage update, the boundary conditions, the adaptive
// time step and the convergence norms are fused around the tiles with warp-shuffle reductions.
//
// Only interior unknowns are contracted.  With n = N+1 nodes the m = N-1 interior rows/columns of a field
// form an m x m matrix (padded with zeros to MP, a multiple of 16); the four boundary lines enter as rank-1
// corrections, which is exact:   (Dx U)[i,j] = sum_{k interior} Dx[i,k] U[k,j] + Dx[i,0] U[0,j] + Dx[i,N] U[N,j].
// Residuals at boundary nodes are never used by the time stepper (the boundary conditions overwrite them).
// The pressure path p_full = Ix p Iy^T, dp/dx = Dx p_full (spectral/sg.py:255-276) folds into one m x m
// operator Gx = (Dx Ix)[1:-1, :], because the interior rows of Ix, Iy are the identity (to rounding).
#pragma once
#include "common.cuh"
#include "peak.cuh"
#include "tiled.cuh"

namespace ldc {

// ---- packed layouts ---------------------------------------------------------------------------------------
// operators (global, r_ops):  NMAT matrices of MP x LD (Dx, Dxx, Gx interior blocks; then Dy, Dyy, Gy unless
// square), then 8 vectors of MP: dxW, dxE, dxxW, dxxE (column 0 / N of Dx, Dxx at interior rows) and
// dyS, dyN, dyyS, dyyN.
// state (global, r_state, per cavity): u, v, p interior blocks MP x LD each, then edge vectors uW, uE, uS, uN,
// vW, vE, vS, vN (MP each, interior-indexed, zero padded), then 8 corner values (u00, u0N, uN0, uNN, v..).
__host__ __device__ inline int res_ld(int MP) { return MP + 4; }  // LD = 4 mod 8: conflict-free fragment loads
__host__ __device__ inline size_t res_state_stride(int MP) { return (size_t)3 * MP * res_ld(MP) + 8 * MP + 8; }
__host__ __device__ inline size_t res_ops_size(int MP, bool square) {
    return (size_t)(square ? 3 : 6) * MP * res_ld(MP) + 8 * MP;
}

__host__ inline int res_pick_mp(int m) {
    if (m <= 16) return 16;
    if (m <= 32) return 32;
    if (m <= 48) return 48;
    if (m <= 64) return 64;
    return 0;
}

// Dx (n1 x n1) -> interior block + boundary columns; G = (D I) full product (n1 x m) -> rows 1..n1-2.
__global__ void k_pack_ops(const double* __restrict__ D, const double* __restrict__ DD, const double* __restrict__ DI,
                           double* __restrict__ mats /* 3 consecutive MP x LD */, double* __restrict__ vecs /* 4 x MP */,
                           int n1, int MP) {
    const int LD = res_ld(MP), m = n1 - 2;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < MP * LD) {
        const int i = e / LD, k = e % LD;
        const bool in = (i < m && k < m);
        mats[e] = in ? D[(i + 1) * n1 + (k + 1)] : 0.0;
        mats[MP * LD + e] = in ? DD[(i + 1) * n1 + (k + 1)] : 0.0;
        mats[2 * MP * LD + e] = in ? DI[(i + 1) * m + k] : 0.0;
    }
    if (e < MP) {
        const bool in = e < m;
        vecs[e] = in ? D[(e + 1) * n1] : 0.0;
        vecs[MP + e] = in ? D[(e + 1) * n1 + n1 - 1] : 0.0;
        vecs[2 * MP + e] = in ? DD[(e + 1) * n1] : 0.0;
        vecs[3 * MP + e] = in ? DD[(e + 1) * n1 + n1 - 1] : 0.0;
    }
}

// canonical (u, v, p) -> packed state
__global__ void k_pack_state(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ p,
                             double* __restrict__ st, size_t stride, int n1, int MP) {
    const int b = blockIdx.y, LD = res_ld(MP), m = n1 - 2;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    double* s = st + (size_t)b * stride;
    const double *ub = u + (size_t)b * n1 * n1, *vb = v + (size_t)b * n1 * n1, *pb = p + (size_t)b * m * m;
    if (e < MP * LD) {
        const int i = e / LD, j = e % LD;
        const bool in = (i < m && j < m);
        s[e] = in ? ub[(i + 1) * n1 + j + 1] : 0.0;
        s[MP * LD + e] = in ? vb[(i + 1) * n1 + j + 1] : 0.0;
        s[2 * MP * LD + e] = in ? pb[i * m + j] : 0.0;
    }
    double* ed = s + 3 * MP * LD;
    if (e < MP) {
        const bool in = e < m;
        const int N = n1 - 1;
        ed[e] = in ? ub[e + 1] : 0.0;                        // uW: row 0
        ed[MP + e] = in ? ub[N * n1 + e + 1] : 0.0;          // uE: row N
        ed[2 * MP + e] = in ? ub[(e + 1) * n1] : 0.0;        // uS: column 0
        ed[3 * MP + e] = in ? ub[(e + 1) * n1 + N] : 0.0;    // uN: column N (lid)
        ed[4 * MP + e] = in ? vb[e + 1] : 0.0;
        ed[5 * MP + e] = in ? vb[N * n1 + e + 1] : 0.0;
        ed[6 * MP + e] = in ? vb[(e + 1) * n1] : 0.0;
        ed[7 * MP + e] = in ? vb[(e + 1) * n1 + N] : 0.0;
    }
    if (e < 8) {
        const int N = n1 - 1;
        const double* f = e < 4 ? ub : vb;
        const int c = e & 3;
        const int i = (c & 2) ? N : 0, j = (c & 1) ? N : 0;
        ed[8 * MP + e] = f[i * n1 + j];
    }
}

// packed state -> canonical (u, v, p)
__global__ void k_unpack_state(double* __restrict__ u, double* __restrict__ v, double* __restrict__ p,
                               const double* __restrict__ st, size_t stride, int n1, int MP) {
    const int b = blockIdx.y, LD = res_ld(MP), m = n1 - 2, N = n1 - 1;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n1 * n1) return;
    const double* s = st + (size_t)b * stride;
    const double* ed = s + 3 * MP * LD;
    const int i = e / n1, j = e % n1;
    double uu, vv;
    const bool ci = (i == 0 || i == N), cj = (j == 0 || j == N);
    if (ci && cj) {
        const int c = (i == N ? 2 : 0) + (j == N ? 1 : 0);
        uu = ed[8 * MP + c];
        vv = ed[8 * MP + 4 + c];
    } else if (i == 0) {
        uu = ed[j - 1];
        vv = ed[4 * MP + j - 1];
    } else if (i == N) {
        uu = ed[MP + j - 1];
        vv = ed[5 * MP + j - 1];
    } else if (j == 0) {
        uu = ed[2 * MP + i - 1];
        vv = ed[6 * MP + i - 1];
    } else if (j == N) {
        uu = ed[3 * MP + i - 1];
        vv = ed[7 * MP + i - 1];
    } else {
        uu = s[(i - 1) * LD + j - 1];
        vv = s[MP * LD + (i - 1) * LD + j - 1];
        p[(size_t)b * m * m + (i - 1) * m + j - 1] = s[2 * MP * LD + (i - 1) * LD + j - 1];
    }
    u[(size_t)b * n1 * n1 + e] = uu;
    v[(size_t)b * n1 * n1 + e] = vv;
}

// NaN-propagating max (np.max propagates NaN; fmax would drop it)
__device__ __forceinline__ double nanmax(double a, double b) { return (b > a || b != b) ? b : a; }

struct ResParams {
    const double* ops;
    int square;
    double* state;
    size_t state_stride;
    const double *Re, *beta2, *CFL, *lidvel, *ulid, *vlid;
    int n1, m, K4;
    double dx_min, dy_min;
    CavityStatus* status;
    const double* tol;
    int warmup, nan_exit, freeze;
    long long max_iter;    // absolute cap on status.iters for the whole solve
    long long launch_end;  // a cavity leaves this launch once status.iters reaches launch_end
    long long chunk;       // steps a CTA runs on one cavity before it hands the cavity back to the queue
    // work queue of cavities (bounded MPMC ring, one slot per cavity): see CavityQueue
    int* q_data;           // [B] cavity index
    unsigned* q_seq;       // [B] slot sequence numbers
    unsigned* q_ctl;       // head, tail, remaining
    int B;
};

// Coherent (L2) access to a cavity's status: cavities migrate between SMs, so L1 must be bypassed.
__device__ __forceinline__ CavityStatus load_status(const CavityStatus* p) {
    CavityStatus s;
    s.iters = __ldcg(&p->iters);
    s.active = __ldcg(&p->active);
    s.converged = __ldcg(&p->converged);
    s.diverged = __ldcg(&p->diverged);
    s.pad = 0;
    s.u_res = __ldcg(&p->u_res);
    s.v_res = __ldcg(&p->v_res);
    s.umax = __ldcg(&p->umax);
    s.vmax = __ldcg(&p->vmax);
    return s;
}
__device__ __forceinline__ void store_status(CavityStatus* p, const CavityStatus& s) {
    __stcg(&p->iters, s.iters);
    __stcg(&p->active, s.active);
    __stcg(&p->converged, s.converged);
    __stcg(&p->diverged, s.diverged);
    __stcg(&p->u_res, s.u_res);
    __stcg(&p->v_res, s.v_res);
    __stcg(&p->umax, s.umax);
    __stcg(&p->vmax, s.vmax);
}

// Work queue of the resident kernel.  A launch runs min(B, #SM) persistent CTAs; each pops a cavity, advances it
// by `chunk` RK4 steps with the state in shared memory, writes it back to L2 and, if the cavity still has steps
// to do in this launch, pushes it to the tail.  B > #SM therefore costs no wave quantisation: every SM stays
// busy until the whole batch is done, and a cavity that converges early frees its SM for the others.
// Protocol: bounded ring with per-slot sequence numbers (slot free for ticket t when seq == t, full when
// seq == t + 1); every CTA of the launch is co-resident, so the spins cannot deadlock.
struct CavityQueue {
    template <class PT>
    __device__ static int pop(const PT& P) {
        const unsigned h = atomicAdd(&P.q_ctl[0], 1u);
        const unsigned slot = h % (unsigned)P.B;
        volatile unsigned* seq = P.q_seq + slot;
        volatile unsigned* remaining = P.q_ctl + 2;
        while (*seq != h + 1u) {
            if (*remaining == 0u) return -1;
            __nanosleep(200);
        }
        __threadfence();
        const int b = *(volatile int*)(P.q_data + slot);
        __threadfence();
        *seq = h + (unsigned)P.B;
        return b;
    }
    template <class PT>
    __device__ static void push(const PT& P, int b) {
        const unsigned t = atomicAdd(&P.q_ctl[1], 1u);
        const unsigned slot = t % (unsigned)P.B;
        volatile unsigned* seq = P.q_seq + slot;
        while (*seq != t) __nanosleep(100);
        *(volatile int*)(P.q_data + slot) = b;
        __threadfence();
        *seq = t + 1u;
    }
    template <class PT>
    __device__ static void retire(const PT& P) { atomicSub(&P.q_ctl[2], 1u); }
};

// Fill the queue with the cavities that still have work in this launch, in index order (one block).
__global__ void __launch_bounds__(1024) k_queue_init(const CavityStatus* __restrict__ status, int B, long long launch_end,
                                                     int* __restrict__ q_data, unsigned* __restrict__ q_seq,
                                                     unsigned* __restrict__ q_ctl) {
    __shared__ int warp_cnt[32];
    __shared__ int base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) base = 0;
    __syncthreads();
    for (int b0 = 0; b0 < B; b0 += blockDim.x) {
        const int b = b0 + tid;
        const bool live = b < B && status[b].active && status[b].iters < launch_end;
        const unsigned bal = __ballot_sync(0xffffffffu, live);
        if (lane == 0) warp_cnt[warp] = __popc(bal);
        __syncthreads();
        int off = base;
        for (int w = 0; w < warp; ++w) off += warp_cnt[w];
        if (live) q_data[off + __popc(bal & ((1u << lane) - 1u))] = b;
        __syncthreads();
        if (tid == 0) {
            int tot = 0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += warp_cnt[w];
            base += tot;
        }
        __syncthreads();
    }
    const int n = base;
    for (int i = tid; i < B; i += blockDim.x) q_seq[i] = i < n ? (unsigned)i + 1u : (unsigned)i;
    if (tid == 0) {
        q_ctl[0] = 0u;
        q_ctl[1] = (unsigned)n;
        q_ctl[2] = (unsigned)n;
    }
}

template <int MP, int WM, int WN>
struct ResCfg {
    static constexpr int LD = MP + 4;
    static constexpr int TM = MP / WM, TN = MP / WN;
    static constexpr int MT = TM / 8, NT = TN / 8;
    static constexpr int WARPS = WM * WN, THREADS = WARPS * 32;
    static constexpr int MAT = MP * LD;
    static_assert(TM % 8 == 0 && TN % 8 == 0, "warp tile must be a multiple of the 8x8 DMMA tile");
    __host__ __device__ static constexpr size_t smem_doubles(bool square) {
        return (size_t)(square ? 3 : 6) * MAT + 8 * MP   // operators + their boundary vectors
               + 3 * MAT                                  // stage state u, v, p
               + 8 * MP + 8                               // current edges + corners
               + 2 * MP + 4                               // lid interior + lid corners (u0N, uNN, v0N, vNN)
               + 8 * WARPS;                               // reduction scratch
    }
};

}  // namespace ldc
