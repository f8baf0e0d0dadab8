// Persistent TMA-fed streamed even-odd path (grids whose parity state does not fit one SM: 34 <= N <= 129).
//
// Same parity algebra and tile shape as stream_eo.cuh (a cavity spread over up to 16 CTAs, each owning a 16 x 16 tile
// of all four parity blocks of both momentum equations), but
//   * ONE cooperative launch runs many RK4 steps: a team of (tiles per cavity) co-resident CTAs pops a cavity from the
//     work queue, advances it `chunk` steps and hands it back.  The stages of a step are separated by a team barrier in
//     global memory (release / acquire counter) instead of a kernel boundary, and the end-of-step reduction (convergence
//     norms, next dt) is done redundantly, in fixed order, by a producer warp of every CTA of the team.
//   * operand panels arrive by TMA bulk copies (cp.async.bulk, SASS UBLKCP) with mbarrier completion into a ring of
//     S2_SLOTS chunk slots: four producer warps issue (one thread's copies are served one after the other, ~360 cycles
//     each whatever their size, copies of different warps side by side: scripts/tma_bw.cu), eight consumer warps wait
//     on the slot's "full" barrier and hand the slot back through its "empty" barrier -- no CTA-wide barrier and no
//     cp.async wait inside the K loop.  setmaxnreg gives the consumers 224 registers and the producers 56.
//   * state and operators live in global memory in a k4-interleaved tile layout: for a 16-line tile and one k4-step (four
//     consecutive k) the 16 x 4 doubles of a block are contiguous, [line][k % 4], and the 18 blocks a side needs (six
//     operator blocks, u, v, p) sit back to back.  A chunk is then ONE contiguous bulk copy per side, and a DMMA fragment
//     load (lane (g, t) -> line g, k t) reads 256 contiguous bytes: conflict-free with no padding.  Every field is stored
//     twice, row side (A operands) and column side (its transpose, B operands).
//   * the two warp groups split K inside every chunk (k4-step 0: group 0, k4-step 1: group 1), swap half of their
//     accumulators through the two last chunk slots, and EACH runs the epilogue of one 8-column half of the tile.
//   * the stage input u, v at the thread's own elements -- at stage 0 also phi^n and p^n, kept for the four stages -- are
//     picked out of the row panels as they pass through shared memory: the epilogue reads nothing from global memory.
// The 16 x 32 shape for large batches is stream_tma_wide.cuh; measurements and the design history are in DESIGN.md 4.6.
#pragma once
#include "stream_eo.cuh"

namespace ldc {

constexpr int S2_HP = ST_HP;                 // padded parity-block edge (64)
constexpr int S2_NK4 = S2_HP / 4;            // k4-steps per block row
constexpr int S2_KC4 = 2;                    // k4-steps per chunk (K-chunk of 8)
constexpr int S2_SLOTS = 4;                  // chunk slots in flight
constexpr int S2_CONSUMERS = 256;            // 8 warps: (k-group) x (u | v equation) x (row half)
constexpr int S2_THREADS = S2_CONSUMERS + 128;   // + the producer's warpgroup (setmaxnreg works on warpgroups)
constexpr int S2_ME = 2 * S2_HP;
// global record of one cavity (doubles): row side | column side | boundary lines (8 x ME) | 8 corners.  A side holds, for
// each 16-line tile and k4-step, S2_NB = 18 blocks of 16 x 4 doubles back to back -- the six operator blocks of the cavity
// (D_SA D_AS nuDD_SS nuDD_AA -G_SA -G_AS; both sides carry the same copy) and the twelve state blocks u(4) v(4) p(4):
//   row side    [rt][kk][blk][r][kq] = X_blk[16 rt + r][4 kk + kq]
//   column side [ct][kk][blk][c][kq] = X_blk[4 kk + kq][16 ct + c]   (state: the transpose; operators: rows 16 ct + c)
// so that ONE bulk copy per side brings everything a K-chunk needs (the TMA unit takes ~350 cycles per copy whatever its
// size: scripts/tma_bw.cu, profiles/r02_tma_bw.txt).
constexpr int S2_NB = 18, S2_BU = 6, S2_BV = 10, S2_BP = 14;
constexpr size_t S2_SIDE = (size_t)4 * S2_NK4 * S2_NB * 64;
constexpr size_t S2_EDGE_OFF = 2 * S2_SIDE;
constexpr size_t S2_STRIDE = S2_EDGE_OFF + 8 * S2_ME + 8;
// one chunk slot (doubles): row side [kk][blk][16][4] | column side
constexpr int S2_SL_A = 0, S2_SL_B = S2_KC4 * S2_NB * 64, S2_SLOT = 2 * S2_KC4 * S2_NB * 64;
constexpr unsigned S2_SIDE_BYTES = S2_KC4 * S2_NB * 64 * sizeof(double);
constexpr unsigned S2_SLOT_BYTES = S2_SLOT * sizeof(double);
constexpr size_t S2_SMEM_DOUBLES = (size_t)S2_SLOTS * S2_SLOT   // chunk ring
                                   + 8 * 256 + 12 * 256         // stage input u, v and phi^n (u, v, p) at the tile's own elements
                                   + 8 * S2_HP + 16 * S2_HP     // boundary-column vectors, boundary lines in parity form
                                   + 4 * 8 * 32                 // dv/dy hand-over v warps -> u warps
                                   + 16 * 8 + 80 + 16           // tile partials, reduction scratch, step info
                                   + 2 * S2_SLOTS + 2;          // mbarriers

__host__ __device__ inline size_t s2_off(int tile, int kk, int blk, int r, int kq) {
    return (((size_t)tile * S2_NK4 + kk) * S2_NB + blk) * 64 + r * 4 + kq;
}

struct S2Params {
    const double* ops;            // eo packed operators (HP = 64): only the 8 boundary-column vectors are read here
    double* rec;                  // [3][B][S2_STRIDE] rotating records
    size_t buf_stride;            // B * S2_STRIDE
    const double *Re, *beta2, *CFL, *lidvel, *ulid, *vlid;
    int n1, m, K4, ntile;         // ntile: tiles per block edge that hold data
    CavityStatus* status;         // status.pad = index of the record holding phi^n
    double* partials;             // [B][16][8]
    const double* tol;
    int warmup, nan_exit, freeze;
    long long max_iter, launch_end, chunk;
    double dx_min, dy_min;
    int* q_data; unsigned* q_seq; unsigned* q_ctl; int B;   // cavity work queue (resident.cuh)
    unsigned* team_bar;           // [teams] arrivals at the team's stage barriers (zero at launch)
    unsigned long long* team_mail;// [teams] (visit number << 32) | (cavity + 1), written by the team's leader
    long long* dbg;               // optional [S2_DBG_STAGES][16] clock samples of CTA 0 (LDC_S2_DEBUG), else null
};
constexpr int S2_DBG_STAGES = 256;

// ---- PTX wrappers ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned s2_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s2_smem(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s2_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s2_smem(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    const unsigned a = s2_smem(bar);
    unsigned ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}
// TMA bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS UBLKCP)
__device__ __forceinline__ void bulk_g2s(double* dst, const double* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(s2_smem(dst)), "l"(src), "r"(bytes), "r"(s2_smem(bar)) : "memory");
}
// the same with shared-memory addresses already converted (the producer keeps them in registers)
__device__ __forceinline__ void mbar_expect_tx_a(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s_a(unsigned dst, const double* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void red_release_add(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- layout conversion kernels -----------------------------------------------------------------------------------------
// canonical (u, v, p) -> record (row-side and column-side copies of the twelve parity blocks + boundary lines)
__global__ void k_s2_pack_state(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ p,
                                double* __restrict__ rec, int n1) {
    const int b = blockIdx.y, HP = S2_HP, m = n1 - 2, h = eo_half(m), N = n1 - 1;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    double* s = rec + (size_t)b * S2_STRIDE;
    const double *ub = u + (size_t)b * n1 * n1, *vb = v + (size_t)b * n1 * n1, *pb = p + (size_t)b * m * m;
    if (e < HP * HP) {
        const int i = e / HP, k = e % HP;
        const bool in = (i < h && k < h);
        const int ip = m - 1 - i, kp = m - 1 - k;
#pragma unroll
        for (int f = 0; f < 3; ++f) {
            double q0 = 0, q1 = 0, q2 = 0, q3 = 0;   // F(i,k) F(i',k) F(i,k') F(i',k')
            if (in) {
                if (f < 2) {
                    const double* F = f ? vb : ub;
                    q0 = F[(i + 1) * n1 + k + 1]; q1 = F[(ip + 1) * n1 + k + 1];
                    q2 = F[(i + 1) * n1 + kp + 1]; q3 = F[(ip + 1) * n1 + kp + 1];
                } else {
                    q0 = pb[i * m + k]; q1 = pb[ip * m + k]; q2 = pb[i * m + kp]; q3 = pb[ip * m + kp];
                }
            }
            const double blk4[4] = {0.25 * ((q0 + q1) + (q2 + q3)), 0.25 * ((q0 + q1) - (q2 + q3)),
                                    0.25 * ((q0 - q1) + (q2 - q3)), 0.25 * ((q0 - q1) - (q2 - q3))};   // SS SA AS AA
            const int b0 = S2_BU + 4 * f;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                s[s2_off(i / 16, k / 4, b0 + q, i % 16, k % 4)] = blk4[q];
                s[S2_SIDE + s2_off(k / 16, i / 4, b0 + q, k % 16, i % 4)] = blk4[q];
            }
        }
    }
    double* ed = s + S2_EDGE_OFF;
    const int ME = S2_ME;
    if (e < ME) {
        const bool in = e < m;
        ed[e] = in ? ub[e + 1] : 0.0;                         // uW: row 0
        ed[ME + e] = in ? ub[N * n1 + e + 1] : 0.0;           // uE: row N
        ed[2 * ME + e] = in ? ub[(e + 1) * n1] : 0.0;         // uS: column 0
        ed[3 * ME + e] = in ? ub[(e + 1) * n1 + N] : 0.0;     // uN: column N (lid)
        ed[4 * ME + e] = in ? vb[e + 1] : 0.0;
        ed[5 * ME + e] = in ? vb[N * n1 + e + 1] : 0.0;
        ed[6 * ME + e] = in ? vb[(e + 1) * n1] : 0.0;
        ed[7 * ME + e] = in ? vb[(e + 1) * n1 + N] : 0.0;
    }
    if (e < 8) {
        const double* f = e < 4 ? ub : vb;
        const int c = e & 3;
        const int i = (c & 2) ? N : 0, j = (c & 1) ? N : 0;
        ed[8 * ME + e] = f[i * n1 + j];
    }
}

// record -> canonical (u, v, p); bufs = base of the three rotating copies, the cavity's current one is status.pad
__global__ void k_s2_unpack_state(double* __restrict__ u, double* __restrict__ v, double* __restrict__ p,
                                  const double* __restrict__ bufs, size_t buf_stride, const CavityStatus* __restrict__ status,
                                  int n1) {
    const int b = blockIdx.y, m = n1 - 2, N = n1 - 1, ME = S2_ME;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n1 * n1) return;
    const double* s = bufs + (size_t)status[b].pad * buf_stride + (size_t)b * S2_STRIDE;
    const double* ed = s + S2_EDGE_OFF;
    const int i = e / n1, j = e % n1;
    double uu, vv;
    const bool ci = (i == 0 || i == N), cj = (j == 0 || j == N);
    if (ci && cj) {
        const int c = (i == N ? 2 : 0) + (j == N ? 1 : 0);
        uu = ed[8 * ME + c];
        vv = ed[8 * ME + 4 + c];
    } else if (i == 0) {
        uu = ed[j - 1];
        vv = ed[4 * ME + j - 1];
    } else if (i == N) {
        uu = ed[ME + j - 1];
        vv = ed[5 * ME + j - 1];
    } else if (j == 0) {
        uu = ed[2 * ME + i - 1];
        vv = ed[6 * ME + i - 1];
    } else if (j == N) {
        uu = ed[3 * ME + i - 1];
        vv = ed[7 * ME + i - 1];
    } else {
        const int ii = i - 1, jj = j - 1, ip = m - 1 - ii, jp = m - 1 - jj;
        const int r = ii < ip ? ii : ip, c = jj < jp ? jj : jp;
        const double sa = ii > ip ? -1.0 : 1.0, sb = jj > jp ? -1.0 : 1.0;
        double f[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int b0 = S2_BU + 4 * q;
            const double ss = s[s2_off(r / 16, c / 4, b0 + 0, r % 16, c % 4)];
            const double sA = s[s2_off(r / 16, c / 4, b0 + 1, r % 16, c % 4)];
            const double as = s[s2_off(r / 16, c / 4, b0 + 2, r % 16, c % 4)];
            const double aa = s[s2_off(r / 16, c / 4, b0 + 3, r % 16, c % 4)];
            f[q] = (ss + sb * sA) + sa * (as + sb * aa);
        }
        uu = f[0];
        vv = f[1];
        p[(size_t)b * m * m + ii * m + jj] = f[2];
    }
    u[(size_t)b * n1 * n1 + e] = uu;
    v[(size_t)b * n1 * n1 + e] = vv;
}

// the cavity's operator blocks into both sides of its three records: D as is, DD scaled by nu = 1 / Re, G negated
__global__ void k_s2_pack_ops(const double* __restrict__ ops, const double* __restrict__ Re, double* __restrict__ rec,
                              size_t buf_stride) {
    const int b = blockIdx.y;
    const double nu = 1.0 / Re[b];
    double* dst = rec + (size_t)b * S2_STRIDE;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < 6 * S2_HP * S2_HP; e += gridDim.x * blockDim.x) {
        const int blk = e / (S2_HP * S2_HP), i = (e / S2_HP) % S2_HP, k = e % S2_HP;
        const double sc = blk < 2 ? 1.0 : (blk < 4 ? nu : -1.0);
        const double val = sc * ops[(size_t)blk * ST_BLK + i * ST_LD + k];
        const size_t o = s2_off(i / 16, k / 4, blk, i % 16, k % 4);
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            dst[q * buf_stride + o] = val;
            dst[q * buf_stride + S2_SIDE + o] = val;
        }
    }
}

// start of a solve: max|u|, max|v| of the record in buffer 0 into the status (the kernel forms dt from them)
__global__ void __launch_bounds__(256) k_s2_begin(const double* __restrict__ buf0, int m, CavityStatus* __restrict__ status) {
    __shared__ double sh[2 * 32];
    const int b = blockIdx.x, h = eo_half(m);
    const double* rec = buf0 + (size_t)b * S2_STRIDE;
    double mx[2] = {0.0, 0.0};
    bool isnan_[2] = {false, false};
    for (int e = threadIdx.x; e < h * h; e += blockDim.x) {
        const int i = e / h, k = e % h;
#pragma unroll
        for (int f = 0; f < 2; ++f) {
            double r[4];
            const size_t o = s2_off(i / 16, k / 4, S2_BU + 4 * f, i % 16, k % 4);
            eo_to_points(rec[o], rec[o + 64], rec[o + 128], rec[o + 192], r);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                isnan_[f] |= (r[q] != r[q]);
                mx[f] = fmax(mx[f], fabs(r[q]));
            }
        }
    }
    const double* ed = rec + S2_EDGE_OFF;
    for (int e = threadIdx.x; e < 4 * S2_ME + 4; e += blockDim.x) {
        const int q = e < 4 * S2_ME ? e : 8 * S2_ME + (e - 4 * S2_ME);
        const double a = ed[q], c = ed[q + (e < 4 * S2_ME ? 4 * S2_ME : 4)];
        isnan_[0] |= (a != a);
        isnan_[1] |= (c != c);
        mx[0] = fmax(mx[0], fabs(a));
        mx[1] = fmax(mx[1], fabs(c));
    }
    double flag[2] = {isnan_[0] ? 1.0 : 0.0, isnan_[1] ? 1.0 : 0.0};
    block_reduce<2, true>(mx, sh);
    block_reduce<2, true>(flag, sh);
    if (threadIdx.x == 0) {
        status[b].umax = flag[0] > 0.0 ? nan("") : mx[0];
        status[b].vmax = flag[1] > 0.0 ? nan("") : mx[1];
    }
}

// ---- the K loop of one chunk -------------------------------------------------------------------------------------------
// Contractions of the warp's field for the k4-steps [kbeg, kend) of one chunk slot (cf. st_chunk / eo_field_pass).
//   ro = (8 tw + g) * 4 + t : the thread's A-fragment offset inside a 16 x 4 block;  co = g * 4 + t : B-fragment offset
template <bool P_LEFT>
__device__ __forceinline__ void s2_chunk(const double* __restrict__ sl, int boff, int kbeg, int kend, int ro, int co, int fld,
                                         double (&a1)[4][2][2], double (&a2)[4][2][2], double (&aL)[4][2][2]) {
    constexpr int NT = 2;
    // every operand of a k4-step sits in one 18-block group of its side: [kk][blk][16][4]
    const double* sOa = sl + S2_SL_A + ro;
    const double* sOb = sl + boff + co;
    const double* sFa = sl + S2_SL_A + (S2_BU + 4 * fld) * 64 + ro;
    const double* sFb = sl + boff + (S2_BU + 4 * fld) * 64 + co;
    const double* sPa = sl + S2_SL_A + S2_BP * 64 + ro;
    const double* sPb = sl + boff + S2_BP * 64 + co;
#pragma unroll 1
    for (int kk = kbeg; kk < kend; ++kk) {
        double FA[4], FB[4][NT];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            FA[q] = sFa[(kk * S2_NB + q) * 64];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) FB[q][nt] = sFb[(kk * S2_NB + q) * 64 + nt * 32];
        }
        {
            double dA[2], ddA[2];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                dA[a] = sOa[(kk * S2_NB + a) * 64];
                ddA[a] = sOa[(kk * S2_NB + 2 + a) * 64];
            }
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        dmma884(a1[2 * a + b][nt][0], a1[2 * a + b][nt][1], dA[a], FB[2 * (1 - a) + b][nt]);
                        dmma884(aL[2 * a + b][nt][0], aL[2 * a + b][nt][1], ddA[a], FB[2 * a + b][nt]);
                    }
        }
        {
            double dB[2][NT], ddB[2][NT];
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    dB[b][nt] = sOb[(kk * S2_NB + b) * 64 + nt * 32];
                    ddB[b][nt] = sOb[(kk * S2_NB + 2 + b) * 64 + nt * 32];
                }
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        dmma884(a2[2 * a + b][nt][0], a2[2 * a + b][nt][1], FA[2 * a + (1 - b)], dB[b][nt]);
                        dmma884(aL[2 * a + b][nt][0], aL[2 * a + b][nt][1], FA[2 * a + b], ddB[b][nt]);
                    }
        }
        if (P_LEFT) {
            double gA[2], PB[4][NT];
#pragma unroll
            for (int a = 0; a < 2; ++a) gA[a] = sOa[(kk * S2_NB + 4 + a) * 64];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) PB[q][nt] = sPb[(kk * S2_NB + q) * 64 + nt * 32];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt)
                        dmma884(aL[2 * a + b][nt][0], aL[2 * a + b][nt][1], gA[a], PB[2 * (1 - a) + b][nt]);
        } else {
            double PA[4], gB[2][NT];
#pragma unroll
            for (int q = 0; q < 4; ++q) PA[q] = sPa[(kk * S2_NB + q) * 64];
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) gB[b][nt] = sOb[(kk * S2_NB + 4 + b) * 64 + nt * 32];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt)
                        dmma884(aL[2 * a + b][nt][0], aL[2 * a + b][nt][1], PA[2 * a + (1 - b)], gB[b][nt]);
        }
    }
}

// Boundary-line corrections + nonlinear assembly at the thread's two elements (columns j, j + 1) of every parity block
// (cf. st_assemble).  On return bL = momentum residual blocks, b1 = d(field)/dx (IS_U) or d(field)/dy (!IS_U) blocks.
template <bool IS_U>
__device__ __forceinline__ void s2_assemble(const double* __restrict__ sVec, const double* __restrict__ sEdgeF, double nu,
                                            bool bc_lines, int i, int j0, const double (&uo)[4][2], const double (&vo)[4][2],
                                            double (&b1)[4][2], double (&b2)[4][2], double (&bL)[4][2]) {
    constexpr int HP = S2_HP;
    double d0i[2], dNi[2], dd0i[2], ddNi[2], Si[2], Ni[2];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        d0i[a] = sVec[(0 + a) * HP + i];
        dNi[a] = sVec[(2 + a) * HP + i];
        dd0i[a] = sVec[(4 + a) * HP + i];
        ddNi[a] = sVec[(6 + a) * HP + i];
        Si[a] = sEdgeF[(4 + a) * HP + i];
        Ni[a] = sEdgeF[(6 + a) * HP + i];
    }
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
        const int j = j0 + hh;
        double dx[4], dy[4];
        if (bc_lines) {   // W = E = S = 0: only the lid line contributes, and only to the y-derivatives
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const double dNj = sVec[(2 + b) * HP + j], ddNj = sVec[(6 + b) * HP + j];
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    const int q = 2 * a + b;
                    dx[q] = b1[q][hh];
                    dy[q] = fma(Ni[a], dNj, b2[q][hh]);
                    bL[q][hh] = fma(nu * Ni[a], ddNj, bL[q][hh]);
                }
            }
        } else {
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const double Wj = sEdgeF[(0 + b) * HP + j], Ej = sEdgeF[(2 + b) * HP + j];
                const double d0j = sVec[(0 + b) * HP + j], dNj = sVec[(2 + b) * HP + j];
                const double dd0j = sVec[(4 + b) * HP + j], ddNj = sVec[(6 + b) * HP + j];
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    const int q = 2 * a + b;
                    dx[q] = b1[q][hh] + d0i[a] * Wj + dNi[a] * Ej;
                    dy[q] = b2[q][hh] + Si[a] * d0j + Ni[a] * dNj;
                    bL[q][hh] += nu * (dd0i[a] * Wj + ddNi[a] * Ej + Si[a] * dd0j + Ni[a] * ddNj);
                }
            }
        }
        // convective term in parity form: Z2 x Z2 group convolution (see eo_assemble)
        double cq[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            double c = uo[q][hh] * dx[0];
#pragma unroll
            for (int bb = 1; bb < 4; ++bb) c = fma(uo[q ^ bb][hh], dx[bb], c);
#pragma unroll
            for (int bb = 0; bb < 4; ++bb) c = fma(vo[q ^ bb][hh], dy[bb], c);
            cq[q] = c;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            bL[q][hh] -= cq[q];
            b1[q][hh] = IS_U ? dx[q] : dy[q];
        }
    }
}

// dt from the start-of-step maxima (spectral/sg.py:387-408).  Python's max(np.max(|u|), lid): a NaN first argument wins.
__device__ __forceinline__ double s2_dt(double umax, double vmax, double lidv, double nu, double b2, double cfl, double dx_min,
                                        double dy_min) {
    const double ua = (lidv > umax) ? lidv : umax, va = (1e-10 > vmax) ? 1e-10 : vmax;
    const double lam_x = (ua + sqrt(ua * ua + b2)) / dx_min + nu / (dx_min * dx_min);
    const double lam_y = (va + sqrt(va * va + b2)) / dy_min + nu / (dy_min * dy_min);
    return cfl / (lam_x + lam_y);
}

// per-visit constants of a cavity, shared by the roles of a CTA
struct S2Visit {
    double nu, beta2, cfl, lidv, tol;
    int first;   // the cavity has taken no step yet: its boundary lines are as stored, not yet the boundary conditions
};

// After the K loop: swap accumulator halves with the other k-group (X = this group: it keeps the 8-column half X of the
// tile), then the epilogue of that half.  Returns nothing; writes the stage output, and at the last stage the tile partials.
template <int X>
__device__ __forceinline__ void s2_finish(const S2Params& P, const S2Visit& V, double* __restrict__ exch_own,
                                          const double* __restrict__ exch_other, unsigned long long* empty_own,
                                          unsigned long long* empty_other,
                                          const double* __restrict__ sUV, const double* __restrict__ sPhi,
                                          const double* __restrict__ sVec,
                                          double* __restrict__ sEdgeP, double* __restrict__ sX, double* __restrict__ sRed,
                                          const double* __restrict__ recN, double* __restrict__ recO, int tile, int b,
                                          int r0, int c0, int stage, bool bc_lines, double dt, int fld, int tw, int lane,
                                          double (&a1)[4][2][2], double (&a2)[4][2][2], double (&aL)[4][2][2],
                                          long long* __restrict__ tdbg) {
    const int g = lane >> 2, t = lane & 3;
    const int tid128 = threadIdx.x & 127;
    const int m = P.m, h = eo_half(m), centre = (m & 1) ? h - 1 : -1;
    const int i = r0 + 8 * tw + g, j0 = c0 + 8 * X + 2 * t;
    const int rt = r0 >> 4, ct = c0 >> 4, ir = i & 15;
    const size_t offR = s2_off(rt, j0 >> 2, S2_BU + 4 * fld, ir, j0 & 3);    // + q * 64
    const size_t offPR = s2_off(rt, j0 >> 2, S2_BP, ir, j0 & 3);
    // ---- split-K hand-over: give away the other half, take the partner's contribution to this half ----
    consumer_sync();   // both groups are done reading their last chunk slots
    if (tdbg) tdbg[0] = clock64();
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int idx = q * 2 + hh;
            exch_own[(0 + idx) * 128 + tid128] = a1[q][1 - X][hh];
            exch_own[(8 + idx) * 128 + tid128] = a2[q][1 - X][hh];
            exch_own[(16 + idx) * 128 + tid128] = aL[q][1 - X][hh];
        }
    consumer_sync();
    if (tdbg) tdbg[1] = clock64();
    double b1[4][2], b2[4][2], bL[4][2];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int idx = q * 2 + hh;
            // group 0's sum first, whichever group does the addition (bitwise the same in both)
            const double o1 = exch_other[(0 + idx) * 128 + tid128], o2 = exch_other[(8 + idx) * 128 + tid128],
                         oL = exch_other[(16 + idx) * 128 + tid128];
            b1[q][hh] = X == 0 ? a1[q][X][hh] + o1 : o1 + a1[q][X][hh];
            b2[q][hh] = X == 0 ? a2[q][X][hh] + o2 : o2 + a2[q][X][hh];
            bL[q][hh] = X == 0 ? aL[q][X][hh] + oL : oL + aL[q][X][hh];
        }
    // the partner's slot has been read -- its generic writes are therefore performed --: hand both slots back to the
    // producers (each slot is refilled once all eight warps have let go of it)
    __syncwarp();
    if (lane == 0) {
        mbar_arrive(empty_own);
        mbar_arrive(empty_other);
    }

    // ---- stage input u, v at the thread's elements (picked out of the row panels during the K loop) ----
    double uo[4][2], vo[4][2];
    {
        const int pos = (X * 2 + tw) * 32 + lane;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            // stage 0: the stage input is phi^n itself (picked into sPhi during the K loop)
            const double* uvsrc = stage == 0 ? sPhi : sUV;
            const double2 a = *reinterpret_cast<const double2*>(uvsrc + ((0 + q) * 128 + pos) * 2);
            const double2 c = *reinterpret_cast<const double2*>(uvsrc + ((4 + q) * 128 + pos) * 2);
            uo[q][0] = a.x; uo[q][1] = a.y;
            vo[q][0] = c.x; vo[q][1] = c.y;
        }
    }
    if (fld == 0) s2_assemble<true>(sVec, sEdgeP, V.nu, bc_lines, i, j0, uo, vo, b1, b2, bL);
    else s2_assemble<false>(sVec, sEdgeP + 8 * S2_HP, V.nu, bc_lines, i, j0, uo, vo, b1, b2, bL);

    const double adt = rk_alpha(stage) * dt;
    const bool last = (stage == 3);
    double s0 = 0.0, s1 = 0.0, mx = 0.0;   // sum d^2, sum phi_n^2, max |phi_new| of this warp's field
    {
        const double wi = (i == centre) ? 0.5 : 1.0;
        const double w0 = 4.0 * wi * (j0 == centre ? 0.5 : 1.0), w1 = 4.0 * wi * (j0 + 1 == centre ? 0.5 : 1.0);
        double fn[4][2];
        double* outR = recO + offR;
        double* outC = recO + S2_SIDE + s2_off(ct, i >> 2, S2_BU + 4 * fld, j0 & 15, i & 3);   // + q * 64; second column + 4
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            // phi^n of the warp's field at the thread's elements: kept in shared memory since stage 0 of this step
            const double2 f0 = *reinterpret_cast<const double2*>(sPhi + ((4 * fld + q) * 128 + (X * 2 + tw) * 32 + lane) * 2);
            const double fx = f0.x, fy = f0.y;
            fn[q][0] = fx + adt * bL[q][0];
            fn[q][1] = fy + adt * bL[q][1];
            if (last) {
                double d;
                d = fn[q][0] - fx; s0 = fma(w0 * d, d, s0); s1 = fma(w0 * fx, fx, s1);
                d = fn[q][1] - fy; s0 = fma(w1 * d, d, s0); s1 = fma(w1 * fy, fy, s1);
            }
            __stcg(reinterpret_cast<double2*>(outR + q * 64), make_double2(fn[q][0], fn[q][1]));
            __stcg(outC + q * 64, fn[q][0]);
            __stcg(outC + q * 64 + 4, fn[q][1]);
            if (fld == 1) {
                sX[((X * 2 + tw) * 8 + q * 2 + 0) * 32 + lane] = b1[q][0];
                sX[((X * 2 + tw) * 8 + q * 2 + 1) * 32 + lane] = b1[q][1];
            }
        }
        if (last) {
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                double fr[4];
                eo_to_points(fn[0][hh], fn[1][hh], fn[2][hh], fn[3][hh], fr);
#pragma unroll
                for (int q = 0; q < 4; ++q) mx = nanmax(mx, fabs(fr[q]));
            }
        }
    }
    if (last) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, o);
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            mx = nanmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if (lane == 0) {
            const int w = threadIdx.x >> 5;
            sRed[w * 4 + 0] = s0;
            sRed[w * 4 + 1] = s1;
            sRed[w * 4 + 2] = mx;
        }
    }
    if (tdbg) tdbg[2] = clock64();
    consumer_sync();   // dv/dy of the v warps is in sX (and the warps' partial sums in sRed)
    if (fld == 0) {
        double* outR = recO + offPR;
        double* outC = recO + S2_SIDE + s2_off(ct, i >> 2, S2_BP, j0 & 15, i & 3);
        // frozen pressure (SGSolver): every stage takes grad(p^n), and a low-storage stage never reads the
        // previous stage's p, so stages 0..2 pass p^n on as the next stage's pressure operand and only the last stage
        // stores the new p; staged pressure (MultigridSmoother): every stage stores its own
        const bool keep_pn = P.freeze && stage != 3;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double dv0 = b1[q][0] + sX[((X * 2 + tw) * 8 + q * 2 + 0) * 32 + lane];
            const double dv1 = b1[q][1] + sX[((X * 2 + tw) * 8 + q * 2 + 1) * 32 + lane];
            const double2 p0 = *reinterpret_cast<const double2*>(sPhi + ((8 + q) * 128 + (X * 2 + tw) * 32 + lane) * 2);
            const double pn0 = keep_pn ? p0.x : p0.x + adt * (-V.beta2 * dv0);
            const double pn1 = keep_pn ? p0.y : p0.y + adt * (-V.beta2 * dv1);
            __stcg(reinterpret_cast<double2*>(outR + q * 64), make_double2(pn0, pn1));
            __stcg(outC + q * 64, pn0);
            __stcg(outC + q * 64 + 4, pn1);
        }
    }
    if (!bc_lines) {
        // first stage after a load done: from now on the boundary lines are the boundary conditions (walls zero, lid line)
        for (int e = threadIdx.x; e < 8 * S2_HP; e += S2_CONSUMERS) {
            const int line = e / S2_HP, k = e % S2_HP;
            double x = 0.0, y = 0.0;
            if (k < h && (line & 3) == 3) {
                const double* lid = (line < 4 ? P.ulid : P.vlid) + (size_t)b * P.n1 + 1;
                x = lid[k];
                y = lid[m - 1 - k];
            }
            sEdgeP[2 * line * S2_HP + k] = 0.5 * (x + y);
            sEdgeP[(2 * line + 1) * S2_HP + k] = 0.5 * (x - y);
        }
    }
    if (last && threadIdx.x == 0) {
        // warps: w = 4 X + 2 fld + tw.  u: warps 0, 1, 4, 5; v: warps 2, 3, 6, 7 -- always in this order
        double* out = P.partials + ((size_t)b * ST_TILES + tile) * 8;
        __stcg(out + 0, ((sRed[0 * 4 + 0] + sRed[1 * 4 + 0]) + sRed[4 * 4 + 0]) + sRed[5 * 4 + 0]);
        __stcg(out + 1, ((sRed[0 * 4 + 1] + sRed[1 * 4 + 1]) + sRed[4 * 4 + 1]) + sRed[5 * 4 + 1]);
        __stcg(out + 2, ((sRed[2 * 4 + 0] + sRed[3 * 4 + 0]) + sRed[6 * 4 + 0]) + sRed[7 * 4 + 0]);
        __stcg(out + 3, ((sRed[2 * 4 + 1] + sRed[3 * 4 + 1]) + sRed[6 * 4 + 1]) + sRed[7 * 4 + 1]);
        __stcg(out + 4, nanmax(nanmax(sRed[0 * 4 + 2], sRed[1 * 4 + 2]), nanmax(sRed[4 * 4 + 2], sRed[5 * 4 + 2])));
        __stcg(out + 5, nanmax(nanmax(sRed[2 * 4 + 2], sRed[3 * 4 + 2]), nanmax(sRed[6 * 4 + 2], sRed[7 * 4 + 2])));
    }
}

// 4 sums + 2 NaN-propagating maxima across the eight consumer warps, fixed order, result in every consumer thread.
__device__ __forceinline__ void s2_consumer_reduce(double (&s)[4], double (&mx)[2], double* sRed, int warp, int lane) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int q = 0; q < 4; ++q) s[q] += __shfl_xor_sync(0xffffffffu, s[q], o);
#pragma unroll
        for (int q = 0; q < 2; ++q) mx[q] = nanmax(mx[q], __shfl_xor_sync(0xffffffffu, mx[q], o));
    }
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) sRed[warp * 8 + q] = s[q];
        sRed[warp * 8 + 4] = mx[0];
        sRed[warp * 8 + 5] = mx[1];
    }
    consumer_sync();
#pragma unroll
    for (int q = 0; q < 4; ++q) s[q] = sRed[q];
    mx[0] = sRed[4];
    mx[1] = sRed[5];
    for (int w = 1; w < 8; ++w) {
#pragma unroll
        for (int q = 0; q < 4; ++q) s[q] += sRed[w * 8 + q];
        mx[0] = nanmax(mx[0], sRed[w * 8 + 4]);
        mx[1] = nanmax(mx[1], sRed[w * 8 + 5]);
    }
    consumer_sync();
}

// shared-memory carve-up of k_stream_tma
struct S2Smem {
    double *slot, *uv, *phin, *vec, *edgep, *x, *part, *red, *step;
    unsigned long long *full, *empty, *stepbar, *teambar;
    int* cavity;
    __device__ explicit S2Smem(double* smem) {
        slot = smem;
        uv = slot + (size_t)S2_SLOTS * S2_SLOT;
        phin = uv + 8 * 256;
        vec = phin + 12 * 256;
        edgep = vec + 8 * S2_HP;
        x = edgep + 16 * S2_HP;
        part = x + 4 * 8 * 32;             // [16 tiles][8]
        red = part + 16 * 8;               // 80
        step = red + 80;                   // dt, active | boundary-line sums: first step 4, max|lid| 2, later steps 4 | cavity
        full = reinterpret_cast<unsigned long long*>(step + 16);
        empty = full + S2_SLOTS;
        stepbar = empty + S2_SLOTS;
        teambar = stepbar + 1;
        cavity = reinterpret_cast<int*>(step + 15);
    }
};

// CTA geometry inside its team
struct S2Geo {
    int ntile, ntc, team_size, team, tin, trow, tcol, tile, r0, c0, m, h, n1, nk4, nch;
    // cw = 16-column tiles per CTA: 1 (k_stream_tma) or 2 (k_stream_tma_wide); tcol = the CTA's first column tile
    __device__ S2Geo(const S2Params& P, int cw) {
        ntile = P.ntile; ntc = (ntile + cw - 1) / cw; team_size = ntile * ntc;
        team = blockIdx.x / team_size; tin = blockIdx.x % team_size;
        trow = tin / ntc; tcol = (tin % ntc) * cw; tile = trow * (S2_HP / 16) + tin % ntc;
        r0 = trow * 16; c0 = tcol * 16;
        m = P.m; h = eo_half(m); n1 = P.n1;
        nk4 = P.K4 / 4; nch = (nk4 + S2_KC4 - 1) / S2_KC4;
    }
};

// named barriers: 1 = the eight consumer warps (consumer_sync); 2 = consumers + the four producer warps, start of a visit;
// 3 = consumers arrive / producer waits: the visit's boundary-line sums are in shared memory
__device__ __forceinline__ void visit_sync() { asm volatile("bar.sync 2, 384;" ::: "memory"); }
__device__ __forceinline__ void sums_arrive() { asm volatile("bar.arrive 3, 288;" ::: "memory"); }
__device__ __forceinline__ void sums_wait() { asm volatile("bar.sync 3, 288;" ::: "memory"); }

// Start of a visit (consumer warps): the boundary lines in parity form for the first stage, and the boundary-line parts
// of the convergence sums into S.step for the producer warp's reductions.
template <class SM>
__device__ __forceinline__ void s2_visit_prepare(const S2Params& P, const SM& S, const S2Visit& V, const double* __restrict__ rec0,
                                                 int cur, const double* __restrict__ ul, const double* __restrict__ vl, int m, int h,
                                                 int n1) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
        const double* ed = rec0 + (size_t)cur * P.buf_stride + S2_EDGE_OFF;
        for (int e = tid; e < 8 * S2_HP; e += S2_CONSUMERS) {
            const int line = e / S2_HP, k = e % S2_HP;
            double x = 0.0, y = 0.0;
            if (k < h) {
                if (V.first) {
                    x = __ldcg(ed + line * S2_ME + k);
                    y = __ldcg(ed + line * S2_ME + m - 1 - k);
                } else if ((line & 3) == 3) {
                    const double* lid = (line < 4 ? ul : vl) + 1;
                    x = lid[k];
                    y = lid[m - 1 - k];
                }
            }
            S.edgep[2 * line * S2_HP + k] = 0.5 * (x + y);
            S.edgep[(2 * line + 1) * S2_HP + k] = 0.5 * (x - y);
        }
        // sums over the boundary lines of ||phi^{n+1} - phi^n||^2, ||phi^n||^2 and max |lid|: with the lines as stored
        // (first step after a load) and with the boundary-condition lines (every later step)
        double sf[4] = {0, 0, 0, 0}, sb[4] = {0, 0, 0, 0}, bm[2] = {0, 0}, dummy[2] = {0, 0};
        for (int e = tid; e < m; e += S2_CONSUMERS) {
#pragma unroll
            for (int f = 0; f < 2; ++f) {
                const double lid = (f ? vl : ul)[e + 1];
                const double* L = ed + f * 4 * S2_ME;
                const double w = __ldcg(L + e), ea = __ldcg(L + S2_ME + e), so = __ldcg(L + 2 * S2_ME + e),
                             no = __ldcg(L + 3 * S2_ME + e);
                const double dn = lid - no;
                sf[2 * f] += w * w + ea * ea + so * so + dn * dn;
                sf[2 * f + 1] += w * w + ea * ea + so * so + no * no;
                sb[2 * f + 1] += lid * lid;
                bm[f] = nanmax(bm[f], fabs(lid));
            }
        }
        if (tid == 0) {
#pragma unroll
            for (int f = 0; f < 2; ++f) {
                const double* c = ed + 8 * S2_ME + 4 * f;   // c00 c0N cN0 cNN
                const double l0 = (f ? vl : ul)[0], lN = (f ? vl : ul)[n1 - 1];
                const double q0 = __ldcg(c), q1 = __ldcg(c + 1), q2 = __ldcg(c + 2), q3 = __ldcg(c + 3);
                const double d1 = l0 - q1, d3 = lN - q3;
                sf[2 * f] += q0 * q0 + d1 * d1 + q2 * q2 + d3 * d3;
                sf[2 * f + 1] += q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3;
                sb[2 * f + 1] += l0 * l0 + lN * lN;
                bm[f] = nanmax(nanmax(bm[f], fabs(l0)), fabs(lN));
            }
        }
        s2_consumer_reduce(sf, bm, S.red, warp, lane);
        s2_consumer_reduce(sb, dummy, S.red, warp, lane);
        if (tid == 0) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                S.step[2 + q] = sf[q];
                S.step[8 + q] = sb[q];
            }
            S.step[6] = bm[0];
            S.step[7] = bm[1];
        }
        __threadfence_block();
        sums_arrive();     // the producer warp reads them at its first reduction of the visit
        consumer_sync();   // S.edgep complete
}

// ================================================ consumer warps ================================================
__device__ __forceinline__ void s2_consumer_main(const S2Params& P, const S2Smem& S, const S2Geo& G) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int X = warp >> 2, fld = (warp >> 1) & 1, tw = warp & 1, g = lane >> 2, t = lane & 3;
    const int ro = (8 * tw + g) * 4 + t, co = g * 4 + t;
    const int m = G.m, h = G.h, n1 = G.n1, nch = G.nch;
    const int cgrab = (G.c0 >> 3) + X;                       // the chunk holding this group's 8 columns of the tile
    // Split K: both k-groups visit every chunk, group X takes the chunk's k4-step X.  The slots of the last two chunks
    // are kept after the loop: the groups swap accumulator halves through them (s2_finish).
    unsigned gcb = 0;        // chunks consumed before the current stage (ring position)
    unsigned nstepinfo = 0;  // step-info hand-overs consumed so far
    const bool dbg_on = P.dbg != nullptr && blockIdx.x == 0 && (tid & 127) == 0;   // warp 0 (group 0) and warp 4 (group 1)
    int dbg_stage = 0;

    for (;;) {
        visit_sync();   // the producer warp has the team's next cavity
        const int b = *S.cavity;
        if (b < 0) return;
        const CavityStatus st = load_status(P.status + b);
        int cur = __ldcg(&P.status[b].pad);
        S2Visit V;
        V.nu = 1.0 / P.Re[b]; V.beta2 = P.beta2[b]; V.cfl = P.CFL[b]; V.lidv = P.lidvel[b];
        V.tol = 0.0;
        V.first = st.iters == 0 ? 1 : 0;
        long long nsteps = P.launch_end - st.iters;
        if (nsteps > P.chunk) nsteps = P.chunk;
        const double *ul = P.ulid + (size_t)b * n1, *vl = P.vlid + (size_t)b * n1;
        double* rec0 = P.rec + (size_t)b * S2_STRIDE;

        s2_visit_prepare(P, S, V, rec0, cur, ul, vl, m, h, n1);
        // dt of the first step of the visit from the status (spectral/sg.py:387-408)
        double dt = s2_dt(st.umax, st.vmax, V.lidv, V.nu, V.beta2, V.cfl, P.dx_min, P.dy_min);
        bool bc_lines = !V.first;
        bool aborted = false;
        for (long long step = 0; step < nsteps && !aborted; ++step) {
            for (int stage = 0; stage < 4; ++stage) {
                const double* recN = rec0 + (size_t)cur * P.buf_stride;
                double* recO = rec0 + (size_t)st_out(cur, stage) * P.buf_stride;
                double a1[4][2][2], a2[4][2][2], aL[4][2][2];
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) a1[q][nt][hh] = a2[q][nt][hh] = aL[q][nt][hh] = 0.0;
                // Stage 0 picks phi^n (the warp's field; the u warps also p^n) out of the passing row panels into S.phin,
                // where it stays for the four stages; stages 1..3 pick the stage input of the warp's field into S.uv.
                double* grab_dst = stage == 0 ? S.phin : S.uv;
                if (cgrab >= nch) {   // the group's columns lie in the zero padding
                    const int pos = (X * 2 + tw) * 32 + lane;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        *reinterpret_cast<double2*>(grab_dst + ((4 * fld + q) * 128 + pos) * 2) = make_double2(0.0, 0.0);
                        if (stage == 0 && fld == 0)
                            *reinterpret_cast<double2*>(S.phin + ((8 + q) * 128 + pos) * 2) = make_double2(0.0, 0.0);
                    }
                }
                long long t_start = 0, t_wait = 0;
                if (dbg_on) t_start = clock64();
                for (int c = 0; c < nch; ++c) {
                    const unsigned gc = gcb + c, slot = gc % S2_SLOTS;
                    long long tw0 = 0;
                    if (dbg_on) tw0 = clock64();
                    mbar_wait(S.full + slot, (gc / S2_SLOTS) & 1);
                    if (dbg_on) {
                        const long long tw1 = clock64();
                        t_wait += tw1 - tw0;
                        if (c == 0 && dbg_stage < S2_DBG_STAGES) P.dbg[dbg_stage * 16 + X * 4 + 1] = tw1 - t_start;
                    }
                    const double* sl = S.slot + (size_t)slot * S2_SLOT;
                    if (c == cgrab) {
                        // the thread's own elements of its field's stage input: columns 8 X + 2 t, + 1 of the tile
                        const int pos = (X * 2 + tw) * 32 + lane;
                        const double* src = sl + S2_SL_A + ((t >> 1) * S2_NB + S2_BU + 4 * fld) * 64 + (8 * tw + g) * 4 + 2 * (t & 1);
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            *reinterpret_cast<double2*>(grab_dst + ((4 * fld + q) * 128 + pos) * 2) =
                                *reinterpret_cast<const double2*>(src + q * 64);
                            if (stage == 0 && fld == 0)   // p^n: the pressure blocks sit (S2_BP - S2_BU) blocks further on
                                *reinterpret_cast<double2*>(S.phin + ((8 + q) * 128 + pos) * 2) =
                                    *reinterpret_cast<const double2*>(src + (S2_BP - S2_BU + q) * 64);
                        }
                    }
                    const int ksteps = min(S2_KC4, G.nk4 - c * S2_KC4);
                    const int kend = ksteps < X + 1 ? ksteps : X + 1;
                    if (fld == 0) s2_chunk<true>(sl, S2_SL_B, X, kend, ro, co, fld, a1, a2, aL);
                    else s2_chunk<false>(sl, S2_SL_B, X, kend, ro, co, fld, a1, a2, aL);
                    if (c < nch - 2) {
                        __syncwarp();
                        if (lane == 0) mbar_arrive(S.empty + slot);
                    }
                }
                const unsigned slot_own = (gcb + nch - 2 + X) % S2_SLOTS, slot_other = (gcb + nch - 1 - X) % S2_SLOTS;
                gcb += nch;
                long long t_loop = 0;
                if (dbg_on) t_loop = clock64();
                if (stage == 0 && step > 0) {
                    // the reduction of the previous step (producer warp): dt of this step, and whether there is one
                    mbar_wait(S.stepbar, nstepinfo & 1);
                    ++nstepinfo;
                    dt = S.step[0];
                    if (S.step[1] == 0.0) {
                        // the cavity went inactive with the previous step: this stage was speculative
                        __syncwarp();
                        if (lane == 0) {
                            mbar_arrive(S.empty + slot_own);
                            mbar_arrive(S.empty + slot_other);
                        }
                        aborted = true;
                        break;
                    }
                }
                long long tfin[4] = {0, 0, 0, 0};
                double* exch_own = S.slot + (size_t)slot_own * S2_SLOT;
                const double* exch_other = S.slot + (size_t)slot_other * S2_SLOT;
                if (X == 0)
                    s2_finish<0>(P, V, exch_own, exch_other, S.empty + slot_own, S.empty + slot_other, S.uv, S.phin, S.vec, S.edgep, S.x, S.red, recN, recO,
                                 G.tile, b, G.r0, G.c0, stage, bc_lines, dt, fld, tw, lane, a1, a2, aL,
                                 (dbg_on && dbg_stage < S2_DBG_STAGES) ? tfin : nullptr);
                else
                    s2_finish<1>(P, V, exch_own, exch_other, S.empty + slot_own, S.empty + slot_other, S.uv, S.phin, S.vec, S.edgep, S.x, S.red, recN, recO,
                                 G.tile, b, G.r0, G.c0, stage, bc_lines, dt, fld, tw, lane, a1, a2, aL, nullptr);
                bc_lines = true;
                long long t_fin = 0;
                if (dbg_on) t_fin = clock64();
                // ---- arrive at the team's stage barrier: this CTA's stage output is complete.  The release of the counter
                // update (cumulative over the CTA barrier) publishes every consumer's stores; the readers are TMA copies,
                // whose issuing thread crosses to the async proxy after its acquire (s2_producer_main) ----
                consumer_sync();
                if (tid == 0) red_release_add(P.team_bar + G.team, 1u);
                if (dbg_on && dbg_stage < S2_DBG_STAGES) {
                    // per group X: [0] stage start (absolute), [1] wait for the first chunk, [2] K loop incl. waits,
                    // [3] all waits inside the K loop; group 0 also: [8] exchange + epilogue, [9] arrive,
                    // [13] wait for the other group, [14] accumulator swap, [15] assembly + momentum stores
                    long long* d = P.dbg + dbg_stage * 16 + X * 4;
                    d[0] = t_start;
                    d[2] = t_loop - t_start;
                    d[3] = t_wait;
                    if (X == 0) {
                        P.dbg[dbg_stage * 16 + 8] = t_fin - t_loop;
                        P.dbg[dbg_stage * 16 + 9] = clock64() - t_fin;
                        P.dbg[dbg_stage * 16 + 13] = tfin[0] - t_loop;
                        P.dbg[dbg_stage * 16 + 14] = tfin[1] - tfin[0];
                        P.dbg[dbg_stage * 16 + 15] = tfin[2] - tfin[1];
                    }
                }
                ++dbg_stage;
                if (stage == 3) cur = (cur + 2) % 3;
            }
        }
    }
}

// ======================================= producer warps (8: leader, 9..11: helpers) =======================================
// One thread's bulk copies are served one after the other, ~360 cycles each whatever their size, while copies issued by
// different warps proceed side by side (scripts/tma_bw.cu, profiles/r02_tma_bw.txt).  So the four warps of the producer
// warpgroup each issue ONE side of every other chunk: warp 8 + pw issues side (pw & 1) of the chunks c = (pw >> 1) mod 2.
// Warp 8 is also the team's link to the cavity queue and the reducer of the end-of-step sums.
// WIDE (k_stream_tma_wide: a chunk has three sides A, B0, B1): warps 8, 9, 10 issue one side of EVERY chunk each; warp 11
// stands in for warp 8 on side A in the stages the leader spends reducing.
template <bool WIDE, class SM>
__device__ __forceinline__ void s2_producer_main(const S2Params& P, const SM& S, const S2Geo& G, int pw) {
    constexpr int SLOTS = WIDE ? 3 : S2_SLOTS;
    constexpr unsigned SLOT_BYTES = (WIDE ? 3 : 2) * S2_SIDE_BYTES;
    const int lane = threadIdx.x & 31;
    const bool leader = pw == 0;
    const int side = WIDE ? (pw == 3 ? 0 : pw) : (pw & 1), cpar = WIDE ? 0 : (pw >> 1), cstep = WIDE ? 1 : 2;
    const int m = G.m, n1 = G.n1, nch = G.nch, team_size = G.team_size;
    unsigned gcb = 0;        // chunks issued before the current stage (ring position)
    unsigned nbar = 0;       // team barriers completed before the current stage
    unsigned nstepinfo = 0;  // step-info hand-overs seen so far (helpers)
    unsigned nteam = 0;      // team-barrier hand-overs seen so far (helpers)
    unsigned visit = 0;
    const bool dbg_on = P.dbg != nullptr && blockIdx.x == 0 && lane == 0 && leader;
    int dbg_stage = 0;
    const unsigned slot0 = s2_smem(S.slot), full0 = s2_smem(S.full), empty0 = s2_smem(S.empty);
    // side s of chunk c of a record: c chunk-strides after the tile's first k4-step (s2_off)
    const size_t offA = (size_t)G.trow * (S2_NK4 * S2_NB * 64), offB = S2_SIDE + (size_t)G.tcol * (S2_NK4 * S2_NB * 64);
    const size_t offB1 = offB + (size_t)(S2_NK4 * S2_NB * 64);   // the second column tile (WIDE)
    const unsigned teambar = s2_smem(S.teambar);

    for (;;) {
        // ---- the team's next cavity ----
        if (leader) {
            int bq = 0;
            if (lane == 0) {
                if (G.tin == 0) {
                    bq = CavityQueue::pop(P);
                    st_release_u64(P.team_mail + G.team, ((unsigned long long)(visit + 1) << 32) | (unsigned)(bq + 1));
                } else {
                    unsigned long long mail;
                    while (((mail = ld_acquire_u64(P.team_mail + G.team)) >> 32) != visit + 1) __nanosleep(100);
                    bq = (int)(unsigned)(mail & 0xffffffffu) - 1;
                }
                *S.cavity = bq;
                __threadfence_block();
            }
        }
        visit_sync();
        const int b = *S.cavity;
        if (b < 0) return;
        ++visit;
        fence_proxy_async();   // the record was written through the generic proxy (another team, or the pack kernel)

        CavityStatus st = load_status(P.status + b);
        int cur = __ldcg(&P.status[b].pad);
        const double nu = 1.0 / P.Re[b], beta2 = P.beta2[b], cfl = P.CFL[b], lidv = P.lidvel[b];
        const double tol = P.tol ? P.tol[b] : -1.0;
        const bool first = st.iters == 0;
        long long nsteps = P.launch_end - st.iters;
        if (nsteps > P.chunk) nsteps = P.chunk;
        const double *ul = P.ulid + (size_t)b * n1, *vl = P.vlid + (size_t)b * n1;
        double* rec0 = P.rec + (size_t)b * S2_STRIDE;
        bool active = true, sums_ready = false;

        // end-of-step reduction (leader): tile partials in fixed order + boundary lines, convergence bookkeeping
        auto reduce_step = [&](bool first_step) {
            if (!sums_ready) {
                sums_wait();   // the consumers' boundary-line sums of this visit
                sums_ready = true;
            }
            if (lane < team_size) {
                const int tl = (lane / G.ntc) * (S2_HP / 16) + lane % G.ntc;
                const double* part = P.partials + ((size_t)b * ST_TILES + tl) * 8;
#pragma unroll
                for (int q = 0; q < 6; ++q) S.part[lane * 8 + q] = __ldcg(part + q);
            }
            __syncwarp();
            if (lane == 0) {
                const double* es = S.step + (first_step ? 2 : 8);
                double s[4] = {es[0], es[1], es[2], es[3]}, mx[2] = {0, 0};
                for (int tl = 0; tl < team_size; ++tl) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) s[q] += S.part[tl * 8 + q];
                    mx[0] = nanmax(mx[0], S.part[tl * 8 + 4]);
                    mx[1] = nanmax(mx[1], S.part[tl * 8 + 5]);
                }
                const double ures = sqrt(s[0]) / (sqrt(s[1]) + 1e-12), vres = sqrt(s[2]) / (sqrt(s[3]) + 1e-12);
                update_status(st, ures, vres, tol, P.warmup, P.max_iter, P.nan_exit);
                st.umax = nanmax(mx[0], S.step[6]);
                st.vmax = nanmax(mx[1], S.step[7]);
            }
            __syncwarp();
        };
        auto wait_team = [&]() {
            const unsigned target = nbar * (unsigned)team_size;
            // relaxed polls (an acquire load invalidates L1 on every turn), one acquire fence once the count is there
            while ((int)(ld_relaxed_u32(P.team_bar + G.team) - target) < 0) {}
            fence_acq_rel_gpu();
        };

        for (long long step = 0; step < nsteps && active; ++step) {
            for (int stage = 0; stage < 4; ++stage) {
                // the stage input: its pressure blocks are p^n in every stage when the pressure is frozen (s2_finish)
                const double* rec_in = rec0 + (size_t)st_in(cur, stage) * P.buf_stride;
                const bool need_bar = !(step == 0 && stage == 0);
                // In a stage that starts with the previous step's reduction the leader warp only reduces: warp 9 then
                // issues both sides of the even chunks.
                const bool reduce_now = stage == 0 && step > 0;
                const int s_lo = (!WIDE && reduce_now && pw == 1) ? 0 : side, s_hi = (!WIDE && reduce_now && pw == 1) ? 2 : side + 1;
                const bool idle = WIDE ? (pw == 3 ? !reduce_now : (reduce_now && leader)) : (reduce_now && leader);
                if (lane == 0) {
                    long long tp0 = 0, tp1 = 0;
                    if (dbg_on) tp0 = clock64();
                    if (need_bar) {
                        // the leader polls the team's counter in global memory and passes the news on in shared memory
                        if (leader) {
                            wait_team();
                            mbar_arrive(S.teambar);
                        } else {
                            mbar_wait_a(teambar, nteam & 1);
                            ++nteam;
                        }
                        fence_proxy_async();
                    }
                    if (dbg_on) tp1 = clock64();
                    if (!idle) {
                        for (int c = cpar; c < nch; c += cstep) {
                            const unsigned gc = gcb + c, slot = gc % SLOTS, fb = full0 + 8 * slot;
                            mbar_wait_a(empty0 + 8 * slot, ((gc / SLOTS) & 1) ^ 1);
                            for (int sd = s_lo; sd < s_hi; ++sd) {
                                mbar_expect_tx_a(fb, S2_SIDE_BYTES);
                                bulk_g2s_a(slot0 + slot * SLOT_BYTES + sd * S2_SIDE_BYTES,
                                           rec_in + (sd == 0 ? offA : (sd == 1 ? offB : offB1)) + (size_t)c * (S2_KC4 * S2_NB * 64),
                                           S2_SIDE_BYTES, fb);
                            }
                            if (dbg_on && c == cpar && dbg_stage < S2_DBG_STAGES) {
                                // [10] start of the wait (absolute), [11] wait at the team barrier, [12] first chunk issued
                                P.dbg[dbg_stage * 16 + 10] = tp0;
                                P.dbg[dbg_stage * 16 + 11] = tp1 - tp0;
                                P.dbg[dbg_stage * 16 + 12] = clock64() - tp1;
                            }
                        }
                    }
                    ++dbg_stage;
                }
                __syncwarp();
                gcb += nch;
                if (reduce_now && leader) {
                    // the previous step's sums (needed by the consumers only at this stage's epilogue)
                    reduce_step(first && step == 1);
                    if (lane == 0) {
                        active = st.active != 0;
                        S.step[0] = s2_dt(st.umax, st.vmax, lidv, nu, beta2, cfl, P.dx_min, P.dy_min);
                        S.step[1] = active ? 1.0 : 0.0;
                        __threadfence_block();
                        mbar_arrive(S.stepbar);
                    }
                    active = __shfl_sync(0xffffffffu, (int)active, 0) != 0;
                }
                if (reduce_now && !leader) {
                    // the leader's verdict on the previous step: is this stage a real one or was it speculative?
                    mbar_wait(S.stepbar, nstepinfo & 1);
                    ++nstepinfo;
                    active = S.step[1] != 0.0;
                }
                if (!active) break;   // the consumers drop this (speculative) stage as well: no barrier follows it
                ++nbar;
                if (stage == 3) cur = (cur + 2) % 3;
            }
        }
        if (!leader) continue;
        if (active) {
            // the last step of the visit: wait for the team, then its reduction
            if (lane == 0) wait_team();
            __syncwarp();
            reduce_step(first && nsteps == 1);
        }
        if (!sums_ready) sums_wait();   // (keeps the arrive / wait pairs of barrier 3 matched)
        // ---- visit end (leader CTA): boundary-condition lines with the new record, status, hand the cavity back ----
        if (G.tin == 0) {
            double* edn = rec0 + (size_t)cur * P.buf_stride + S2_EDGE_OFF;
            for (int e = lane; e < S2_ME; e += 32) {
#pragma unroll
                for (int f = 0; f < 2; ++f) {
                    double* Ln = edn + f * 4 * S2_ME;
                    __stcg(Ln + e, 0.0);
                    __stcg(Ln + S2_ME + e, 0.0);
                    __stcg(Ln + 2 * S2_ME + e, 0.0);
                    __stcg(Ln + 3 * S2_ME + e, e < m ? (f ? vl : ul)[e + 1] : 0.0);
                }
            }
            if (lane < 2) {
                double* cn = edn + 8 * S2_ME + 4 * lane;
                __stcg(cn + 0, 0.0);
                __stcg(cn + 1, (lane ? vl : ul)[0]);
                __stcg(cn + 2, 0.0);
                __stcg(cn + 3, (lane ? vl : ul)[n1 - 1]);
            }
            __syncwarp();
            if (lane == 0) {
                store_status(P.status + b, st);
                __stcg(&P.status[b].pad, cur);
                __threadfence();
                if (st.active && st.iters < P.launch_end) CavityQueue::push(P, b);
                else CavityQueue::retire(P);
            }
        }
    }
}

// The persistent kernel.  grid = teams x (tiles per cavity), cooperative launch (every CTA co-resident: the team barriers
// and the queue spin).  384 threads: warps 0..7 consume (224 registers each after setmaxnreg), warps 8..11 issue the bulk
// copies (56 registers; setmaxnreg works on whole warpgroups), warp 8 also polls the team barrier and reduces.
__global__ void __launch_bounds__(S2_THREADS, 1) k_stream_tma(const S2Params P) {
    extern __shared__ __align__(16) double smem[];
    const S2Smem S(smem);
    const S2Geo G(P, 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        for (int s = 0; s < S2_SLOTS; ++s) {
            mbar_init(S.full + s, 2);   // one arrival (with its byte count) per side
            mbar_init(S.empty + s, 8);   // the eight consumer warps
        }
        mbar_init(S.stepbar, 1);
        mbar_init(S.teambar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int e = tid; e < 8 * S2_HP; e += S2_THREADS) S.vec[e] = __ldg(P.ops + 6 * ST_BLK + e);
    __syncthreads();
    if (warp >= 8) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
        s2_producer_main<false>(P, S, G, warp - 8);
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
        s2_consumer_main(P, S, G);
    }
}

// After a run: move every cavity's record into buffer 0 so that the next run (or a continued solve) finds it there.
__global__ void k_s2_normalise(double* __restrict__ bufs, size_t buf_stride, CavityStatus* __restrict__ status) {
    const int b = blockIdx.y, cur = status[b].pad;
    if (cur == 0) return;
    const double2* src = reinterpret_cast<const double2*>(bufs + cur * buf_stride + (size_t)b * S2_STRIDE);
    double2* dst = reinterpret_cast<double2*>(bufs + (size_t)b * S2_STRIDE);
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < S2_STRIDE / 2; e += (size_t)gridDim.x * blockDim.x)
        dst[e] = src[e];
}

}  // namespace ldc
