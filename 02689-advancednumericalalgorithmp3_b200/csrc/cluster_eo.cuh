// Cluster-resident even-odd path for few cavities on the largest resident grids (50 <= N <= 65, parity blocks of 32 x 32).
//
// A cavity is spread over a thread-block CLUSTER of 8 CTAs (4 row tiles x 2 column tiles; a CTA owns an 8 x 16 tile of all
// four parity blocks of u, v and p).  Nothing of the stepping path goes through global memory:
//   * every CTA keeps the operand panels of its tile in its own shared memory for the whole visit -- the row panel
//     (operator rows and state rows of its 8 lines, all k) and the column panel (its 16 columns, all k), in the
//     k4-interleaved block layout of stream_tma.cuh, so that the DMMA fragment loads are the same conflict-free 256-byte reads;
//   * the output of an RK stage goes straight into the panels of the CTAs that need it -- the row panels of the two CTAs
//     of the tile's row, the column panels of the four CTAs of its column: the CTA's own by plain stores, the others'
//     through distributed shared memory by st.async stores (SASS STAS) that count their bytes on an mbarrier of the
//     RECEIVING CTA, which waits for its byte count instead of for a barrier round trip.  The one hardware cluster barrier
//     of a stage (barrier.cluster arrive.relaxed / wait) says that every CTA has finished reading its panels;
//   * the K loop of a stage runs in two parts: the half-steps on state the CTA produced itself right after its own stores,
//     with the remote stores issued between them, the others once the mbarrier says that everything has arrived;
//   * the end-of-step sums (convergence norms, maxima for the next dt) travel the same way ahead of the last stage's output
//     and are reduced, in fixed order, by warps 0 and 1 of every CTA while the other warps send; status and state go back to the
//     record in L2 only at the end of a visit.
// What bounds it: the SM-to-SM network takes ~20 B/clk out of an SM (scripts/dsmem_bw.cu, profiles/r02_dsmem_bw.txt), and
// a stage sends 48 KB per CTA; the K loop of a stage is 80 DMMAs per warp, two warps per SM sub-partition.
// Eight warps per CTA: (k-group X: even | odd k4-steps) x (u | v equation) x (8-column half of the tile); the two k-groups
// swap accumulator halves -- each thread keeps ONE of its two columns -- and both run the epilogue, one element per thread.
// Records, pack / unpack kernels, queue and host-side driver are those of stream_tma.cuh (ldc_abi.cu picks this kernel as a
// third "shape" of the persistent streamed path when every cavity still iterating can have a cluster).
#pragma once
#include "stream_tma.cuh"

namespace ldc {

constexpr int CE_NK4 = 8;                        // k4-steps of a 32-wide parity block
constexpr int CE_TR = 8, CE_TC = 16;             // tile of a CTA: rows x columns
constexpr int CE_CR = 32 / CE_TR, CE_CC = 32 / CE_TC;
constexpr int CE_CLUSTER = CE_CR * CE_CC;        // 8 CTAs per cavity; rank = row tile * CE_CC + column tile
constexpr int CE_THREADS = 256;
constexpr int CE_ABLK = CE_TR * 4, CE_BBLK = CE_TC * 4;      // doubles per block of the row / column panel
constexpr int CE_RP = CE_NK4 * S2_NB * CE_ABLK;              // row panel    [kk][blk][8 r][4]
constexpr int CE_CP = CE_NK4 * S2_NB * CE_BBLK;              // column panel [kk][blk][16 c][4]
constexpr int CE_EXCH = 2 * 12 * 128;                        // accumulator swap: [group][12 values][128 threads]
constexpr size_t CE_SMEM_DOUBLES = (size_t)CE_RP + CE_CP + CE_EXCH
                                   + 8 * S2_HP + 16 * S2_HP  // boundary-column vectors, boundary lines in parity form
                                   + 4 * 4 * 32              // dv/dy hand-over v warps -> u warps
                                   + CE_CLUSTER * 8 + 64 + 16 + 3;   // partials of the 8 tiles, reduction scratch, step info, cavity + mbarriers

struct CESmem {
    double *rp, *cp, *exch, *vec, *edgep, *x, *part, *red, *step;
    int* cavity;
    unsigned long long* databar;   // mbarrier: the other CTAs' part of a stage output has arrived (byte count)
    unsigned long long* partbar;   // mbarrier: the other CTAs' end-of-step sums have arrived
    __device__ explicit CESmem(double* smem) {
        rp = smem;
        cp = rp + CE_RP;
        exch = cp + CE_CP;
        vec = exch + CE_EXCH;
        edgep = vec + 8 * S2_HP;
        x = edgep + 16 * S2_HP;
        part = x + 4 * 4 * 32;
        red = part + CE_CLUSTER * 8;
        step = red + 64;   // [2..5] boundary-line sums of a first step, [6..7] max |lid|, [8..11] sums of later steps
        cavity = reinterpret_cast<int*>(step + 16);
        databar = reinterpret_cast<unsigned long long*>(step + 17);
        partbar = databar + 1;
    }
};

// ---- cluster PTX ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned ce_rank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
// shared::cluster address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ unsigned ce_mapa(const void* p, unsigned rank) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(s2_smem(p)), "r"(rank));
    return r;
}
// stores into another CTA's shared memory that count their bytes on an mbarrier there once they have landed (STAS)
__device__ __forceinline__ void ce_sta(unsigned addr, double v, unsigned bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(addr), "d"(v), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void ce_sta2(unsigned addr, double v0, double v1, unsigned bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];" ::"r"(addr), "d"(v0),
                 "d"(v1), "r"(bar) : "memory");
}
__device__ __forceinline__ void ce_st_s32(unsigned addr, int v) {
    asm volatile("st.shared::cluster.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void ce_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
// arrival that publishes nothing ("this CTA has read its panels"): no release fence
__device__ __forceinline__ void ce_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void ce_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// ---- one k4-step of the warp's contractions (cf. s2_chunk), in two halves ---------------------------------------------
// A warp computes an 8 x 8 sub-tile of every parity block of its field (accumulators [block][2]: the thread's two columns).
//   A = row panel at this k4-step + the thread's A-fragment offset (g * 4 + t), block b at A + b * CE_ABLK
//   Bp = column panel likewise, + 32 for the second 8-column half,               block b at Bp + b * CE_BBLK
// ce_kstep_opsA: operator rows (row panel) x state columns (column panel) -- needs the state rows 4 kk .. 4 kk + 3, which
//   the CTAs of row tile kk / 2 produce;
// ce_kstep_opsB: state rows (row panel) x operator rows at the tile's columns (column panel) -- needs the state columns
//   4 kk .. 4 kk + 3, which the CTAs of column tile kk / 4 produce.
// So three of a warp's eight half-steps work on state this CTA has produced itself: opsA at k4-step 2 trow + X, opsB at
// 4 tcol + X and 4 tcol + 2 + X (k-group X takes the k4-steps X, X + 2, X + 4, X + 6).  They run right after the CTA's own
// stores of the previous stage, while the other CTAs' stores are still crossing the SM-to-SM network; the other five follow
// once everything has arrived.  (The order of the accumulation differs from tile to tile but is fixed for a tile.)
template <bool P_LEFT>
__device__ __forceinline__ void ce_kstep_opsA(const double* __restrict__ A, const double* __restrict__ Bp, int fld,
                                              double (&a1)[4][2], double (&aL)[4][2]) {
    const int bf = S2_BU + 4 * fld;
    double FB[4], dA[2], ddA[2];
#pragma unroll
    for (int q = 0; q < 4; ++q) FB[q] = Bp[(bf + q) * CE_BBLK];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        dA[a] = A[a * CE_ABLK];
        ddA[a] = A[(2 + a) * CE_ABLK];
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            dmma884(a1[2 * a + b][0], a1[2 * a + b][1], dA[a], FB[2 * (1 - a) + b]);
            dmma884(aL[2 * a + b][0], aL[2 * a + b][1], ddA[a], FB[2 * a + b]);
        }
    if (P_LEFT) {
        double gA[2], PB[4];
#pragma unroll
        for (int a = 0; a < 2; ++a) gA[a] = A[(4 + a) * CE_ABLK];
#pragma unroll
        for (int q = 0; q < 4; ++q) PB[q] = Bp[(S2_BP + q) * CE_BBLK];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b) dmma884(aL[2 * a + b][0], aL[2 * a + b][1], gA[a], PB[2 * (1 - a) + b]);
    }
}
template <bool P_LEFT>
__device__ __forceinline__ void ce_kstep_opsB(const double* __restrict__ A, const double* __restrict__ Bp, int fld,
                                              double (&a2)[4][2], double (&aL)[4][2]) {
    const int bf = S2_BU + 4 * fld;
    double FA[4], dB[2], ddB[2];
#pragma unroll
    for (int q = 0; q < 4; ++q) FA[q] = A[(bf + q) * CE_ABLK];
#pragma unroll
    for (int b = 0; b < 2; ++b) {
        dB[b] = Bp[b * CE_BBLK];
        ddB[b] = Bp[(2 + b) * CE_BBLK];
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            dmma884(a2[2 * a + b][0], a2[2 * a + b][1], FA[2 * a + (1 - b)], dB[b]);
            dmma884(aL[2 * a + b][0], aL[2 * a + b][1], FA[2 * a + b], ddB[b]);
        }
    if (!P_LEFT) {
        double PA[4], gB[2];
#pragma unroll
        for (int q = 0; q < 4; ++q) PA[q] = A[(S2_BP + q) * CE_ABLK];
#pragma unroll
        for (int b = 0; b < 2; ++b) gB[b] = Bp[(4 + b) * CE_BBLK];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b) dmma884(aL[2 * a + b][0], aL[2 * a + b][1], PA[2 * a + (1 - b)], gB[b]);
    }
}
// the five half-steps of k-group X that need other CTAs' output
template <bool P_LEFT>
__device__ __forceinline__ void ce_remote_part(const double* __restrict__ rp, const double* __restrict__ cp, int X, int trow,
                                               int tcol, int fld, double (&a1)[4][2], double (&a2)[4][2], double (&aL)[4][2]) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const int kk = 4 * (1 - tcol) + 2 * c + X;
        ce_kstep_opsB<P_LEFT>(rp + (size_t)kk * (S2_NB * CE_ABLK), cp + (size_t)kk * (S2_NB * CE_BBLK), fld, a2, aL);
    }
#pragma unroll
    for (int c = 1; c < CE_CR; ++c) {
        const int kk = 2 * ((trow + c) % CE_CR) + X;
        ce_kstep_opsA<P_LEFT>(rp + (size_t)kk * (S2_NB * CE_ABLK), cp + (size_t)kk * (S2_NB * CE_BBLK), fld, a1, aL);
    }
}

// s2_dt with the two quotients that do not change during a visit, nu / dx_min^2 and nu / dy_min^2, taken out (the same
// operations on the same operands: the same bits).  Two square roots and three divisions are left: ~1 k cycles in a row.
__device__ __forceinline__ double ce_dt(double umax, double vmax, double lidv, double b2, double cfl, double dx_min, double dy_min,
                                        double nu_dx2, double nu_dy2) {
    const double ua = (lidv > umax) ? lidv : umax, va = (1e-10 > vmax) ? 1e-10 : vmax;
    const double lam_x = (ua + sqrt(ua * ua + b2)) / dx_min + nu_dx2;
    const double lam_y = (va + sqrt(va * va + b2)) / dy_min + nu_dy2;
    return cfl / (lam_x + lam_y);
}

// 4 sums + 2 NaN-propagating maxima across the eight warps, fixed order, result in every thread.
__device__ __forceinline__ void ce_block_reduce(double (&s)[4], double (&mx)[2], double* sRed, int warp, int lane) {
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
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 4; ++q) s[q] = sRed[q];
    mx[0] = sRed[4];
    mx[1] = sRed[5];
    for (int w = 1; w < CE_THREADS / 32; ++w) {
#pragma unroll
        for (int q = 0; q < 4; ++q) s[q] += sRed[w * 8 + q];
        mx[0] = nanmax(mx[0], sRed[w * 8 + 4]);
        mx[1] = nanmax(mx[1], sRed[w * 8 + 5]);
    }
    __syncthreads();
}

// Start of a visit: the boundary lines in parity form for the first stage, and the boundary-line parts of the convergence
// sums into S.step (every CTA of the cluster computes the same numbers in the same order).  cf. s2_visit_prepare.
__device__ __forceinline__ void ce_visit_prepare(const CESmem& S, bool first, const double* __restrict__ ed,
                                                 const double* __restrict__ ul, const double* __restrict__ vl, int m, int h, int n1) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int e = tid; e < 8 * S2_HP; e += CE_THREADS) {
        const int line = e / S2_HP, k = e % S2_HP;
        double x = 0.0, y = 0.0;
        if (k < h) {
            if (first) {
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
    // sums over the boundary lines of ||phi^{n+1} - phi^n||^2, ||phi^n||^2 and max |lid|: with the lines as stored (first
    // step after a load) and with the boundary-condition lines (every later step)
    double sf[4] = {0, 0, 0, 0}, sb[4] = {0, 0, 0, 0}, bm[2] = {0, 0}, dummy[2] = {0, 0};
    for (int e = tid; e < m; e += CE_THREADS) {
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
    ce_block_reduce(sf, bm, S.red, warp, lane);
    ce_block_reduce(sb, dummy, S.red, warp, lane);
    if (tid == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            S.step[2 + q] = sf[q];
            S.step[8 + q] = sb[q];
        }
        S.step[6] = bm[0];
        S.step[7] = bm[1];
    }
    __syncthreads();
}

// geometry of a CTA inside its cluster and the shared::cluster addresses it stores to
struct CEGeo {
    int rank, trow, tcol, r0, c0, m, h, n1;
    unsigned rrow[CE_CC];        // row panels of the CTAs of this tile row
    unsigned rcol[CE_CR];        // column panels of the CTAs of this tile column
    unsigned bar_off;            // data barrier of a CTA = its panel address + bar_off (one window per CTA, same layout)
    __device__ CEGeo(const S2Params& P, const CESmem& S) {
        rank = (int)ce_rank();
        trow = rank / CE_CC; tcol = rank % CE_CC;
        r0 = trow * CE_TR; c0 = tcol * CE_TC;
        m = P.m; h = eo_half(m); n1 = P.n1;
        // entry 0 is this CTA itself (stored to directly), entries 1.. the other CTAs of the row / column, starting with the next
#pragma unroll
        for (int c = 0; c < CE_CC; ++c) rrow[c] = ce_mapa(S.rp, trow * CE_CC + (tcol + c) % CE_CC);
#pragma unroll
        for (int r = 0; r < CE_CR; ++r) rcol[r] = ce_mapa(S.cp, ((trow + r) % CE_CR) * CE_CC + tcol);
        bar_off = ce_mapa(S.databar, 0) - ce_mapa(S.rp, 0);
    }
};

// Boundary-line corrections + nonlinear assembly at ONE element (row i, column j) of every parity block: s2_assemble for a
// single column.
//   in: b1 = D f, b2 = f D^T, bL = linear part (the thread's values at this column);  out: bL = momentum residual blocks,
//   b1 = d(field)/dx (u warps) or d(field)/dy (v warps)
__device__ __forceinline__ void ce_assemble_col(const double* __restrict__ sVec, const double* __restrict__ sEdgeF, double nu,
                                                bool bc_lines, bool is_u, int i, int j, const double (&uo)[4], const double (&vo)[4],
                                                double (&b1)[4], const double (&b2)[4], double (&bL)[4]) {
    constexpr int HP = S2_HP;
    double dx[4], dy[4];
    if (bc_lines) {   // W = E = S = 0: only the lid line contributes, and only to the y-derivatives
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const double dNj = sVec[(2 + b) * HP + j], ddNj = sVec[(6 + b) * HP + j];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const int q = 2 * a + b;
                const double Ni = sEdgeF[(6 + a) * HP + i];
                dx[q] = b1[q];
                dy[q] = fma(Ni, dNj, b2[q]);
                bL[q] = fma(nu * Ni, ddNj, bL[q]);
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
                const double d0i = sVec[(0 + a) * HP + i], dNi = sVec[(2 + a) * HP + i];
                const double dd0i = sVec[(4 + a) * HP + i], ddNi = sVec[(6 + a) * HP + i];
                const double Si = sEdgeF[(4 + a) * HP + i], Ni = sEdgeF[(6 + a) * HP + i];
                dx[q] = b1[q] + d0i * Wj + dNi * Ej;
                dy[q] = b2[q] + Si * d0j + Ni * dNj;
                bL[q] += nu * (dd0i * Wj + ddNi * Ej + Si * dd0j + Ni * ddNj);
            }
        }
    }
    // convective term in parity form: Z2 x Z2 group convolution (see eo_assemble)
    double cq[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        double c = uo[q] * dx[0];
#pragma unroll
        for (int bb = 1; bb < 4; ++bb) c = fma(uo[q ^ bb], dx[bb], c);
#pragma unroll
        for (int bb = 0; bb < 4; ++bb) c = fma(vo[q ^ bb], dy[bb], c);
        cq[q] = c;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        bL[q] -= cq[q];
        b1[q] = is_u ? dx[q] : dy[q];
    }
}

// the thread's value of one block (row i, column j) into the CTA's own panels / into the other CTAs' that hold it
//   offR / offC: its double offsets inside a row / column panel
__device__ __forceinline__ void ce_publish_local(const CESmem& S, unsigned offR, unsigned offC, double x) {
    S.rp[offR] = x;
    S.cp[offC] = x;
}
// The CTA's part of a stage output to the other CTAs' panels, out of its own (complete) panels: its tile's region of the row
// panel (k4-steps 4 tcol .. + 3: 64 nb double2) to the other CTA of the tile row, of the column panel (k4-steps 2 trow, + 1:
// 64 nb double2) to the three other CTAs of the tile column; nb = 8 state blocks (u, v) or 12 (u, v, p), which sit back to
// back.  ce_send_slot sends the double2 number e of each region (ce_send: thread tid takes e = 256 it + tid): 16-byte st.async
// stores, 512 contiguous bytes per warp instruction, bytes counted on the receiver's mbarrier.
__device__ __forceinline__ void ce_send_slot(const CESmem& S, const CEGeo& G, int e, int nb, int blk0 = S2_BU) {
    const unsigned cp_to_bar = G.bar_off - (unsigned)(CE_RP * sizeof(double));   // (the column panel follows the row panel)
    {
        const int kk = 4 * G.tcol + e / (16 * nb), blk = blk0 + (e / 16) % nb, w = e % 16;
        const unsigned off = (unsigned)((kk * S2_NB + blk) * CE_ABLK + 2 * w);
        const double2 x = *reinterpret_cast<const double2*>(S.rp + off);
#pragma unroll
        for (int c = 1; c < CE_CC; ++c) ce_sta2(G.rrow[c] + off * 8u, x.x, x.y, G.rrow[c] + G.bar_off);
    }
    {
        const int kk = 2 * G.trow + e / (32 * nb), blk = blk0 + (e / 32) % nb, w = e % 32;
        const unsigned off = (unsigned)((kk * S2_NB + blk) * CE_BBLK + 2 * w);
        const double2 x = *reinterpret_cast<const double2*>(S.cp + off);
#pragma unroll
        for (int r = 1; r < CE_CR; ++r) ce_sta2(G.rcol[r] + off * 8u, x.x, x.y, G.rcol[r] + cp_to_bar);
    }
}

__device__ __forceinline__ void ce_send(const CESmem& S, const CEGeo& G, int it, int nb, int blk0 = S2_BU) {
    ce_send_slot(S, G, it * CE_THREADS + (int)threadIdx.x, nb, blk0);
}

// After the K loop (and after the first arrival at the cluster barrier: this CTA has read its panels): the two k-groups swap
// accumulator halves -- group X keeps column X of every thread's pair -- then the epilogue of the thread's ONE element
// (row i, column j) of the four parity blocks: boundary corrections + nonlinear assembly, RK stage update, pressure update on
// the u warps, the CTA's own panels, at the last stage the tile's partial sums.  cf. s2_finish.  The stores to the other
// CTAs' panels are the caller's (ce_send, out of the CTA's own panels), between the half-steps of the next K loop's local part.
//   uo, vo: stage input at the thread's element; fN, pN: phi^n of the warp's field and p^n there (u warps)
template <int X>
__device__ __forceinline__ void ce_finish(const S2Params& P, const S2Visit& V, const CESmem& S, const CEGeo& G, int b, int stage,
                                          bool bc_lines, double dt, int fld, int nt, int lane, const double (&uo)[4],
                                          const double (&vo)[4], const double (&fN)[4], const double (&pN)[4],
                                          const double (&a1)[4][2], const double (&a2)[4][2], const double (&aL)[4][2],
                                          long long* __restrict__ tdbg) {
    const int g = lane >> 2, t = lane & 3, tid128 = (fld * 2 + nt) * 32 + lane, warp = (X * 2 + fld) * 2 + nt;
    const int m = P.m, h = G.h, centre = (m & 1) ? h - 1 : -1;
    const int i = G.r0 + g, j = G.c0 + 8 * nt + 2 * t + X;
    // ---- split-K hand-over: give away the other column, take the partner's contribution to this one ----
    double* exch_own = S.exch + X * 12 * 128;
    const double* exch_other = S.exch + (1 - X) * 12 * 128;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        exch_own[(0 + q) * 128 + tid128] = a1[q][1 - X];
        exch_own[(4 + q) * 128 + tid128] = a2[q][1 - X];
        exch_own[(8 + q) * 128 + tid128] = aL[q][1 - X];
    }
    __syncthreads();
    if (tdbg) tdbg[3] = clock64();
    double b1[4], b2[4], bL[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        // group 0's sum first, whichever group does the addition (bitwise the same in both)
        const double o1 = exch_other[(0 + q) * 128 + tid128], o2 = exch_other[(4 + q) * 128 + tid128],
                     oL = exch_other[(8 + q) * 128 + tid128];
        b1[q] = X == 0 ? a1[q][X] + o1 : o1 + a1[q][X];
        b2[q] = X == 0 ? a2[q][X] + o2 : o2 + a2[q][X];
        bL[q] = X == 0 ? aL[q][X] + oL : oL + aL[q][X];
    }
    ce_assemble_col(S.vec, S.edgep + (fld ? 8 * S2_HP : 0), V.nu, bc_lines, fld == 0, i, j, uo, vo, b1, b2, bL);

    const double adt = rk_alpha(stage) * dt;
    const bool last = (stage == 3);
    double s0 = 0.0, s1 = 0.0, mx = 0.0;   // sum d^2, sum phi_n^2, max |phi_new| of this warp's field
    double fn[4];
    {
        const double w = 4.0 * ((i == centre) ? 0.5 : 1.0) * (j == centre ? 0.5 : 1.0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double f0 = fN[q];
            fn[q] = f0 + adt * bL[q];
            if (last) {
                const double d = fn[q] - f0;
                s0 = fma(w * d, d, s0);
                s1 = fma(w * f0, f0, s1);
            }
            if (fld == 1) S.x[((X * 2 + nt) * 4 + q) * 32 + lane] = b1[q];   // dv/dy to the u warp of the same elements
        }
        if (last) {
            double fr[4];
            eo_to_points(fn[0], fn[1], fn[2], fn[3], fr);
#pragma unroll
            for (int q = 0; q < 4; ++q) mx = nanmax(mx, fabs(fr[q]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                s0 += __shfl_xor_sync(0xffffffffu, s0, o);
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                mx = nanmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            }
            if (lane == 0) {
                S.red[warp * 4 + 0] = s0;
                S.red[warp * 4 + 1] = s1;
                S.red[warp * 4 + 2] = mx;
            }
        }
    }
    // frozen pressure (SGSolver): every stage takes grad(p^n), so stages 0..2 leave the pressure blocks alone
    const bool keep_pn = P.freeze && stage != 3;
    const unsigned offR = (unsigned)(((j >> 2) * S2_NB + S2_BU + 4 * fld) * CE_ABLK + g * 4 + (j & 3));
    const unsigned offC = (unsigned)(((i >> 2) * S2_NB + S2_BU + 4 * fld) * CE_BBLK + (j & 15) * 4 + (i & 3));
    // the CTA's own panels (its own warps have all read them: the barrier after the swap); the other CTAs' below / by the caller
#pragma unroll
    for (int q = 0; q < 4; ++q) ce_publish_local(S, offR + q * CE_ABLK, offC + q * CE_BBLK, fn[q]);
    if (tdbg) tdbg[4] = clock64();
    __syncthreads();   // dv/dy of the v warps is in S.x, the warps' partial sums in S.red, u and v in the CTA's own panels
    ce_wait();         // every CTA of the cluster has finished reading its panels: they may be overwritten
    if (tdbg) tdbg[6] = clock64();
    if (last && threadIdx.x < CE_CLUSTER) {
        // this tile's partial sums to every CTA of the cluster.  warps: w = (2 X + fld) 2 + nt; u: warps 0, 1, 4, 5; v: 2, 3, 6, 7
        const double* r = S.red;
        const double o[6] = {((r[0 * 4 + 0] + r[1 * 4 + 0]) + r[4 * 4 + 0]) + r[5 * 4 + 0],
                             ((r[0 * 4 + 1] + r[1 * 4 + 1]) + r[4 * 4 + 1]) + r[5 * 4 + 1],
                             ((r[2 * 4 + 0] + r[3 * 4 + 0]) + r[6 * 4 + 0]) + r[7 * 4 + 0],
                             ((r[2 * 4 + 1] + r[3 * 4 + 1]) + r[6 * 4 + 1]) + r[7 * 4 + 1],
                             nanmax(nanmax(r[0 * 4 + 2], r[1 * 4 + 2]), nanmax(r[4 * 4 + 2], r[5 * 4 + 2])),
                             nanmax(nanmax(r[2 * 4 + 2], r[3 * 4 + 2]), nanmax(r[6 * 4 + 2], r[7 * 4 + 2]))};
        if ((int)threadIdx.x == G.rank) {
#pragma unroll
            for (int q = 0; q < 6; ++q) S.part[G.rank * 8 + q] = o[q];
        } else {
            const unsigned dst = ce_mapa(S.part + G.rank * 8, threadIdx.x), bar = ce_mapa(S.partbar, threadIdx.x);
#pragma unroll
            for (int q = 0; q < 6; ++q) ce_sta(dst + 8 * q, o[q], bar);
        }
    }
    // u and v leave for the other CTAs' panels ahead of the pressure update (not at the last stage of a step, where all the
    // sends run beside warp 0's reduction: the caller's)
    if (!last) {
        ce_send(S, G, 0, 8);
        ce_send(S, G, 1, 8);
    }
    if (fld == 0 && !keep_pn) {
        const unsigned pR = offR + (S2_BP - S2_BU) * CE_ABLK, pC = offC + (S2_BP - S2_BU) * CE_BBLK;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double dv = b1[q] + S.x[((X * 2 + nt) * 4 + q) * 32 + lane];
            ce_publish_local(S, pR + q * CE_ABLK, pC + q * CE_BBLK, pN[q] + adt * (-V.beta2 * dv));
        }
    }
    if (!bc_lines) {
        // first stage after a load done: from now on the boundary lines are the boundary conditions (walls zero, lid line)
        for (int e = threadIdx.x; e < 8 * S2_HP; e += CE_THREADS) {
            const int line = e / S2_HP, k = e % S2_HP;
            double x = 0.0, y = 0.0;
            if (k < h && (line & 3) == 3) {
                const double* lid = (line < 4 ? P.ulid : P.vlid) + (size_t)b * P.n1 + 1;
                x = lid[k];
                y = lid[m - 1 - k];
            }
            S.edgep[2 * line * S2_HP + k] = 0.5 * (x + y);
            S.edgep[(2 * line + 1) * S2_HP + k] = 0.5 * (x - y);
        }
    }
    __syncthreads();   // the CTA's own panels are complete: the caller runs the local part of the next K loop
    if (tdbg) tdbg[5] = clock64();
}

// The persistent kernel.  grid = clusters x 8 (cluster dimension 8 set at launch); a cluster pops a cavity from the queue,
// advances it `chunk` steps and hands it back.  256 threads: warp w = (2 X + fld) 2 + nt.
__global__ void __launch_bounds__(CE_THREADS, 1) k_cluster_eo(const S2Params P) {
    extern __shared__ __align__(16) double smem[];
    const CESmem S(smem);
    const CEGeo G(P, S);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int X = warp >> 2, fld = (warp >> 1) & 1, nt = warp & 1, g = lane >> 2, t = lane & 3;
    const int fo = g * 4 + t;   // the thread's fragment offset inside a block (A: row g; B: column 8 nt + g; k = t)
    const int m = G.m, h = G.h, n1 = G.n1;
    // LDC_S2_DEBUG: clock samples of warp 0 of CTA 0 per stage -- [0] stage start, [1] remote part of the K loop done, [2] own
    // elements read, [3] accumulator swap done, [4] assembly done, [5] own panels written, [6] every CTA has read its panels,
    // [10] local part of the next K loop + the stores to the other CTAs done, [8] the others' output has arrived; [9] cycles
    // of the end-of-step reduction in front of the stage
    const bool dbg_on = P.dbg != nullptr && blockIdx.x == 0 && tid == 0;
    int dbg_stage = 0;
    for (int e = tid; e < 8 * S2_HP; e += CE_THREADS) S.vec[e] = __ldg(P.ops + 6 * ST_BLK + e);
    if (tid == 0) {
        mbar_init(S.databar, 1);   // one arrival per stage (with the byte count the other CTAs will deliver)
        mbar_init(S.partbar, 1);   // one per step
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    unsigned nphase = 0;   // stages whose output has been waited for (the data barrier's phase)
    unsigned npart = 0;    // steps whose sums have been waited for
    ce_arrive();
    ce_wait();   // every CTA of the cluster is running: its shared memory may be written

    for (;;) {
        // ---- the cluster's next cavity ----
        if (G.rank == 0 && tid == 0) {
            const int bq = CavityQueue::pop(P);
            for (unsigned r = 0; r < CE_CLUSTER; ++r) ce_st_s32(ce_mapa(S.cavity, r), bq);
        }
        ce_arrive();
        ce_wait();
        const int b = *reinterpret_cast<volatile int*>(S.cavity);
        if (b < 0) return;

        CavityStatus st = load_status(P.status + b);
        const int cur = __ldcg(&P.status[b].pad);
        S2Visit V;
        V.nu = 1.0 / P.Re[b]; V.beta2 = P.beta2[b]; V.cfl = P.CFL[b]; V.lidv = P.lidvel[b];
        V.tol = P.tol ? P.tol[b] : -1.0;
        V.first = st.iters == 0 ? 1 : 0;
        long long nsteps = P.launch_end - st.iters;
        if (nsteps > P.chunk) nsteps = P.chunk;
        const double *ul = P.ulid + (size_t)b * n1, *vl = P.vlid + (size_t)b * n1;
        double* rec = P.rec + (size_t)cur * P.buf_stride + (size_t)b * S2_STRIDE;

        // ---- the tile's panels out of the record (L2): operator blocks and state blocks, row side and column side ----
        {
            const double* srcA = rec + s2_off(G.trow >> 1, 0, 0, (G.trow & 1) * 8, 0);
            for (int e = tid; e < CE_RP / 2; e += CE_THREADS) {
                const int grp = e >> 4, w = e & 15;   // 16 double2 per 8 x 4 block
                const int kk = grp / S2_NB, blk = grp % S2_NB;
                reinterpret_cast<double2*>(S.rp)[e] =
                    __ldcg(reinterpret_cast<const double2*>(srcA + ((size_t)kk * S2_NB + blk) * 64) + w);
            }
            const double* srcB = rec + S2_SIDE + s2_off(G.tcol, 0, 0, 0, 0);
            for (int e = tid; e < CE_CP / 2; e += CE_THREADS)   // the 18 blocks of a k4-step are contiguous on both sides
                reinterpret_cast<double2*>(S.cp)[e] = __ldcg(reinterpret_cast<const double2*>(srcB) + e);
        }
        ce_visit_prepare(S, V.first != 0, rec + S2_EDGE_OFF, ul, vl, m, h, n1);   // ends with a CTA barrier
        const double nu_dx2 = V.nu / (P.dx_min * P.dx_min), nu_dy2 = V.nu / (P.dy_min * P.dy_min);
        double dt = ce_dt(st.umax, st.vmax, V.lidv, V.beta2, V.cfl, P.dx_min, P.dy_min, nu_dx2, nu_dy2);
        bool bc_lines = !V.first, active = true;
        double fN[4], pN[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) fN[q] = pN[q] = 0.0;

        // end of a step: the eight tiles' partial sums + boundary lines, in fixed order, by every thread (the same numbers
        // everywhere): convergence bookkeeping, dt of the next step
        long long* red_dbg = nullptr;   // LDC_S2_DEBUG: [11..15] clock samples inside the reduction
        // Two warps share it at the end of a step (start_stage): warp 0 the residuals and the bookkeeping, warp 1 the next dt.
        auto reduce_maxima = [&]() {
            double mx[2] = {0.0, 0.0};
#pragma unroll
            for (int tl = 0; tl < CE_CLUSTER; ++tl) {
                mx[0] = nanmax(mx[0], S.part[tl * 8 + 4]);
                mx[1] = nanmax(mx[1], S.part[tl * 8 + 5]);
            }
            st.umax = nanmax(mx[0], S.step[6]);
            st.vmax = nanmax(mx[1], S.step[7]);
        };
        auto reduce_status = [&](bool first_step) {
            if (red_dbg) red_dbg[11] = clock64();
            const double* es = S.step + (first_step ? 2 : 8);
            double s[4] = {es[0], es[1], es[2], es[3]};
#pragma unroll
            for (int tl = 0; tl < CE_CLUSTER; ++tl)
#pragma unroll
                for (int q = 0; q < 4; ++q) s[q] += S.part[tl * 8 + q];
            reduce_maxima();
            if (red_dbg) red_dbg[12] = clock64();
            const double ures = sqrt(s[0]) / (sqrt(s[1]) + 1e-12), vres = sqrt(s[2]) / (sqrt(s[3]) + 1e-12);
            if (red_dbg) red_dbg[14] = clock64();
            update_status(st, ures, vres, V.tol, P.warmup, P.max_iter, P.nan_exit);
            active = st.active != 0;
            if (red_dbg) red_dbg[15] = clock64();
        };
        auto reduce_dt = [&]() {
            reduce_maxima();
            dt = ce_dt(st.umax, st.vmax, V.lidv, V.beta2, V.cfl, P.dx_min, P.dy_min, nu_dx2, nu_dy2);
        };
        // Start of a stage: accumulators to zero, then the part of the K loop that needs no one else's output (three
        // half-steps) -- with the st.async stores of the previous stage's output to the other CTAs' panels (ce_send, out of
        // this CTA's own panels, which ce_finish has just completed) issued in between: the SM-to-SM network takes ~20 B/clk
        // out of an SM, and the DMMAs run while the stores queue up for it.
        double a1[4][2], a2[4][2], aL[4][2];
        const int j_own = G.c0 + 8 * nt + 2 * t + X;   // the thread's element in the epilogue: row r0 + g, column j_own
        const double* rpf = S.rp + fo;                  // fragment bases: row panel, column panel at the warp's 8-column half
        const double* cpf = S.cp + fo + nt * 32;
        auto start_stage = [&](bool publish, bool with_p, bool reduce_now, bool first_step) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) a1[q][hh] = a2[q][hh] = aL[q][hh] = 0.0;
            const int ka = 2 * G.trow + X, kb0 = 4 * G.tcol + X, kb1 = kb0 + 2;
            const int nb = with_p ? 12 : 8;
            if (reduce_now) {
                // Last stage of a step: warps 0 and 1 run the step's reduction (residuals + bookkeeping, and the next dt:
                // ~1.2 k cycles of dependent FP64 operations each; the others' sums left ahead of their stage output) while the
                // other six warps push the stage output into the network.
                if (warp < 2) {
                    mbar_wait(S.partbar, npart & 1);
                    if (warp == 0) {
                        reduce_status(first_step);
                        if (lane == 0) S.step[1] = active ? 1.0 : 0.0;
                    } else {
                        reduce_dt();
                        if (lane == 0) S.step[0] = dt;
                    }
                } else {
                    for (int e = tid - 64; e < 64 * nb; e += CE_THREADS - 64) ce_send_slot(S, G, e, nb);
                }
                ++npart;
            }
            // (stages 0..2: u and v have left in ce_finish; the pressure blocks, one double2 of each region per thread, now)
            if (publish && !reduce_now && with_p) ce_send(S, G, 0, 4, S2_BP);
            if (fld == 0) ce_kstep_opsA<true>(rpf + (size_t)ka * (S2_NB * CE_ABLK), cpf + (size_t)ka * (S2_NB * CE_BBLK), 0, a1, aL);
            else ce_kstep_opsA<false>(rpf + (size_t)ka * (S2_NB * CE_ABLK), cpf + (size_t)ka * (S2_NB * CE_BBLK), 1, a1, aL);
            if (fld == 0) ce_kstep_opsB<true>(rpf + (size_t)kb0 * (S2_NB * CE_ABLK), cpf + (size_t)kb0 * (S2_NB * CE_BBLK), 0, a2, aL);
            else ce_kstep_opsB<false>(rpf + (size_t)kb0 * (S2_NB * CE_ABLK), cpf + (size_t)kb0 * (S2_NB * CE_BBLK), 1, a2, aL);
            if (fld == 0) ce_kstep_opsB<true>(rpf + (size_t)kb1 * (S2_NB * CE_ABLK), cpf + (size_t)kb1 * (S2_NB * CE_BBLK), 0, a2, aL);
            else ce_kstep_opsB<false>(rpf + (size_t)kb1 * (S2_NB * CE_ABLK), cpf + (size_t)kb1 * (S2_NB * CE_BBLK), 1, a2, aL);
        };
        start_stage(false, false, false, false);
        for (long long step = 0; step < nsteps; ++step) {
            for (int stage = 0; stage < 4; ++stage) {
                long long* tdbg = (dbg_on && dbg_stage < S2_DBG_STAGES) ? P.dbg + dbg_stage * 16 : nullptr;
                ++dbg_stage;
                if (tdbg) tdbg[0] = clock64();
                if (tid == 0) {
                    // what the other CTAs will deliver for this stage: u, v (8 blocks) and, unless the pressure is frozen, p
                    // (4 blocks) of one tile into the row panel and of three into the column panel; at the last stage 6 sums
                    // from each of the 7 others
                    const unsigned blocks = (P.freeze && stage != 3) ? 8u : 12u;
                    mbar_expect_tx(S.databar, 4u * blocks * (CE_TR * CE_TC * 8u));
                    if (stage == 3) mbar_expect_tx(S.partbar, 7u * 48u);
                }
                // the rest of the K loop: everything has arrived
                if (fld == 0) ce_remote_part<true>(rpf, cpf, X, G.trow, G.tcol, 0, a1, a2, aL);
                else ce_remote_part<false>(rpf, cpf, X, G.trow, G.tcol, 1, a1, a2, aL);
                if (tdbg) tdbg[1] = clock64();
                // the stage input at the thread's own element (row i_own, column j_own), before anyone overwrites it
                double uo[4], vo[4];
                {
                    const double* own = S.rp + (size_t)(j_own >> 2) * (S2_NB * CE_ABLK) + g * 4 + (j_own & 3);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        uo[q] = own[(S2_BU + q) * CE_ABLK];
                        vo[q] = own[(S2_BV + q) * CE_ABLK];
                        if (stage == 0) {   // phi^n of the warp's field (u warps: also p^n), kept for the four stages
                            fN[q] = fld == 0 ? uo[q] : vo[q];
                            if (fld == 0) pN[q] = own[(S2_BP + q) * CE_ABLK];
                        }
                    }
                }
                ce_arrive_relaxed();   // this CTA has read its panels (the wait is in ce_finish, behind the assembly)
                if (tdbg) tdbg[2] = clock64();
                if (X == 0) ce_finish<0>(P, V, S, G, b, stage, bc_lines, dt, fld, nt, lane, uo, vo, fN, pN, a1, a2, aL, tdbg);
                else ce_finish<1>(P, V, S, G, b, stage, bc_lines, dt, fld, nt, lane, uo, vo, fN, pN, a1, a2, aL, nullptr);
                bc_lines = true;
                red_dbg = tdbg;
                start_stage(true, !(P.freeze && stage != 3), stage == 3, V.first && step == 0);
                if (stage == 3) {
                    __syncthreads();   // warp 0's verdict on the step
                    dt = S.step[0];
                    active = S.step[1] != 0.0;
                }
                if (tdbg) tdbg[10] = clock64();
                mbar_wait(S.databar, nphase & 1);   // everyone else's part of the stage output is in this CTA's panels
                ++nphase;
                if (tdbg) tdbg[8] = clock64();
            }
            if (!active) break;   // (the local part run for a next step was speculative: it only reads)
        }

        // ---- visit end: the state back into the record (rows by the CTAs of column tile 0, columns by those of row tile
        // 0), boundary-condition lines and status by CTA 0, then the cavity goes back to the queue ----
        if (G.tcol == 0) {
            double* dstA = rec + s2_off(G.trow >> 1, 0, S2_BU, (G.trow & 1) * 8, 0);
            for (int e = tid; e < CE_NK4 * 12 * 16; e += CE_THREADS) {
                const int grp = e >> 4, w = e & 15;
                const int kk = grp / 12, fb = grp % 12;
                __stcg(reinterpret_cast<double2*>(dstA + ((size_t)kk * S2_NB + fb) * 64) + w,
                       reinterpret_cast<const double2*>(S.rp + ((size_t)kk * S2_NB + S2_BU + fb) * CE_ABLK)[w]);
            }
        }
        if (G.trow == 0) {
            double* dstB = rec + S2_SIDE + s2_off(G.tcol, 0, S2_BU, 0, 0);
            for (int e = tid; e < CE_NK4 * 12 * 32; e += CE_THREADS) {
                const int kk = e / (12 * 32), w = e % (12 * 32);
                __stcg(reinterpret_cast<double2*>(dstB + (size_t)kk * S2_NB * 64) + w,
                       reinterpret_cast<const double2*>(S.cp + ((size_t)kk * S2_NB + S2_BU) * CE_BBLK)[w]);
            }
        }
        if (G.rank == 0 && warp == 0) {
            double* edn = rec + S2_EDGE_OFF;
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
        }
        __threadfence();
        ce_arrive();
        ce_wait();   // the whole record is written
        if (G.rank == 0 && tid == 0) {
            store_status(P.status + b, st);
            __threadfence();
            if (st.active && st.iters < P.launch_end) CavityQueue::push(P, b);
            else CavityQueue::retire(P);
        }
    }
}

}  // namespace ldc
