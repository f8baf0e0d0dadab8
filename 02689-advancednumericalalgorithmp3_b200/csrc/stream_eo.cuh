// Streamed even-odd path for grids whose state does not fit one SM (65 <= nx - 1 <= 128, e.g. N = 128).
//
// Same parity algebra as the resident even-odd kernel (resident_eo.cuh), but a cavity is spread over
// (HP/16)^2 CTAs: each CTA owns a 16 x 16 tile of all four parity blocks, streams the operand panels it needs
// from L2 through shared memory in K-chunks of 16 and runs one RK stage; one launch per stage, state in three
// rotating global buffers (phi^n, stage in, stage out).  Eight cavities at N = 128 thus occupy 128 CTAs instead
// of 8 SMs.  Per-cavity reductions (dt, convergence norms) are per-tile partials reduced in fixed order by
// k_stream_end, so results do not depend on scheduling.
#pragma once
#include "resident_eo_kernel.cuh"

namespace ldc {

constexpr int ST_HP = 64, ST_LD = ST_HP + 4, ST_BLK = ST_HP * ST_LD;   // global parity blocks (eo layout, HP = 64)
constexpr int ST_T = 16;                 // tile edge (rows and columns of each parity block)
constexpr int ST_KC = 16, ST_LDK = 20;   // K-chunk and padded chunk row (= 4 mod 8: conflict-free fragments)
constexpr int ST_PANEL = ST_T * ST_LDK;  // doubles per staged chunk panel (row-type: T x KC, column-type: KC x T)
constexpr int ST_TILES = (ST_HP / ST_T) * (ST_HP / ST_T);
constexpr int ST_THREADS = 128;          // 4 warps: (row half) x (u equation | v equation)
constexpr int ST_ME = 2 * ST_HP;
// shared memory: 36 chunk panels (x 2 when double-buffered: few CTAs, one per SM; single-buffered two CTAs share an
// SM and overlap each other's loads) + boundary-column vectors (8 x HP) + boundary lines in parity form (16 x HP)
// + divergence exchange (2 x 16 x 32) + reduction scratch
constexpr size_t st_smem_doubles(bool db) { return (size_t)(db ? 2 : 1) * 36 * ST_PANEL + 8 * ST_HP + 16 * ST_HP + 2 * 16 * 32 + 32; }

// Programmatic dependent launch (sm_90+): launch_dependents lets the next kernel of the stream start once every CTA of this
// grid has issued it (or exited); wait blocks until the preceding grid has completed and its writes are visible.  Both are
// no-ops for a kernel launched without the programmatic-serialization attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

struct StreamParams {
    const double* ops;            // eo packed operators, HP = 64
    const double* opsc;           // [B][6 blocks] per-cavity copies: D, nu DD, -G (cp.async sources)
    double* buf[3];               // rotating state buffers (eo packed records)
    size_t stride;                // doubles per cavity record
    const double *Re, *beta2, *CFL, *lidvel, *ulid, *vlid;
    double* dt;                   // [B]
    int n1, m, K4;
    CavityStatus* status;         // status.pad = index of the buffer holding phi^n
    double* partials;             // [B][ST_TILES][8]: s0..s3, max|u|, max|v|
    int stage, freeze, bc_edges;  // bc_edges = 0: boundary lines as stored with phi^n (first stage after a load)
    // end-of-step bookkeeping (k_stream_end)
    const double* tol;            // [B] or null
    int first, warmup, nan_exit;  // first = 1: first step after a load (boundary lines as stored)
    long long max_iter;
    double dx_min, dy_min;
};

__device__ __forceinline__ void cp_async16(double* smem_dst, const double* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}

// per-cavity operator copies for the streamed kernel: D as is, DD scaled by nu = 1/Re, G negated
__global__ void k_stream_scale_ops(const double* __restrict__ ops, const double* __restrict__ Re, double* __restrict__ opsc) {
    const int b = blockIdx.y;
    const double nu = 1.0 / Re[b];
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < 6 * ST_BLK; e += gridDim.x * blockDim.x) {
        const int blk = e / ST_BLK;
        const double sc = blk < 2 ? 1.0 : (blk < 4 ? nu : -1.0);
        opsc[(size_t)b * 6 * ST_BLK + e] = sc * ops[e];
    }
}

__device__ __forceinline__ int st_in(int cur, int stage) { return stage == 0 ? cur : (stage == 2 ? (cur + 2) % 3 : (cur + 1) % 3); }
__device__ __forceinline__ int st_out(int cur, int stage) { return (stage & 1) ? (cur + 2) % 3 : (cur + 1) % 3; }

// One K-chunk of the contractions of the warp's field (see eo_field_pass): operands come from staged panels.
//   sOa: 6 operator row panels at the tile rows   sOb: 6 operator row panels at the tile columns
//   sFa: 4 row panels of the field                 sFb: 4 column panels of the field
//   sPa / sPb: 4 row / column panels of the pressure
template <bool P_LEFT>
__device__ __forceinline__ void st_chunk(const double* __restrict__ sOa, const double* __restrict__ sOb,
                                         const double* __restrict__ sFa, const double* __restrict__ sFb,
                                         const double* __restrict__ sPa, const double* __restrict__ sPb, int ks0, int ks1,
                                         int wr, int g, int t, double (&a1)[4][2][2], double (&a2)[4][2][2],
                                         double (&aL)[4][2][2]) {
    constexpr int NT = 2;
    const int arow = (wr + g) * ST_LDK + t;   // row-type panel: X[wr + g][k + t]
    const int brow = g * ST_LDK + t;          // operator rows at the tile columns: Op[8 nt + g][k + t]
    const int bfld = t * ST_LDK + g;          // column-type panel: F[k + t][8 nt + g]
#pragma unroll 1
    for (int ks = ks0; ks < ks1; ++ks) {
        const int k0 = 4 * ks;
        double FA[4], FB[4][NT];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            FA[q] = sFa[q * ST_PANEL + arow + k0];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) FB[q][nt] = sFb[q * ST_PANEL + bfld + k0 * ST_LDK + nt * 8];
        }
        {
            double dA[2], ddA[2];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                dA[a] = sOa[a * ST_PANEL + arow + k0];
                ddA[a] = sOa[(2 + a) * ST_PANEL + arow + k0];
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
                    dB[b][nt] = sOb[b * ST_PANEL + brow + nt * 8 * ST_LDK + k0];
                    ddB[b][nt] = sOb[(2 + b) * ST_PANEL + brow + nt * 8 * ST_LDK + k0];
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
            for (int a = 0; a < 2; ++a) gA[a] = sOa[(4 + a) * ST_PANEL + arow + k0];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) PB[q][nt] = sPb[q * ST_PANEL + bfld + k0 * ST_LDK + nt * 8];
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
            for (int q = 0; q < 4; ++q) PA[q] = sPa[q * ST_PANEL + arow + k0];
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) gB[b][nt] = sOb[(4 + b) * ST_PANEL + brow + nt * 8 * ST_LDK + k0];
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

}  // namespace ldc

namespace ldc {

// Boundary-line corrections + nonlinear assembly for the warp's field (cf. eo_assemble), with the thread's own
// stage values of u and v in registers.  On return aL = momentum residual blocks, a1 = d(field)/dx (IS_U) or
// d(field)/dy (!IS_U) blocks.
template <bool IS_U>
__device__ __forceinline__ void st_assemble(const double* __restrict__ sVec, const double* __restrict__ sEdgeF, double nu,
                                            bool bc_lines, int i, int c0, int t, const double (&uo)[4][2][2], const double (&vo)[4][2][2],
                                            double (&a1)[4][2][2], double (&a2)[4][2][2], double (&aL)[4][2][2]) {
    constexpr int HP = ST_HP;
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
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int j = c0 + nt * 8 + 2 * t + hh;
            double dx[4], dy[4];
            if (bc_lines) {   // W = E = S = 0: only the lid line contributes, and only to the y-derivatives
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const double dNj = sVec[(2 + b) * HP + j], ddNj = sVec[(6 + b) * HP + j];
#pragma unroll
                    for (int a = 0; a < 2; ++a) {
                        const int q = 2 * a + b;
                        dx[q] = a1[q][nt][hh];
                        dy[q] = fma(Ni[a], dNj, a2[q][nt][hh]);
                        aL[q][nt][hh] = fma(nu * Ni[a], ddNj, aL[q][nt][hh]);
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
                        dx[q] = a1[q][nt][hh] + d0i[a] * Wj + dNi[a] * Ej;
                        dy[q] = a2[q][nt][hh] + Si[a] * d0j + Ni[a] * dNj;
                        aL[q][nt][hh] += nu * (dd0i[a] * Wj + ddNi[a] * Ej + Si[a] * dd0j + Ni[a] * ddNj);
                    }
                }
            }
            // convective term in parity form: Z2 x Z2 group convolution (see eo_assemble)
            double cq[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                double c = uo[q][nt][hh] * dx[0];
#pragma unroll
                for (int bb = 1; bb < 4; ++bb) c = fma(uo[q ^ bb][nt][hh], dx[bb], c);
#pragma unroll
                for (int bb = 0; bb < 4; ++bb) c = fma(vo[q ^ bb][nt][hh], dy[bb], c);
                cq[q] = c;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                aL[q][nt][hh] -= cq[q];
                a1[q][nt][hh] = IS_U ? dx[q] : dy[q];
            }
        }
}

// End of an RK4 step: reduce the tile partials in fixed order, add the boundary lines, convergence bookkeeping,
// next dt, rotate the buffers and store the boundary-condition lines with the new state.
// 128 threads per cavity; sh = 32 doubles of reduction scratch, shp = ST_TILES x 6 doubles for the tile partials.
// lidv, b2, nu, cfl, tolb: the cavity's constants, fetched by the caller before it waits for the stage kernels.
__device__ __forceinline__ void stream_end_body(const StreamParams& P, int b, double* sh, double* shp, double lidv, double b2,
                                                double nu, double cfl, double tolb) {
    const int first = P.first, warmup = P.warmup, nan_exit = P.nan_exit;
    const long long max_iter = P.max_iter;
    const double dx_min = P.dx_min, dy_min = P.dy_min;
    double* dt = P.dt;
    CavityStatus st = P.status[b];
    // the tile partials arrive in parallel (one L2 round trip); thread 0 adds them in fixed order below
    const double* part = P.partials + (size_t)b * ST_TILES * 8;
    for (int e = threadIdx.x; e < ST_TILES * 6; e += blockDim.x) shp[e] = __ldcg(part + (e / 6) * 8 + e % 6);
    const int m = P.m, n1 = P.n1, cur = st.pad, nxt = (cur + 2) % 3;
    const double* ed = P.buf[cur] + (size_t)b * P.stride + 12 * ST_BLK;   // boundary lines of phi^n
    double* edn = P.buf[nxt] + (size_t)b * P.stride + 12 * ST_BLK;
    const double *ul = P.ulid + (size_t)b * n1, *vl = P.vlid + (size_t)b * n1;
    double s[4] = {0, 0, 0, 0}, bm[2] = {0, 0};
    for (int e = threadIdx.x; e < m; e += blockDim.x) {
#pragma unroll
        for (int f = 0; f < 2; ++f) {
            const double lid = (f ? vl : ul)[e + 1];
            const double* L = ed + f * 4 * ST_ME;
            const double w = first ? L[e] : 0.0, ea = first ? L[ST_ME + e] : 0.0, so = first ? L[2 * ST_ME + e] : 0.0;
            const double no = first ? L[3 * ST_ME + e] : lid;
            const double dn = lid - no;
            s[2 * f] += w * w + ea * ea + so * so + dn * dn;
            s[2 * f + 1] += w * w + ea * ea + so * so + no * no;
            bm[f] = nanmax(bm[f], fabs(lid));
            double* Ln = edn + f * 4 * ST_ME;
            Ln[e] = 0.0; Ln[ST_ME + e] = 0.0; Ln[2 * ST_ME + e] = 0.0; Ln[3 * ST_ME + e] = lid;
        }
    }
    if (threadIdx.x == 0) {
#pragma unroll
        for (int f = 0; f < 2; ++f) {
            const double* c = ed + 8 * ST_ME + 4 * f;   // c00 c0N cN0 cNN
            const double l0 = (f ? vl : ul)[0], lN = (f ? vl : ul)[n1 - 1];
            const double c0 = first ? c[0] : 0.0, c1 = first ? c[1] : l0, c2 = first ? c[2] : 0.0, c3 = first ? c[3] : lN;
            const double d1 = l0 - c1, d3 = lN - c3;
            s[2 * f] += c0 * c0 + d1 * d1 + c2 * c2 + d3 * d3;
            s[2 * f + 1] += c0 * c0 + c1 * c1 + c2 * c2 + c3 * c3;
            bm[f] = nanmax(nanmax(bm[f], fabs(l0)), fabs(lN));
            double* cn = edn + 8 * ST_ME + 4 * f;
            cn[0] = 0.0; cn[1] = l0; cn[2] = 0.0; cn[3] = lN;
        }
    }
    // deterministic CTA reduction of the 4 sums + 2 maxima
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    cta_reduce<4>(s, bm, sh, warp, lane);
    if (threadIdx.x == 0) {
        double mx[2] = {0, 0};
        for (int tl = 0; tl < ST_TILES; ++tl) {
#pragma unroll
            for (int q = 0; q < 4; ++q) s[q] += shp[tl * 6 + q];
            mx[0] = nanmax(mx[0], shp[tl * 6 + 4]);
            mx[1] = nanmax(mx[1], shp[tl * 6 + 5]);
        }
        const double ures = sqrt(s[0]) / (sqrt(s[1]) + 1e-12), vres = sqrt(s[2]) / (sqrt(s[3]) + 1e-12);
        update_status(st, ures, vres, tolb, warmup, max_iter, nan_exit);
        st.umax = nanmax(mx[0], bm[0]);
        st.vmax = nanmax(mx[1], bm[1]);
        st.pad = nxt;
        P.status[b] = st;
        const double ua = (lidv > st.umax) ? lidv : st.umax, va = (1e-10 > st.vmax) ? 1e-10 : st.vmax;
        const double lam_x = (ua + sqrt(ua * ua + b2)) / dx_min + nu / (dx_min * dx_min);
        const double lam_y = (va + sqrt(va * va + b2)) / dy_min + nu / (dy_min * dy_min);
        dt[b] = cfl / (lam_x + lam_y);
    }
}

// One RK stage of every active cavity: grid (ST_TILES, B).
// DB = false (many CTAs, two per SM): 128 threads, single panel buffer, the two CTAs of an SM overlap each other.
// DB = true (few CTAs, one per SM): double-buffered panels and 256 threads -- a second group of four warps takes the
// upper half of the k-steps of every chunk (split-K), so that each SM sub-partition has two warps issuing DMMAs (one warp
// alone is latency-bound: 25 instead of 16 cycles per DMMA, profiles/README.md visit 6); its accumulators are added to
// the first group's through shared memory, in fixed order, and the first group runs the epilogue.
template <bool DB>
__global__ void __launch_bounds__(DB ? 2 * ST_THREADS : ST_THREADS, DB ? 1 : 2) k_stream_stage(const StreamParams P) {
    constexpr int NTHR = DB ? 2 * ST_THREADS : ST_THREADS;
    extern __shared__ __align__(16) double smem[];
    double* sPan = smem;
    double* sVec = sPan + (DB ? 2 : 1) * 36 * ST_PANEL;
    double* sEdgeP = sVec + 8 * ST_HP;
    double* sX = sEdgeP + 16 * ST_HP;
    double* sRed = sX + 2 * 16 * 32;

    const int b = blockIdx.y;
    const int tile = blockIdx.x, r0 = (tile / (ST_HP / ST_T)) * ST_T, c0 = (tile % (ST_HP / ST_T)) * ST_T;
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int kgrp = DB ? (threadIdx.x >> 7) : 0;          // split-K group (DB only)
    const int tid = threadIdx.x & (ST_THREADS - 1);         // thread index inside the group
    const int warp = tid >> 5, tw = warp & 1, fld = warp >> 1, wr = 8 * tw;
    const int m = P.m, h = eo_half(m), centre = (m & 1) ? h - 1 : -1;
    if (r0 >= h || c0 >= h) return;   // tile lies in the zero padding of the blocks

    // Programmatic dependent launch: the stage kernels and k_stream_end of a run form a chain in which each launch may
    // start while its predecessor drains.  Everything up to pdl_wait() touches only data no kernel of the chain writes
    // (operators, lid profiles): the boundary vectors and the operator panels of the first chunk are on their way while
    // the predecessor's last tiles finish.
    pdl_launch_dependents();
    const double nu = 1.0 / __ldg(P.Re + b), beta2 = __ldg(P.beta2 + b);   // per-cavity constants: not written by the chain
    const int lr = tid >> 3, lc = 2 * (tid & 7);   // this thread's (row, column pair) inside a 16 x 16 chunk panel
    const double* opsc = P.opsc + (size_t)b * 6 * ST_BLK;
    // operators come pre-scaled per cavity from P.opsc: D, nu DD, -G
    auto issue_ops = [&](int k0, double* buf) {
#pragma unroll
        for (int p = 0; p < 12; ++p) {
            if (DB && (p & 1) != kgrp) continue;   // each group fetches every other panel
            const double* src = p < 6 ? opsc + (size_t)p * ST_BLK + (r0 + lr) * ST_LD + k0 + lc            // rows at the tile rows
                                      : opsc + (size_t)(p - 6) * ST_BLK + (c0 + lr) * ST_LD + k0 + lc;     // ... at the tile columns
            cp_async16(buf + p * ST_PANEL + lr * ST_LDK + lc, src);
        }
    };
    issue_ops(0, sPan);
    for (int e = threadIdx.x; e < 8 * ST_HP; e += NTHR) sVec[e] = __ldg(P.ops + 6 * ST_BLK + e);
    // ---- boundary lines in parity form: uW uE uS uN vW vE vS vN ----
    if (P.bc_edges) {   // W, E, S are zero and the lid line is the prescribed profile (every stage but the first after a load)
        for (int e = threadIdx.x; e < 8 * ST_HP; e += NTHR) {
            const int line = e / ST_HP, k = e % ST_HP;
            double x = 0.0, y = 0.0;
            if (k < h && (line & 3) == 3) {
                const double* lid = (line < 4 ? P.ulid : P.vlid) + (size_t)b * P.n1 + 1;
                x = lid[k];
                y = lid[m - 1 - k];
            }
            sEdgeP[2 * line * ST_HP + k] = 0.5 * (x + y);
            sEdgeP[(2 * line + 1) * ST_HP + k] = 0.5 * (x - y);
        }
    }
    pdl_wait();   // the predecessor's state, status and dt are complete and visible from here on

    const CavityStatus st = P.status[b];
    const double dt_b = __ldcg(P.dt + b);   // on its way with the status (used by the epilogue only)
    if (!st.active) {
        asm volatile("cp.async.wait_all;" ::: "memory");
        return;
    }
    const int cur = st.pad;
    const double* recN = P.buf[cur] + (size_t)b * P.stride;                       // phi^n
    const double* recI = P.buf[st_in(cur, P.stage)] + (size_t)b * P.stride;       // stage input
    double* recO = P.buf[st_out(cur, P.stage)] + (size_t)b * P.stride;            // stage output
    const double* recP = P.freeze ? recN : recI;                                  // pressure operand

    // ---- K loop: the 36 chunk panels arrive by cp.async into two alternating buffers while the previous chunk is
    // contracted ----
    auto issue_state = [&](int k0, double* buf) {
#pragma unroll
        for (int p = 12; p < 36; ++p) {
            if (DB && (p & 1) != kgrp) continue;
            const double* src;
            if (p < 24) src = (p < 20 ? recI : recP) + (size_t)(p - 12) * ST_BLK + (r0 + lr) * ST_LD + k0 + lc;   // row panels
            else src = (p < 32 ? recI : recP) + (size_t)(p - 24) * ST_BLK + (k0 + lr) * ST_LD + c0 + lc;          // column panels
            cp_async16(buf + p * ST_PANEL + lr * ST_LDK + lc, src);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto issue = [&](int k0, double* buf) {
        issue_ops(k0, buf);
        issue_state(k0, buf);
    };
    const int nchunks = (P.K4 + ST_KC - 1) / ST_KC;
    issue_state(0, sPan);

    if (!P.bc_edges) {   // first stage after a load: the lines as stored with phi^n
        for (int e = threadIdx.x; e < 8 * ST_HP; e += NTHR) {
            const int line = e / ST_HP, k = e % ST_HP;
            double x = 0.0, y = 0.0;
            if (k < h) {
                const double* src = recN + 12 * ST_BLK + line * ST_ME;
                x = __ldcg(src + k);
                y = __ldcg(src + m - 1 - k);
            }
            sEdgeP[2 * line * ST_HP + k] = 0.5 * (x + y);
            sEdgeP[(2 * line + 1) * ST_HP + k] = 0.5 * (x - y);
        }
    }

    double a1[4][2][2], a2[4][2][2], aL[4][2][2];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) a1[q][nt][hh] = a2[q][nt][hh] = aL[q][nt][hh] = 0.0;

    for (int c = 0; c < nchunks; ++c) {
        double* cur_buf = sPan + (DB ? (c & 1) * 36 * ST_PANEL : 0);
        if (DB && c + 1 < nchunks) {
            issue((c + 1) * ST_KC, sPan + ((c + 1) & 1) * 36 * ST_PANEL);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();   // chunk c has landed for every thread (also orders the sVec / sEdgeP writes before their use)
        if (DB && c + 1 == nchunks && kgrp == 0) {
            // Last chunk: the other panel buffer is idle.  The epilogue's operands (this thread's own elements of the
            // stage input u, v, of phi^n of the warp's field and of p^n) are fetched into it now, behind the DMMAs.
            double* ep = sPan + ((c + 1) & 1) * 36 * ST_PANEL;
            const int ei = (r0 + wr + g) * ST_LD + c0 + 2 * t;
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int sl = q * 2 + nt, e = ei + nt * 8;
                    cp_async16(ep + ((0 + sl) * ST_THREADS + tid) * 2, recI + (size_t)q * ST_BLK + e);
                    cp_async16(ep + ((8 + sl) * ST_THREADS + tid) * 2, recI + (size_t)(4 + q) * ST_BLK + e);
                    cp_async16(ep + ((16 + sl) * ST_THREADS + tid) * 2, recN + (size_t)(4 * fld + q) * ST_BLK + e);
                    if (fld == 0) cp_async16(ep + ((24 + sl) * ST_THREADS + tid) * 2, recN + (size_t)(8 + q) * ST_BLK + e);
                }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        const int ksteps = min(ST_KC, P.K4 - c * ST_KC) / 4;
        const int ksplit = DB ? (ksteps + 1) / 2 : ksteps;          // group 0: [0, ksplit), group 1: [ksplit, ksteps)
        const int ks0 = kgrp ? ksplit : 0, ks1 = kgrp ? ksteps : ksplit;
        if (fld == 0)
            st_chunk<true>(cur_buf, cur_buf + 6 * ST_PANEL, cur_buf + 12 * ST_PANEL, cur_buf + 24 * ST_PANEL,
                           cur_buf + 20 * ST_PANEL, cur_buf + 32 * ST_PANEL, ks0, ks1, wr, g, t, a1, a2, aL);
        else
            st_chunk<false>(cur_buf, cur_buf + 6 * ST_PANEL, cur_buf + 16 * ST_PANEL, cur_buf + 28 * ST_PANEL,
                            cur_buf + 20 * ST_PANEL, cur_buf + 32 * ST_PANEL, ks0, ks1, wr, g, t, a1, a2, aL);
        __syncthreads();   // chunk c consumed: its buffer may be refilled by the next issue
        if (!DB && c + 1 < nchunks) issue((c + 1) * ST_KC, sPan);
    }

    if (DB) {
        // split-K hand-off: group 1 parks its accumulators in the panel buffer of the last chunk (free since the barrier
        // above; the prefetched epilogue operands sit in the other one) and leaves; group 0 adds them, always in this order
        double* sAcc = sPan + ((nchunks - 1) & 1) * 36 * ST_PANEL;
        if (kgrp == 1) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int idx = (q * 2 + nt) * 2 + hh;
                        sAcc[(0 + idx) * ST_THREADS + tid] = a1[q][nt][hh];
                        sAcc[(16 + idx) * ST_THREADS + tid] = a2[q][nt][hh];
                        sAcc[(32 + idx) * ST_THREADS + tid] = aL[q][nt][hh];
                    }
        }
        __syncthreads();
        if (kgrp == 1) return;
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int idx = (q * 2 + nt) * 2 + hh;
                    a1[q][nt][hh] += sAcc[(0 + idx) * ST_THREADS + tid];
                    a2[q][nt][hh] += sAcc[(16 + idx) * ST_THREADS + tid];
                    aL[q][nt][hh] += sAcc[(32 + idx) * ST_THREADS + tid];
                }
    }
    // the epilogue's barriers: group 0 only in DB mode (named barrier 1, 128 threads), the whole CTA otherwise
    auto epi_barrier = [&]() {
        if (DB) asm volatile("bar.sync 1, 128;" ::: "memory");
        else __syncthreads();
    };

    // ---- epilogue: the thread's own elements of all four blocks ----
    const int i = r0 + wr + g;
    // prefetched operands (DB): each thread reads back what it copied itself, so wait_group is all the ordering needed
    const double* ep = sPan + (nchunks & 1) * 36 * ST_PANEL;
    if (DB) asm volatile("cp.async.wait_group 0;" ::: "memory");
    double uo[4][2][2], vo[4][2][2];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            const int e = i * ST_LD + c0 + nt * 8 + 2 * t;
            const int sl = q * 2 + nt;
            const double2 a = DB ? *reinterpret_cast<const double2*>(ep + ((0 + sl) * ST_THREADS + tid) * 2)
                                 : __ldcg(reinterpret_cast<const double2*>(recI + (size_t)q * ST_BLK + e));
            const double2 c = DB ? *reinterpret_cast<const double2*>(ep + ((8 + sl) * ST_THREADS + tid) * 2)
                                 : __ldcg(reinterpret_cast<const double2*>(recI + (size_t)(4 + q) * ST_BLK + e));
            uo[q][nt][0] = a.x; uo[q][nt][1] = a.y;
            vo[q][nt][0] = c.x; vo[q][nt][1] = c.y;
        }
    if (fld == 0) st_assemble<true>(sVec, sEdgeP, nu, P.bc_edges != 0, i, c0, t, uo, vo, a1, a2, aL);
    else st_assemble<false>(sVec, sEdgeP + 8 * ST_HP, nu, P.bc_edges != 0, i, c0, t, uo, vo, a1, a2, aL);

    const double adt = rk_alpha(P.stage) * dt_b;
    const bool last = (P.stage == 3);
    double s0 = 0.0, s1 = 0.0, mx = 0.0;   // sum d^2, sum phi_n^2, max |phi_new| of this warp's field
    const double wi = (i == centre) ? 0.5 : 1.0;
    // momentum update of the warp's field; the v warps hand dv/dy to the u warps for the pressure update
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const int j = c0 + nt * 8 + 2 * t;
        const int e = i * ST_LD + j;
        const double w0 = 4.0 * wi * (j == centre ? 0.5 : 1.0), w1 = 4.0 * wi * (j + 1 == centre ? 0.5 : 1.0);
        double fn[4][2];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const size_t off = (size_t)(4 * fld + q) * ST_BLK + e;
            const double2 f0 = DB ? *reinterpret_cast<const double2*>(ep + ((16 + q * 2 + nt) * ST_THREADS + tid) * 2)
                                  : __ldcg(reinterpret_cast<const double2*>(recN + off));
            fn[q][0] = f0.x + adt * aL[q][nt][0];
            fn[q][1] = f0.y + adt * aL[q][nt][1];
            if (last) {
                double d;
                d = fn[q][0] - f0.x; s0 = fma(w0 * d, d, s0); s1 = fma(w0 * f0.x, f0.x, s1);
                d = fn[q][1] - f0.y; s0 = fma(w1 * d, d, s0); s1 = fma(w1 * f0.y, f0.y, s1);
            }
            __stcg(reinterpret_cast<double2*>(recO + off), make_double2(fn[q][0], fn[q][1]));
            if (fld == 1) {
                sX[(tw * 16 + (q * 2 + nt) * 2 + 0) * 32 + lane] = a1[q][nt][0];
                sX[(tw * 16 + (q * 2 + nt) * 2 + 1) * 32 + lane] = a1[q][nt][1];
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
    epi_barrier();
    if (fld == 0) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            const int e = i * ST_LD + c0 + nt * 8 + 2 * t;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const size_t off = (size_t)(8 + q) * ST_BLK + e;
                const double2 p0 = DB ? *reinterpret_cast<const double2*>(ep + ((24 + q * 2 + nt) * ST_THREADS + tid) * 2)
                                      : __ldcg(reinterpret_cast<const double2*>(recN + off));
                const double dv0 = a1[q][nt][0] + sX[(tw * 16 + (q * 2 + nt) * 2 + 0) * 32 + lane];
                const double dv1 = a1[q][nt][1] + sX[(tw * 16 + (q * 2 + nt) * 2 + 1) * 32 + lane];
                __stcg(reinterpret_cast<double2*>(recO + off),
                       make_double2(p0.x + adt * (-beta2 * dv0), p0.y + adt * (-beta2 * dv1)));
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
            sRed[warp * 4 + 0] = s0;
            sRed[warp * 4 + 1] = s1;
            sRed[warp * 4 + 2] = mx;
        }
        epi_barrier();
        if (tid == 0) {
            double* out = P.partials + ((size_t)b * ST_TILES + tile) * 8;
            out[0] = sRed[0] + sRed[4];                 // u: sum d^2   (warps 0, 1)
            out[1] = sRed[1] + sRed[5];                 // u: sum u_n^2
            out[2] = sRed[8] + sRed[12];                // v            (warps 2, 3)
            out[3] = sRed[9] + sRed[13];
            out[4] = nanmax(sRed[2], sRed[6]);
            out[5] = nanmax(sRed[10], sRed[14]);
        }
    }
}

}  // namespace ldc

namespace ldc {

// Start of a run: max|u|, max|v| of the packed state (buffer 0) and the first dt.
__global__ void __launch_bounds__(256) k_stream_begin(const double* __restrict__ buf0, size_t stride, int m,
                                                      const double* __restrict__ Re, const double* __restrict__ beta2,
                                                      const double* __restrict__ CFL, const double* __restrict__ lidvel,
                                                      double dx_min, double dy_min, double* __restrict__ dt,
                                                      CavityStatus* __restrict__ status) {
    __shared__ double sh[2 * 32];
    const int b = blockIdx.x, h = eo_half(m);
    const double* rec = buf0 + (size_t)b * stride;
    double mx[2] = {0.0, 0.0};
    bool isnan_[2] = {false, false};
    for (int e = threadIdx.x; e < h * h; e += blockDim.x) {
        const int o = (e / h) * ST_LD + e % h;
#pragma unroll
        for (int f = 0; f < 2; ++f) {
            const double* B = rec + (size_t)4 * f * ST_BLK + o;
            double r[4];
            eo_to_points(B[0], B[ST_BLK], B[2 * ST_BLK], B[3 * ST_BLK], r);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                isnan_[f] |= (r[q] != r[q]);
                mx[f] = fmax(mx[f], fabs(r[q]));
            }
        }
    }
    const double* ed = rec + 12 * ST_BLK;
    for (int e = threadIdx.x; e < 4 * ST_ME + 4; e += blockDim.x) {
        const int q = e < 4 * ST_ME ? e : 8 * ST_ME + (e - 4 * ST_ME);
        const double a = ed[q], c = ed[q + (e < 4 * ST_ME ? 4 * ST_ME : 4)];
        isnan_[0] |= (a != a);
        isnan_[1] |= (c != c);
        mx[0] = fmax(mx[0], fabs(a));
        mx[1] = fmax(mx[1], fabs(c));
    }
    double flag[2] = {isnan_[0] ? 1.0 : 0.0, isnan_[1] ? 1.0 : 0.0};
    block_reduce<2, true>(mx, sh);
    block_reduce<2, true>(flag, sh);
    if (threadIdx.x == 0) {
        const double um = flag[0] > 0.0 ? nan("") : mx[0], vm = flag[1] > 0.0 ? nan("") : mx[1];
        status[b].umax = um;
        status[b].vmax = vm;
        const double ua = (lidvel[b] > um) ? lidvel[b] : um, va = (1e-10 > vm) ? 1e-10 : vm;
        const double nu = 1.0 / Re[b];
        const double lam_x = (ua + sqrt(ua * ua + beta2[b])) / dx_min + nu / (dx_min * dx_min);
        const double lam_y = (va + sqrt(va * va + beta2[b])) / dy_min + nu / (dy_min * dy_min);
        dt[b] = CFL[b] / (lam_x + lam_y);
    }
}

// End of an RK4 step: one CTA per cavity.  (Folding this into the last stage-3 tile of each cavity was tried: the
// inlined body cost the stage kernel 20 % at every batch size, see profiles/README.md visit 6.)
__global__ void __launch_bounds__(ST_THREADS) k_stream_end(const StreamParams P) {
    __shared__ double sh[32], shp[ST_TILES * 6];
    const int b = blockIdx.x;
    pdl_launch_dependents();
    // per-cavity constants (nothing in the chain writes them): on their way before the wait
    const double lidv = __ldg(P.lidvel + b), b2 = __ldg(P.beta2 + b), nu = 1.0 / __ldg(P.Re + b), cfl = __ldg(P.CFL + b);
    const double tolb = P.tol ? __ldg(P.tol + b) : -1.0;
    pdl_wait();
    if (!P.status[b].active) return;
    stream_end_body(P, b, sh, shp, lidv, b2, nu, cfl, tolb);
}

// After a run: move every cavity's record into buffer 0 so that the next run (or an unpack) finds it there.
__global__ void k_stream_normalise(double* __restrict__ bufs, size_t buf_stride, size_t stride, CavityStatus* __restrict__ status) {
    const int b = blockIdx.y, cur = status[b].pad;
    if (cur == 0) return;
    const double2* src = reinterpret_cast<const double2*>(bufs + cur * buf_stride + (size_t)b * stride);
    double2* dst = reinterpret_cast<double2*>(bufs + (size_t)b * stride);
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < stride / 2; e += (size_t)gridDim.x * blockDim.x)
        dst[e] = src[e];
}
__global__ void k_stream_clear_pad(CavityStatus* __restrict__ status, int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) status[b].pad = 0;
}

}  // namespace ldc
