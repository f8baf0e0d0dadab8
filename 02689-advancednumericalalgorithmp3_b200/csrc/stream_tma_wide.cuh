// Wide-tile shape of the persistent TMA-fed streamed kernel: a CTA owns a 16 x 32 tile (two 16-column tiles) of all four
// parity blocks, so a cavity takes half as many CTAs (8 instead of 16 at N = 128) and each of them does twice the DMMA
// work per RK stage.  The per-stage costs that do not scale with the tile -- the team barrier, the first-chunk latency,
// the fixed part of the epilogue -- are paid once per double the work, and the operand traffic per DMMA drops by a
// quarter (one row panel serves two column tiles).  Used when there are enough cavities to keep the chip busy with the
// smaller teams (ldc_abi.cu: batch >= 1.6 x the number of 16 x 16 teams); few cavities stay on k_stream_tma, which
// spreads each of them over more SMs.
//
// Differences from stream_tma.cuh:
//   * warps: (column tile) x (u | v equation) x (row half); every warp takes ALL k4-steps of a chunk against its own
//     column tile -- no split-K, no accumulator swap;
//   * a chunk slot has three sides, row side A and column sides B0, B1, each one bulk copy issued by its own warp;
//     three slots in the ring;
//   * phi^n at the thread's own elements is read back from the record (L2) instead of being kept in shared memory
//     (no room for it beside the wider ring).
// Layouts, records, pack / unpack kernels, producer logic and the team protocol are those of stream_tma.cuh.
#pragma once
#include "stream_tma.cuh"

namespace ldc {

constexpr int S2W_SLOTS = 3;
constexpr int S2W_SIDE = S2_KC4 * S2_NB * 64;        // doubles per side and chunk
constexpr int S2W_SLOT = 3 * S2W_SIDE;               // A | B0 | B1
constexpr size_t S2W_SMEM_DOUBLES = (size_t)S2W_SLOTS * S2W_SLOT   // chunk ring
                                    + 8 * 512                      // stage input u, v at the tile's own elements
                                    + 8 * S2_HP + 16 * S2_HP       // boundary-column vectors, boundary lines in parity form
                                    + 4 * 16 * 32                  // dv/dy hand-over v warps -> u warps
                                    + 16 * 8 + 80 + 16             // tile partials, reduction scratch, step info
                                    + 2 * S2W_SLOTS + 2;           // mbarriers

struct S2WSmem {
    double *slot, *uv, *vec, *edgep, *x, *part, *red, *step;
    unsigned long long *full, *empty, *stepbar, *teambar;
    int* cavity;
    __device__ explicit S2WSmem(double* smem) {
        slot = smem;
        uv = slot + (size_t)S2W_SLOTS * S2W_SLOT;
        vec = uv + 8 * 512;
        edgep = vec + 8 * S2_HP;
        x = edgep + 16 * S2_HP;
        part = x + 4 * 16 * 32;
        red = part + 16 * 8;
        step = red + 80;
        full = reinterpret_cast<unsigned long long*>(step + 16);
        empty = full + S2W_SLOTS;
        stepbar = empty + S2W_SLOTS;
        teambar = stepbar + 1;
        cavity = reinterpret_cast<int*>(step + 15);
    }
};

// Epilogue of a warp's 8 x 16 part of the tile (both 8-column halves nt = 0, 1): boundary corrections + nonlinear
// assembly, RK stage update, stores of the stage output (row side and column side), the pressure update on the u warps,
// and at the last stage the tile's partial sums.  c0w = first column of the warp's column tile.
__device__ __forceinline__ void s2w_finish(const S2Params& P, const S2Visit& V, const double* __restrict__ sUV,
                                           const double* __restrict__ sVec, double* __restrict__ sEdgeP, double* __restrict__ sX,
                                           double* __restrict__ sRed, const double* __restrict__ recN, double* __restrict__ recO,
                                           int tile, int b, int r0, int c0w, int stage, bool bc_lines, double dt, int ch, int fld,
                                           int tw, int lane, double (&a1)[4][2][2], double (&a2)[4][2][2], double (&aL)[4][2][2]) {
    const int g = lane >> 2, t = lane & 3;
    const int m = P.m, h = eo_half(m), centre = (m & 1) ? h - 1 : -1;
    const int i = r0 + 8 * tw + g, rt = r0 >> 4, ct = c0w >> 4, ir = i & 15;
    const int pos = (ch * 2 + tw) * 32 + lane;
    const double adt = rk_alpha(stage) * dt;
    const bool last = (stage == 3);
    const double wi = (i == centre) ? 0.5 : 1.0;
    double s0 = 0.0, s1 = 0.0, mx = 0.0;   // sum d^2, sum phi_n^2, max |phi_new| of this warp's field
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const int j0 = c0w + 8 * nt + 2 * t;
        // phi^n of the warp's field at the thread's elements of this half (stage 0: it is the stage input, already in
        // sUV): issued here, used after the assembly
        double2 f0[4];
        if (stage != 0) {
            const size_t offR = s2_off(rt, j0 >> 2, S2_BU + 4 * fld, ir, j0 & 3);
#pragma unroll
            for (int q = 0; q < 4; ++q) f0[q] = __ldcg(reinterpret_cast<const double2*>(recN + offR + q * 64));
        }
        double uo[4][2], vo[4][2], b1[4][2], b2[4][2], bL[4][2];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double2 a = *reinterpret_cast<const double2*>(sUV + (((0 + q) * 2 + nt) * 128 + pos) * 2);
            const double2 c = *reinterpret_cast<const double2*>(sUV + (((4 + q) * 2 + nt) * 128 + pos) * 2);
            uo[q][0] = a.x; uo[q][1] = a.y;
            vo[q][0] = c.x; vo[q][1] = c.y;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                b1[q][hh] = a1[q][nt][hh];
                b2[q][hh] = a2[q][nt][hh];
                bL[q][hh] = aL[q][nt][hh];
            }
        }
        if (fld == 0) s2_assemble<true>(sVec, sEdgeP, V.nu, bc_lines, i, j0, uo, vo, b1, b2, bL);
        else s2_assemble<false>(sVec, sEdgeP + 8 * S2_HP, V.nu, bc_lines, i, j0, uo, vo, b1, b2, bL);
        const double w0 = 4.0 * wi * (j0 == centre ? 0.5 : 1.0), w1 = 4.0 * wi * (j0 + 1 == centre ? 0.5 : 1.0);
        double fn[4][2];
        double* outR = recO + s2_off(rt, j0 >> 2, S2_BU + 4 * fld, ir, j0 & 3);
        double* outC = recO + S2_SIDE + s2_off(ct, i >> 2, S2_BU + 4 * fld, j0 & 15, i & 3);   // + q * 64; second column + 4
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double fx = stage == 0 ? (fld == 0 ? uo[q][0] : vo[q][0]) : f0[q].x;
            const double fy = stage == 0 ? (fld == 0 ? uo[q][1] : vo[q][1]) : f0[q].y;
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
            if (fld == 1) {   // dv/dy to the u warp of the same (column tile, row half)
                sX[((ch * 2 + tw) * 16 + nt * 8 + q * 2 + 0) * 32 + lane] = b1[q][0];
                sX[((ch * 2 + tw) * 16 + nt * 8 + q * 2 + 1) * 32 + lane] = b1[q][1];
            } else {          // du/dx stays in the accumulator registers for the pressure update
                a1[q][nt][0] = b1[q][0];
                a1[q][nt][1] = b1[q][1];
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
    // p^n at the u warps' elements: on its way while the warps meet
    double2 p0[4][2];
    if (fld == 0) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            const int j0 = c0w + 8 * nt + 2 * t;
            const size_t offPR = s2_off(rt, j0 >> 2, S2_BP, ir, j0 & 3);
#pragma unroll
            for (int q = 0; q < 4; ++q) p0[q][nt] = __ldcg(reinterpret_cast<const double2*>(recN + offPR + q * 64));
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
    consumer_sync();   // dv/dy of the v warps is in sX (and the warps' partial sums in sRed)
    if (fld == 0) {
        // frozen pressure (SGSolver): stages 0..2 pass p^n on as the next stage's pressure operand (s2_finish)
        const bool keep_pn = P.freeze && stage != 3;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            const int j0 = c0w + 8 * nt + 2 * t;
            double* outR = recO + s2_off(rt, j0 >> 2, S2_BP, ir, j0 & 3);
            double* outC = recO + S2_SIDE + s2_off(ct, i >> 2, S2_BP, j0 & 15, i & 3);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double dv0 = a1[q][nt][0] + sX[((ch * 2 + tw) * 16 + nt * 8 + q * 2 + 0) * 32 + lane];
                const double dv1 = a1[q][nt][1] + sX[((ch * 2 + tw) * 16 + nt * 8 + q * 2 + 1) * 32 + lane];
                const double pn0 = keep_pn ? p0[q][nt].x : p0[q][nt].x + adt * (-V.beta2 * dv0);
                const double pn1 = keep_pn ? p0[q][nt].y : p0[q][nt].y + adt * (-V.beta2 * dv1);
                __stcg(reinterpret_cast<double2*>(outR + q * 64), make_double2(pn0, pn1));
                __stcg(outC + q * 64, pn0);
                __stcg(outC + q * 64 + 4, pn1);
            }
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
        // warps: w = 4 ch + 2 fld + tw.  u: warps 0, 1, 4, 5; v: warps 2, 3, 6, 7 -- always in this order
        double* out = P.partials + ((size_t)b * ST_TILES + tile) * 8;
        __stcg(out + 0, ((sRed[0 * 4 + 0] + sRed[1 * 4 + 0]) + sRed[4 * 4 + 0]) + sRed[5 * 4 + 0]);
        __stcg(out + 1, ((sRed[0 * 4 + 1] + sRed[1 * 4 + 1]) + sRed[4 * 4 + 1]) + sRed[5 * 4 + 1]);
        __stcg(out + 2, ((sRed[2 * 4 + 0] + sRed[3 * 4 + 0]) + sRed[6 * 4 + 0]) + sRed[7 * 4 + 0]);
        __stcg(out + 3, ((sRed[2 * 4 + 1] + sRed[3 * 4 + 1]) + sRed[6 * 4 + 1]) + sRed[7 * 4 + 1]);
        __stcg(out + 4, nanmax(nanmax(sRed[0 * 4 + 2], sRed[1 * 4 + 2]), nanmax(sRed[4 * 4 + 2], sRed[5 * 4 + 2])));
        __stcg(out + 5, nanmax(nanmax(sRed[2 * 4 + 2], sRed[3 * 4 + 2]), nanmax(sRed[6 * 4 + 2], sRed[7 * 4 + 2])));
    }
}


// ================================================ consumer warps ================================================
__device__ __forceinline__ void s2w_consumer_main(const S2Params& P, const S2WSmem& S, const S2Geo& G) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ch = warp >> 2, fld = (warp >> 1) & 1, tw = warp & 1, g = lane >> 2, t = lane & 3;
    const int ro = (8 * tw + g) * 4 + t, co = g * 4 + t;
    const int m = G.m, h = G.h, n1 = G.n1, nch = G.nch;
    const int c0w = G.c0 + 16 * ch;                       // first column of the warp's column tile
    const int boff = (1 + ch) * S2W_SIDE;                 // its column side in a chunk slot
    const int cgrab = c0w >> 3;                           // chunks cgrab, cgrab + 1 hold the tile's columns as k
    const int pos = (ch * 2 + tw) * 32 + lane;
    unsigned gcb = 0;        // chunks consumed before the current stage (ring position)
    unsigned nstepinfo = 0;  // step-info hand-overs consumed so far

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
#pragma unroll
                for (int nt = 0; nt < 2; ++nt)
                    if (cgrab + nt >= nch) {   // these columns lie in the zero padding
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            *reinterpret_cast<double2*>(S.uv + (((4 * fld + q) * 2 + nt) * 128 + pos) * 2) = make_double2(0.0, 0.0);
                    }
                for (int c = 0; c < nch; ++c) {
                    const unsigned gc = gcb + c, slot = gc % S2W_SLOTS;
                    mbar_wait(S.full + slot, (gc / S2W_SLOTS) & 1);
                    const double* sl = S.slot + (size_t)slot * S2W_SLOT;
                    if (c == cgrab || c == cgrab + 1) {
                        // the thread's own elements of its field's stage input: columns 8 nt + 2 t, + 1 of the column tile
                        const int nt = c - cgrab;
                        const double* src = sl + ((t >> 1) * S2_NB + S2_BU + 4 * fld) * 64 + (8 * tw + g) * 4 + 2 * (t & 1);
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            *reinterpret_cast<double2*>(S.uv + (((4 * fld + q) * 2 + nt) * 128 + pos) * 2) =
                                *reinterpret_cast<const double2*>(src + q * 64);
                    }
                    const int ksteps = min(S2_KC4, G.nk4 - c * S2_KC4);
                    if (fld == 0) s2_chunk<true>(sl, boff, 0, ksteps, ro, co, fld, a1, a2, aL);
                    else s2_chunk<false>(sl, boff, 0, ksteps, ro, co, fld, a1, a2, aL);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(S.empty + slot);
                }
                gcb += nch;
                if (stage == 0 && step > 0) {
                    // the reduction of the previous step (producer warp): dt of this step, and whether there is one
                    mbar_wait(S.stepbar, nstepinfo & 1);
                    ++nstepinfo;
                    dt = S.step[0];
                    if (S.step[1] == 0.0) {   // the cavity went inactive with the previous step: this stage was speculative
                        aborted = true;
                        break;
                    }
                }
                consumer_sync();   // the stage input picked by the u and v warps is complete in S.uv
                s2w_finish(P, V, S.uv, S.vec, S.edgep, S.x, S.red, recN, recO, G.tile, b, G.r0, c0w, stage, bc_lines, dt, ch, fld, tw,
                           lane, a1, a2, aL);
                bc_lines = true;
                // ---- arrive at the team's stage barrier: this CTA's stage output is complete (see s2_consumer_main) ----
                consumer_sync();
                if (tid == 0) red_release_add(P.team_bar + G.team, 1u);
                if (stage == 3) cur = (cur + 2) % 3;
            }
        }
    }
}

// The wide persistent kernel: same launch protocol as k_stream_tma (cooperative, teams of ntile x ceil(ntile / 2) CTAs).
__global__ void __launch_bounds__(S2_THREADS, 1) k_stream_tma_wide(const S2Params P) {
    extern __shared__ __align__(16) double smem[];
    const S2WSmem S(smem);
    const S2Geo G(P, 2);
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        for (int s = 0; s < S2W_SLOTS; ++s) {
            mbar_init(S.full + s, 3);    // one arrival (with its byte count) per side
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
        s2_producer_main<true>(P, S, G, warp - 8);
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
        s2w_consumer_main(P, S, G);
    }
}

}  // namespace ldc
