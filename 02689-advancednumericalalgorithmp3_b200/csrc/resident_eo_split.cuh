// Even-odd resident kernel, equation-split variant for small grids (HP <= 24, i.e. N <= 48).
//
// Same formulation, layouts, queue and results as k_resident_eo (resident_eo_main.cuh), tuned for LATENCY instead of
// register reuse: a single cavity at N = 30 occupies only 4 warps of k_resident_eo<16, 1> (one 8 x 8 tile of the four
// parity blocks per warp, both momentum equations in sequence), i.e. one warp per SM sub-partition issuing 160 serial
// DMMAs per RK stage.  Here the u equation and the v equation of a tile go to different warps (2 x WR x WC warps), so
// the DMMA chain per warp halves and twice as many warps hide the epilogue latencies; the divergence contribution of
// the v warps reaches the u warps (which own the pressure update) through shared memory, piggy-backed on the barrier
// that already separates "stage state read" from "stage state written".  The base state phi^n also stays in shared
// memory (it fits for HP <= 24), so the stage update reads no L2 and the state goes back to L2 once per queue visit.
#pragma once
#include "resident_eo_kernel.cuh"

namespace ldc {

template <int HP>
struct EoSplitCfg {
    static constexpr int NT = 1;
    static constexpr int LD = HP + 4, BLK = HP * LD;
    static constexpr int WR = HP / 8, WC = HP / 8, TW = WR * WC, WARPS = 2 * TW, THREADS = WARPS * 32;
    static constexpr int ME = 2 * HP;
    __host__ __device__ static constexpr size_t smem_doubles() {
        return (size_t)6 * BLK + 8 * HP     // operators + boundary-column vectors
               + 12 * BLK                   // stage state u, v, p in parity blocks
               + 12 * BLK                   // base state phi^n
               + 8 * ME + 8                 // boundary lines (real space) + corners
               + 16 * HP                    // boundary lines in parity form
               + 2 * ME + 4                 // lid line (real space) + lid corners
               + 4 * HP                     // lid line in parity form
               + TW * 8 * 32                // divergence exchange: v warps -> u warps
               + 8 * WARPS;                 // reduction scratch
    }
};

template <int HP>
__global__ void __launch_bounds__(EoSplitCfg<HP>::THREADS, 1) k_resident_eo_split(const ResParams P) {
    using C = EoSplitCfg<HP>;
    constexpr int NT = 1, LD = C::LD, BLK = C::BLK, ME = C::ME;
    extern __shared__ __align__(16) double smem[];
    __shared__ int s_cavity;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int fld = warp / C::TW, tw = warp % C::TW;   // fld 0: u equation (+ pressure), 1: v equation
    const int r0 = (tw / C::WC) * 8, c0 = (tw % C::WC) * 8;
    const int m = P.m, h = eo_half(m);
    const int centre = (m & 1) ? h - 1 : -1;

    double* sOps = smem;
    double* sVec = sOps + 6 * BLK;
    double* sU = sVec + 8 * HP;
    double* sV = sU + 4 * BLK;
    double* sP = sV + 4 * BLK;
    double* sBase = sP + 4 * BLK;          // u^n v^n p^n, 4 blocks each
    double* sEdge = sBase + 12 * BLK;
    double* sEdgeP = sEdge + 8 * ME + 8;
    double* sLid = sEdgeP + 16 * HP;
    double* sLidP = sLid + 2 * ME + 4;
    double* sX = sLidP + 4 * HP;           // [TW][q 4][hh 2][lane 32]
    double* sRed = sX + C::TW * 8 * 32;
    const double *sD = sOps, *sDD = sOps + 2 * BLK, *sG = sOps + 4 * BLK;

    for (int e = tid; e < 2 * BLK; e += C::THREADS) {
        sOps[e] = __ldg(P.ops + e);
        sOps[4 * BLK + e] = -__ldg(P.ops + 4 * BLK + e);
    }
    for (int e = tid; e < 8 * HP; e += C::THREADS) sVec[e] = __ldg(P.ops + 6 * BLK + e);

    for (;;) {
        __syncthreads();
        if (tid == 0) s_cavity = CavityQueue::pop(P);
        __syncthreads();
        const int b = s_cavity;
        if (b < 0) return;

        CavityStatus st = load_status(P.status + b);
        const double Re = P.Re[b], nu = 1.0 / Re, beta2 = P.beta2[b], CFL = P.CFL[b], lidvel = P.lidvel[b];
        const double tol = P.tol ? P.tol[b] : -1.0;
        double* gS = P.state + (size_t)b * P.state_stride;
        double* gEdge = gS + 12 * BLK;

        for (int e = tid; e < 2 * BLK; e += C::THREADS) sOps[2 * BLK + e] = nu * __ldg(P.ops + 2 * BLK + e);
        {
            const double2* s2 = reinterpret_cast<const double2*>(gS);
            double2* d2 = reinterpret_cast<double2*>(sU);
            double2* b2 = reinterpret_cast<double2*>(sBase);
            for (int e = tid; e < 12 * BLK / 2; e += C::THREADS) {
                const double2 x = __ldcg(s2 + e);
                d2[e] = x;
                b2[e] = x;
            }
            double2* e2 = reinterpret_cast<double2*>(sEdge);
            for (int e = tid; e < (8 * ME + 8) / 2; e += C::THREADS) e2[e] = __ldcg(s2 + 12 * BLK / 2 + e);
            const int n1 = P.n1;
            for (int e = tid; e < ME; e += C::THREADS) {
                sLid[e] = e < m ? P.ulid[(size_t)b * n1 + e + 1] : 0.0;
                sLid[ME + e] = e < m ? P.vlid[(size_t)b * n1 + e + 1] : 0.0;
            }
            if (tid == 0) {
                sLid[2 * ME + 0] = P.ulid[(size_t)b * n1];
                sLid[2 * ME + 1] = P.ulid[(size_t)b * n1 + n1 - 1];
                sLid[2 * ME + 2] = P.vlid[(size_t)b * n1];
                sLid[2 * ME + 3] = P.vlid[(size_t)b * n1 + n1 - 1];
            }
        }
        __syncthreads();
        bool bc_lines;
        {
            int dirty = 0;
            for (int e = tid; e < ME; e += C::THREADS) {
#pragma unroll
                for (int f = 0; f < 2; ++f) {
                    const double* ed = sEdge + f * 4 * ME;
                    dirty |= (ed[e] != 0.0) | (ed[ME + e] != 0.0) | (ed[2 * ME + e] != 0.0) | (ed[3 * ME + e] != sLid[f * ME + e]);
                }
            }
            bc_lines = !__syncthreads_or(dirty);
        }
        for (int e = tid; e < 10 * HP; e += C::THREADS) {
            const int line = e / HP, k = e % HP;
            const double* src = line < 8 ? sEdge + line * ME : sLid + (line - 8) * ME;
            double S = 0.0, A = 0.0;
            if (k < h) {
                const double x = src[k], y = src[m - 1 - k];
                S = 0.5 * (x + y);
                A = 0.5 * (x - y);
            }
            double* dst = line < 8 ? sEdgeP + 2 * line * HP : sLidP + 2 * (line - 8) * HP;
            dst[k] = S;
            dst[HP + k] = A;
        }
        __syncthreads();

        // ---- max|u|, max|v| of the incoming state (each warp looks at its own field) and of the BC lines ----
        double bcmax_u, bcmax_v, umax, vmax;
        {
            double s[4] = {0, 0, 0, 0}, mx[2] = {0, 0};
            const double* sF = fld ? sV : sU;
            double fmx = 0.0;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int o = (r0 + g) * LD + c0 + 2 * t + hh;
                double fr[4];
                eo_to_points(sF[o], sF[BLK + o], sF[2 * BLK + o], sF[3 * BLK + o], fr);
#pragma unroll
                for (int q = 0; q < 4; ++q) fmx = nanmax(fmx, fabs(fr[q]));
            }
            if (fld == 0) mx[0] = fmx;
            else mx[1] = fmx;
            for (int e = tid; e < 4 * ME + 4; e += C::THREADS) {
                const int q = e < 4 * ME ? e : 8 * ME + (e - 4 * ME);
                mx[0] = nanmax(mx[0], fabs(sEdge[q]));
                mx[1] = nanmax(mx[1], fabs(sEdge[q + (e < 4 * ME ? 4 * ME : 4)]));
            }
            double bm[2] = {0, 0};
            for (int e = tid; e < ME + 2; e += C::THREADS) {
                const double a = e < ME ? sLid[e] : sLid[2 * ME + (e - ME)];
                const double c = e < ME ? sLid[ME + e] : sLid[2 * ME + 2 + (e - ME)];
                bm[0] = nanmax(bm[0], fabs(a));
                bm[1] = nanmax(bm[1], fabs(c));
            }
            cta_reduce<C::WARPS>(s, mx, sRed, warp, lane);
            umax = mx[0];
            vmax = mx[1];
            __syncthreads();
            cta_reduce<C::WARPS>(s, bm, sRed, warp, lane);
            bcmax_u = bm[0];
            bcmax_v = bm[1];
            __syncthreads();
        }

        long long nsteps = P.launch_end - st.iters;
        if (nsteps > P.chunk) nsteps = P.chunk;
        long long step = 0;
        for (; step < nsteps && st.active; ++step) {
#include "resident_eo_split_step.inc"
        }

        if (step > 0) {
            // the new base state and its boundary lines go back to L2 once per visit
            const double2* b2 = reinterpret_cast<const double2*>(sBase);
            double2* g2 = reinterpret_cast<double2*>(gS);
            for (int e = tid; e < 12 * BLK / 2; e += C::THREADS) __stcg(g2 + e, b2[e]);
            for (int e = tid; e < 8 * ME + 8; e += C::THREADS) __stcg(gEdge + e, sEdge[e]);
        }
        if (tid == 0) {
            st.umax = umax;
            st.vmax = vmax;
            store_status(P.status + b, st);
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            if (st.active && st.iters < P.launch_end) CavityQueue::push(P, b);
            else CavityQueue::retire(P);
        }
    }
}

}  // namespace ldc
