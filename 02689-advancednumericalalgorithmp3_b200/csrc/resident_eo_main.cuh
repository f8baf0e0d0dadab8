// Even-odd resident kernel: persistent CTAs + cavity work queue (as k_resident_rk4), state and operators in
// parity blocks, half the DMMA work per residual evaluation.
#pragma once
#include "resident_eo_kernel.cuh"

namespace ldc {

template <int HP, int NT>
__global__ void __launch_bounds__(EoCfg<HP, NT>::THREADS, 1) k_resident_eo(const ResParams P) {
    using C = EoCfg<HP, NT>;
    constexpr int LD = C::LD, BLK = C::BLK, ME = C::ME;
    extern __shared__ __align__(16) double smem[];
    __shared__ int s_cavity;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int r0 = (warp / C::WC) * 8, c0 = (warp % C::WC) * 8 * NT;
    const int m = P.m, h = eo_half(m);
    const int centre = (m & 1) ? h - 1 : -1;   // index of the self-mirrored line, if any

    // ---- shared-memory carve-up ----
    double* sOps = smem;                 // D_SA D_AS | DD_SS DD_AA (x nu) | G_SA G_AS (negated)
    double* sVec = sOps + 6 * BLK;       // 8 x HP
    double* sU = sVec + 8 * HP;          // 4 blocks each
    double* sV = sU + 4 * BLK;
    double* sP = sV + 4 * BLK;
    double* sEdge = sP + 4 * BLK;        // real space: uW uE uS uN vW vE vS vN (ME each), 8 corners
    double* sEdgeP = sEdge + 8 * ME + 8; // parity forms: u: W_S W_A E_S E_A S_S S_A N_S N_A, then v
    double* sLid = sEdgeP + 16 * HP;     // lid_u[ME], lid_v[ME], corners u0N uNN v0N vNN
    double* sLidP = sLid + 2 * ME + 4;   // u_S u_A v_S v_A
    double* sRed = sLidP + 4 * HP;
    const double *sD = sOps, *sDD = sOps + 2 * BLK, *sG = sOps + 4 * BLK;

    // ---- cavity-independent operators: D blocks, -G blocks, boundary-column vectors ----
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
        double* gS = P.state + (size_t)b * P.state_stride;   // packed state = base state phi^n
        double* gEdge = gS + 12 * BLK;

        // ---- nu-scaled second-derivative blocks, state, boundary lines, lid ----
        for (int e = tid; e < 2 * BLK; e += C::THREADS) sOps[2 * BLK + e] = nu * __ldg(P.ops + 2 * BLK + e);
        {
            const double2* s2 = reinterpret_cast<const double2*>(gS);
            double2* d2 = reinterpret_cast<double2*>(sU);
            for (int e = tid; e < (12 * BLK + 8 * ME + 8) / 2; e += C::THREADS) d2[e] = __ldcg(s2 + e);
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
        // do the stored boundary lines already equal the boundary conditions (bitwise)?  True for every visit but the
        // first one after a load from a general state (cold start is BC-clean too).
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
        // parity forms of the eight boundary lines and of the lid line
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

        // ---- max|u|, max|v| of the incoming state and of the boundary-condition lines ----
        double bcmax_u, bcmax_v, umax, vmax;
        {
            double s[4] = {0, 0, 0, 0}, mx[2] = {0, 0};
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int o = (r0 + g) * LD + c0 + nt * 8 + 2 * t + hh;
                    double ur[4], vr[4];
                    eo_to_points(sU[o], sU[BLK + o], sU[2 * BLK + o], sU[3 * BLK + o], ur);
                    eo_to_points(sV[o], sV[BLK + o], sV[2 * BLK + o], sV[3 * BLK + o], vr);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        mx[0] = nanmax(mx[0], fabs(ur[q]));
                        mx[1] = nanmax(mx[1], fabs(vr[q]));
                    }
                }
            for (int e = tid; e < 4 * ME + 4; e += C::THREADS) {
                const int q = e < 4 * ME ? e : 8 * ME + (e - 4 * ME);  // edges, then corners of u
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
#include "resident_eo_step.inc"
        }

        if (step > 0) {
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
