// Main resident kernel: persistent CTAs pull cavities from a work queue and advance each by `chunk` RK4 steps
// with the stage state and the operators resident in shared memory (see resident.cuh).
#pragma once
#include "resident_kernel.cuh"

namespace ldc {

template <int MP, int WM, int WN>
__global__ void __launch_bounds__(ResCfg<MP, WM, WN>::THREADS, 1) k_resident_rk4(const ResParams P) {
    using C = ResCfg<MP, WM, WN>;
    constexpr int LD = C::LD, MT = C::MT, NT = C::NT, MAT = C::MAT;
    extern __shared__ __align__(16) double smem[];
    __shared__ int s_cavity;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int row0 = (warp / WN) * C::TM, col0 = (warp % WN) * C::TN;

    // ---- shared-memory carve-up ----
    const int nmat = P.square ? 3 : 6;
    double* sOps = smem;
    double* sVec = sOps + nmat * MAT;
    double* sU = sVec + 8 * MP;
    double* sV = sU + MAT;
    double* sP = sV + MAT;
    double* sEdge = sP + MAT;          // uW uE uS uN vW vE vS vN, then 8 corners
    double* sLid = sEdge + 8 * MP + 8; // lid_u[MP], lid_v[MP], corners u0N uNN v0N vNN
    double* sRed = sLid + 2 * MP + 4;
    const double *sDx = sOps, *sDxx = sOps + MAT, *sGx = sOps + 2 * MAT;
    const double *sDy = P.square ? sDx : sOps + 3 * MAT, *sDyy = P.square ? sDxx : sOps + 4 * MAT,
                 *sGy = P.square ? sGx : sOps + 5 * MAT;

    // ---- operators that do not depend on the cavity: Dx, -Gx (Dy, -Gy), boundary vectors ----
    {
        const int ngrp = P.square ? 1 : 2;
        for (int q = 0; q < ngrp; ++q) {
            const double2* src = reinterpret_cast<const double2*>(P.ops + (size_t)q * 3 * MAT);
            double2* dst = reinterpret_cast<double2*>(sOps + (size_t)q * 3 * MAT);
            for (int e = tid; e < MAT / 2; e += C::THREADS) {
                dst[e] = __ldg(src + e);
                const double2 gneg = __ldg(src + MAT + e);   // + MAT double2 = + 2 MAT doubles: third matrix of the group, G
                dst[MAT + e] = make_double2(-gneg.x, -gneg.y);
            }
        }
        const double2* vsrc = reinterpret_cast<const double2*>(P.ops + (size_t)nmat * MAT);
        double2* vdst = reinterpret_cast<double2*>(sVec);
        for (int e = tid; e < 4 * MP; e += C::THREADS) vdst[e] = __ldg(vsrc + e);
    }

    for (;;) {
        __syncthreads();  // everyone is done with the previous cavity (and with s_cavity)
        if (tid == 0) s_cavity = CavityQueue::pop(P);
        __syncthreads();
        const int b = s_cavity;
        if (b < 0) return;

        CavityStatus st = load_status(P.status + b);
        const double Re = P.Re[b], nu = 1.0 / Re, beta2 = P.beta2[b], CFL = P.CFL[b], lidvel = P.lidvel[b];
        const double tol = P.tol ? P.tol[b] : -1.0;
        double* gS = P.state + (size_t)b * P.state_stride;  // packed state = base state phi^n
        double *gU = gS, *gV = gS + MAT, *gP = gS + 2 * MAT, *gEdge = gS + 3 * MAT;

        // ---- nu-scaled second-derivative operators, state, edges, lid ----
        {
            const int ngrp = P.square ? 1 : 2;
            for (int q = 0; q < ngrp; ++q) {
                const double2* src = reinterpret_cast<const double2*>(P.ops + (size_t)(3 * q + 1) * MAT);
                double2* dst = reinterpret_cast<double2*>(sOps + (size_t)(3 * q + 1) * MAT);
                for (int e = tid; e < MAT / 2; e += C::THREADS) {
                    const double2 d = __ldg(src + e);
                    dst[e] = make_double2(nu * d.x, nu * d.y);
                }
            }
            const double2* s2 = reinterpret_cast<const double2*>(gS);
            double2* d2 = reinterpret_cast<double2*>(sU);
            const int m2 = (3 * MAT + 8 * MP + 8) / 2;
            for (int e = tid; e < m2; e += C::THREADS) d2[e] = __ldcg(s2 + e);
            const int m = P.m, n1 = P.n1;
            for (int e = tid; e < MP; e += C::THREADS) {
                sLid[e] = e < m ? P.ulid[(size_t)b * n1 + e + 1] : 0.0;
                sLid[MP + e] = e < m ? P.vlid[(size_t)b * n1 + e + 1] : 0.0;
            }
            if (tid == 0) {
                sLid[2 * MP + 0] = P.ulid[(size_t)b * n1];
                sLid[2 * MP + 1] = P.ulid[(size_t)b * n1 + n1 - 1];
                sLid[2 * MP + 2] = P.vlid[(size_t)b * n1];
                sLid[2 * MP + 3] = P.vlid[(size_t)b * n1 + n1 - 1];
            }
        }
        __syncthreads();

        // ---- max|u|, max|v| of the incoming state and of the boundary-condition lines ----
        double bcmax_u, bcmax_v, umax, vmax;
        {
            double s[4] = {0, 0, 0, 0}, mx[2] = {0, 0};
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int e = (row0 + mt * 8 + g) * LD + col0 + nt * 8 + 2 * t + h;
                        mx[0] = nanmax(mx[0], fabs(sU[e]));
                        mx[1] = nanmax(mx[1], fabs(sV[e]));
                    }
            for (int e = tid; e < 4 * MP + 4; e += C::THREADS) {
                const int q = e < 4 * MP ? e : 8 * MP + (e - 4 * MP);  // edges, then corners of u
                mx[0] = nanmax(mx[0], fabs(sEdge[q]));
                mx[1] = nanmax(mx[1], fabs(sEdge[q + (e < 4 * MP ? 4 * MP : 4)]));
            }
            double bm[2] = {0, 0};
            for (int e = tid; e < MP + 2; e += C::THREADS) {
                const double a = e < MP ? sLid[e] : sLid[2 * MP + (e - MP)];
                const double c = e < MP ? sLid[MP + e] : sLid[2 * MP + 2 + (e - MP)];
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
#include "resident_step.inc"
        }

        if (step > 0) {
            // the boundary lines of the packed state are the BC lines from the first stage on
            for (int e = tid; e < 8 * MP + 8; e += C::THREADS) __stcg(gEdge + e, sEdge[e]);
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
