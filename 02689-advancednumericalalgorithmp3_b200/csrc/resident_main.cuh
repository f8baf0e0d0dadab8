// Main resident kernel: one CTA per cavity, up to P.n_steps RK4 steps per launch.
#pragma once
#include "resident_kernel.cuh"

namespace ldc {

template <int MP, int WM, int WN>
__global__ void __launch_bounds__(ResCfg<MP, WM, WN>::THREADS, 1) k_resident_rk4(const ResParams P) {
    using C = ResCfg<MP, WM, WN>;
    constexpr int LD = C::LD, MT = C::MT, NT = C::NT, MAT = C::MAT;
    extern __shared__ __align__(16) double smem[];

    const int b = blockIdx.x;
    CavityStatus st = P.status[b];
    if (!st.active) return;  // uniform across the CTA

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

    double* gS = P.state + (size_t)b * P.state_stride;  // packed state = base state phi^n
    double *gU = gS, *gV = gS + MAT, *gP = gS + 2 * MAT, *gEdge = gS + 3 * MAT;

    // ---- load operators, state, edges (16-byte vectors; every block is a multiple of 2 doubles) ----
    {
        const double2* src = reinterpret_cast<const double2*>(P.ops);
        double2* dst = reinterpret_cast<double2*>(sOps);
        const int n2 = (nmat * MAT + 8 * MP) / 2;
        for (int e = tid; e < n2; e += C::THREADS) dst[e] = __ldg(src + e);
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

    const double Re = P.Re[b], nu = 1.0 / Re, beta2 = P.beta2[b], CFL = P.CFL[b], lidvel = P.lidvel[b];
    const double tol = P.tol ? P.tol[b] : -1.0;

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
        // s[0], s[1] carry the maxima of the boundary-condition lines (non-negative, so max == max of sums' parts)
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

    long long step = 0;
    for (; step < P.n_steps && st.active; ++step) {
        // dt from the start-of-step state (spectral/sg.py:387-408, Q9)
        double dt;
        {
            const double ua = (lidvel > umax) ? lidvel : umax;   // Python max(np.max(|u|), lid): NaN first wins
            const double va = (1e-10 > vmax) ? 1e-10 : vmax;
            const double lam_x = (ua + sqrt(ua * ua + beta2)) / P.dx_min + nu / (P.dx_min * P.dx_min);
            const double lam_y = (va + sqrt(va * va + beta2)) / P.dy_min + nu / (P.dy_min * P.dy_min);
            dt = CFL / (lam_x + lam_y);
        }
        // boundary-line parts of ||phi^{n+1} - phi^n||^2 and ||phi^n||^2 (the new lines are the BC lines)
        double s[4] = {0, 0, 0, 0}, mx[2] = {0, 0};
        for (int e = tid; e < MP; e += C::THREADS) {
#pragma unroll
            for (int f = 0; f < 2; ++f) {
                const double* ed = sEdge + f * 4 * MP;
                const double w = ed[e], ea = ed[MP + e], so = ed[2 * MP + e], no = ed[3 * MP + e];
                const double dn = sLid[f * MP + e] - no;
                s[2 * f] += w * w + ea * ea + so * so + dn * dn;
                s[2 * f + 1] += w * w + ea * ea + so * so + no * no;
            }
        }
        if (tid == 0) {
#pragma unroll
            for (int f = 0; f < 2; ++f) {
                const double* c = sEdge + 8 * MP + 4 * f;  // c00 c0N cN0 cNN
                const double d1 = sLid[2 * MP + 2 * f] - c[1], d3 = sLid[2 * MP + 2 * f + 1] - c[3];
                s[2 * f] += c[0] * c[0] + d1 * d1 + c[2] * c[2] + d3 * d3;
                s[2 * f + 1] += c[0] * c[0] + c[1] * c[1] + c[2] * c[2] + c[3] * c[3];
            }
        }

#pragma unroll 1
        for (int stage = 0; stage < 4; ++stage) {
            double Ru[MT][NT][2], Rv[MT][NT][2], dv[MT][NT][2];
            {
                double a1[MT][NT][2], a2[MT][NT][2], aL[MT][NT][2];
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                        for (int h = 0; h < 2; ++h) a1[mt][nt][h] = a2[mt][nt][h] = aL[mt][nt][h] = 0.0;
                field_pass<MT, NT, LD, true>(sDx, sDxx, sDy, sDyy, sGx, sU, sP, nu, P.K4, row0, col0, g, t, a1, a2, aL);
                assemble<MT, NT, LD, MP, true>(sVec, sEdge, sU, sV, nu, row0, col0, g, t, a1, a2, aL, Ru, dv);
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                        for (int h = 0; h < 2; ++h) a1[mt][nt][h] = a2[mt][nt][h] = aL[mt][nt][h] = 0.0;
                field_pass<MT, NT, LD, false>(sDx, sDxx, sDy, sDyy, sGy, sV, sP, nu, P.K4, row0, col0, g, t, a1, a2, aL);
                assemble<MT, NT, LD, MP, false>(sVec, sEdge + 4 * MP, sU, sV, nu, row0, col0, g, t, a1, a2, aL, Rv, dv);
            }
            // stage update phi = phi^n + alpha dt R at the thread's own elements; phi^n comes from L2
            const double adt = rk_alpha(stage) * dt;
            const bool last = (stage == 3);
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const int e = (row0 + mt * 8 + g) * LD + col0 + nt * 8 + 2 * t;
                    const double2 u0 = __ldcg(reinterpret_cast<const double2*>(gU + e));
                    const double2 v0 = __ldcg(reinterpret_cast<const double2*>(gV + e));
                    const double2 p0 = __ldcg(reinterpret_cast<const double2*>(gP + e));
                    const double un0 = u0.x + adt * Ru[mt][nt][0], un1 = u0.y + adt * Ru[mt][nt][1];
                    const double vn0 = v0.x + adt * Rv[mt][nt][0], vn1 = v0.y + adt * Rv[mt][nt][1];
                    const double pn0 = p0.x + adt * (-beta2 * dv[mt][nt][0]), pn1 = p0.y + adt * (-beta2 * dv[mt][nt][1]);
                    if (last) {
                        double d;
                        d = un0 - u0.x; s[0] = fma(d, d, s[0]); s[1] = fma(u0.x, u0.x, s[1]);
                        d = un1 - u0.y; s[0] = fma(d, d, s[0]); s[1] = fma(u0.y, u0.y, s[1]);
                        d = vn0 - v0.x; s[2] = fma(d, d, s[2]); s[3] = fma(v0.x, v0.x, s[3]);
                        d = vn1 - v0.y; s[2] = fma(d, d, s[2]); s[3] = fma(v0.y, v0.y, s[3]);
                        mx[0] = nanmax(nanmax(mx[0], fabs(un0)), fabs(un1));
                        mx[1] = nanmax(nanmax(mx[1], fabs(vn0)), fabs(vn1));
                        __stcg(reinterpret_cast<double2*>(gU + e), make_double2(un0, un1));
                        __stcg(reinterpret_cast<double2*>(gV + e), make_double2(vn0, vn1));
                        __stcg(reinterpret_cast<double2*>(gP + e), make_double2(pn0, pn1));
                    }
                    Ru[mt][nt][0] = un0; Ru[mt][nt][1] = un1;
                    Rv[mt][nt][0] = vn0; Rv[mt][nt][1] = vn1;
                    dv[mt][nt][0] = pn0; dv[mt][nt][1] = pn1;
                }
            __syncthreads();  // every warp is done reading the stage state
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const int e = (row0 + mt * 8 + g) * LD + col0 + nt * 8 + 2 * t;
                    *reinterpret_cast<double2*>(sU + e) = make_double2(Ru[mt][nt][0], Ru[mt][nt][1]);
                    *reinterpret_cast<double2*>(sV + e) = make_double2(Rv[mt][nt][0], Rv[mt][nt][1]);
                    // frozen pressure (SGSolver): stages keep grad(p^n); p itself moves at the last stage
                    if (!P.freeze || last)
                        *reinterpret_cast<double2*>(sP + e) = make_double2(dv[mt][nt][0], dv[mt][nt][1]);
                }
            if (stage == 0) {
                // boundary conditions: W, E, S walls zero, lid line and its two corners (spectral/sg.py:348-385)
                for (int e = tid; e < MP; e += C::THREADS) {
#pragma unroll
                    for (int f = 0; f < 2; ++f) {
                        double* ed = sEdge + f * 4 * MP;
                        ed[e] = 0.0; ed[MP + e] = 0.0; ed[2 * MP + e] = 0.0; ed[3 * MP + e] = sLid[f * MP + e];
                    }
                }
                if (tid < 2) {
                    double* c = sEdge + 8 * MP + 4 * tid;
                    c[0] = 0.0; c[1] = sLid[2 * MP + 2 * tid]; c[2] = 0.0; c[3] = sLid[2 * MP + 2 * tid + 1];
                }
            }
            __syncthreads();
        }

        cta_reduce<C::WARPS>(s, mx, sRed, warp, lane);
        const double ures = sqrt(s[0]) / (sqrt(s[1]) + 1e-12);
        const double vres = sqrt(s[2]) / (sqrt(s[3]) + 1e-12);
        umax = nanmax(mx[0], bcmax_u);
        vmax = nanmax(mx[1], bcmax_v);
        update_status(st, ures, vres, tol, P.warmup, P.max_iter, P.nan_exit);
    }

    if (step > 0) {
        // the boundary lines of the packed state are the BC lines from the first stage on
        for (int e = tid; e < 8 * MP + 8; e += C::THREADS) gEdge[e] = sEdge[e];
    }
    if (tid == 0) {
        st.umax = umax;
        st.vmax = vmax;
        P.status[b] = st;
    }
}

}  // namespace ldc
