// The resident RK4 kernel (see resident.cuh for the formulation and layouts).
#pragma once
#include "resident.cuh"

namespace ldc {

// One velocity field F (u or v) through its five contractions, as 8x8x4 DMMA tiles:
//   a1 = Dx F      a2 = F Dy^T      aL = (nu Dxx) F + F (nu Dyy)^T + [ (-Gx) p  |  p (-Gy)^T ]
// aL collects every term of the momentum residual that is linear in the fields; P_LEFT picks the
// pressure-gradient component (d/dx for the u equation, d/dy for the v equation).  The shared-memory copies
// of Dxx, Dyy are pre-scaled by the cavity's nu = 1/Re and those of Gx, Gy are negated when they are loaded,
// so the k loop is LDS + DMMA only.
// Fragment layout of mma.m8n8k4.f64 (g = lane/4, t = lane%4): A[g][t], B[t][g], C[g][2t], C[g][2t+1].
template <int MT, int NT, int LD, bool P_LEFT>
__device__ __forceinline__ void field_pass(const double* __restrict__ sDa, const double* __restrict__ sDDa,
                                           const double* __restrict__ sDb, const double* __restrict__ sDDb,
                                           const double* __restrict__ sG, const double* __restrict__ sF,
                                           const double* __restrict__ sP, int K4, int row0, int col0, int g,
                                           int t, double (&a1)[MT][NT][2], double (&a2)[MT][NT][2],
                                           double (&aL)[MT][NT][2]) {
    const int arow = (row0 + g) * LD + t;   // A fragment: element [row0 + 8 mt + g][k0 + t]
    const int brow = (col0 + g) * LD + t;   // B fragment of a transposed operator: Op[col0 + 8 nt + g][k0 + t]
    const int bfld = t * LD + col0 + g;     // B fragment of a field: F[k0 + t][col0 + 8 nt + g]
#pragma unroll 1
    for (int k0 = 0; k0 < K4; k0 += 4) {
        double fa[MT], fb[NT];
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) fa[mt] = sF[arow + mt * 8 * LD + k0];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) fb[nt] = sF[bfld + k0 * LD + nt * 8];
        {   // a1 += Dx F
            double d[MT];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) d[mt] = sDa[arow + mt * 8 * LD + k0];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma884(a1[mt][nt][0], a1[mt][nt][1], d[mt], fb[nt]);
        }
        {   // a2 += F Dy^T
            double d[NT];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) d[nt] = sDb[brow + nt * 8 * LD + k0];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma884(a2[mt][nt][0], a2[mt][nt][1], fa[mt], d[nt]);
        }
        {   // aL += (nu Dxx) F
            double d[MT];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) d[mt] = sDDa[arow + mt * 8 * LD + k0];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma884(aL[mt][nt][0], aL[mt][nt][1], d[mt], fb[nt]);
        }
        {   // aL += F (nu Dyy)^T
            double d[NT];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) d[nt] = sDDb[brow + nt * 8 * LD + k0];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma884(aL[mt][nt][0], aL[mt][nt][1], fa[mt], d[nt]);
        }
        if (P_LEFT) {   // aL += (-Gx) p
            double d[MT], q[NT];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) d[mt] = sG[arow + mt * 8 * LD + k0];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) q[nt] = sP[bfld + k0 * LD + nt * 8];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma884(aL[mt][nt][0], aL[mt][nt][1], d[mt], q[nt]);
        } else {        // aL += p (-Gy)^T
            double d[NT], q[MT];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) q[mt] = sP[arow + mt * 8 * LD + k0];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) d[nt] = sG[brow + nt * 8 * LD + k0];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma884(aL[mt][nt][0], aL[mt][nt][1], q[mt], d[nt]);
        }
    }
}

// Boundary-line corrections + nonlinear assembly of one momentum equation at the thread's own elements.
//   R = aL + nu (boundary part of lap) - (u dF/dx + v dF/dy);  ddir returns dF/dx (IS_U) or dF/dy.
template <int MT, int NT, int LD, int MP, bool IS_U>
__device__ __forceinline__ void assemble(const double* __restrict__ sVec, const double* __restrict__ sEdgeF,
                                         const double* __restrict__ sU, const double* __restrict__ sV, double nu,
                                         int row0, int col0, int g, int t, double (&a1)[MT][NT][2],
                                         double (&a2)[MT][NT][2], double (&aL)[MT][NT][2], double (&R)[MT][NT][2],
                                         double (&div)[MT][NT][2]) {
    const double *dW = sVec, *dE = sVec + MP, *ddW = sVec + 2 * MP, *ddE = sVec + 3 * MP;
    const double *dS = sVec + 4 * MP, *dN = sVec + 5 * MP, *ddS = sVec + 6 * MP, *ddN = sVec + 7 * MP;
    const double *fW = sEdgeF, *fE = sEdgeF + MP, *fS = sEdgeF + 2 * MP, *fN = sEdgeF + 3 * MP;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
        const int i = row0 + mt * 8 + g;
        const double rW = dW[i], rE = dE[i], rrW = ddW[i], rrE = ddE[i], vS = fS[i], vN = fN[i];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int j = col0 + nt * 8 + 2 * t + h;
                const double cW = fW[j], cE = fE[j];
                const double dx = a1[mt][nt][h] + rW * cW + rE * cE;
                const double dy = a2[mt][nt][h] + vS * dS[j] + vN * dN[j];
                const double lin = aL[mt][nt][h] + nu * (rrW * cW + rrE * cE + vS * ddS[j] + vN * ddN[j]);
                const double uu = sU[i * LD + j], vv = sV[i * LD + j];
                R[mt][nt][h] = lin - (uu * dx + vv * dy);
                if (IS_U) div[mt][nt][h] = dx;
                else div[mt][nt][h] += dy;
            }
        }
    }
}

// 4 sums + 2 NaN-propagating maxima across the CTA, fixed order (bit-reproducible), result in every thread.
template <int WARPS>
__device__ __forceinline__ void cta_reduce(double (&s)[4], double (&mx)[2], double* sRed, int warp, int lane) {
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
    for (int w = 1; w < WARPS; ++w) {
#pragma unroll
        for (int q = 0; q < 4; ++q) s[q] += sRed[w * 8 + q];
        mx[0] = nanmax(mx[0], sRed[w * 8 + 4]);
        mx[1] = nanmax(mx[1], sRed[w * 8 + 5]);
    }
}

}  // namespace ldc
