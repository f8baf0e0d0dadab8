// Device pieces of the even-odd resident kernel (see resident_eo.cuh for the formulation and layouts).
#pragma once
#include "resident_eo.cuh"
#include "resident_kernel.cuh"

namespace ldc {

// Contractions of one velocity field F (four parity blocks, index 2a+b) for the warp's 8 x (8 NT) tile of every
// output block:
//   a1[ab] = D_{a,1-a} F_{1-a,b}                         (d/dx)
//   a2[ab] = F_{a,1-b} D_{b,1-b}^T                       (d/dy)
//   aL[ab] = (nu DD_aa) F_ab + F_ab (nu DD_bb)^T + [ (-G_{a,1-a}) P_{1-a,b} | P_{a,1-b} (-G_{b,1-b})^T ]
// sD/sDD/sG point at two consecutive blocks (index = parity of the OUTPUT rows / columns); DD is pre-scaled
// by nu and G negated in shared memory.  Every F fragment serves a first- and a second-derivative DMMA.
template <int NT, int LD, int BLK, bool P_LEFT>
__device__ __forceinline__ void eo_field_pass(const double* __restrict__ sD, const double* __restrict__ sDD,
                                              const double* __restrict__ sG, const double* __restrict__ sF,
                                              const double* __restrict__ sP, int K4, int r0, int c0, int g, int t,
                                              double (&a1)[4][NT][2], double (&a2)[4][NT][2], double (&aL)[4][NT][2]) {
    const int arow = (r0 + g) * LD + t;   // A fragment: X[r0 + g][k0 + t]
    const int brow = (c0 + g) * LD + t;   // B fragment of a transposed operator: Op[c0 + 8 nt + g][k0 + t]
    const int bfld = t * LD + c0 + g;     // B fragment of a field block: F[k0 + t][c0 + 8 nt + g]
#pragma unroll 1
    for (int k0 = 0; k0 < K4; k0 += 4) {
        double FA[4], FB[4][NT];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            FA[q] = sF[q * BLK + arow + k0];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) FB[q][nt] = sF[q * BLK + bfld + k0 * LD + nt * 8];
        }
        {   // left multiplications: a1 += D_a F_{1-a,b},  aL += DD_a F_ab
            double dA[2], ddA[2];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                dA[a] = sD[a * BLK + arow + k0];
                ddA[a] = sDD[a * BLK + arow + k0];
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
        {   // right multiplications: a2 += F_{a,1-b} D_b^T,  aL += F_ab DD_b^T
            double dB[2][NT], ddB[2][NT];
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    dB[b][nt] = sD[b * BLK + brow + nt * 8 * LD + k0];
                    ddB[b][nt] = sDD[b * BLK + brow + nt * 8 * LD + k0];
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
        if (P_LEFT) {   // aL += (-G_a) P_{1-a,b}
            double gA[2], PB[4][NT];
#pragma unroll
            for (int a = 0; a < 2; ++a) gA[a] = sG[a * BLK + arow + k0];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) PB[q][nt] = sP[q * BLK + bfld + k0 * LD + nt * 8];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt)
                        dmma884(aL[2 * a + b][nt][0], aL[2 * a + b][nt][1], gA[a], PB[2 * (1 - a) + b][nt]);
        } else {        // aL += P_{a,1-b} (-G_b)^T
            double PA[4], gB[2][NT];
#pragma unroll
            for (int q = 0; q < 4; ++q) PA[q] = sP[q * BLK + arow + k0];
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) gB[b][nt] = sG[b * BLK + brow + nt * 8 * LD + k0];
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

// parity blocks (SS, SA, AS, AA) -> the four mirror points (k,l) (k',l) (k,l') (k',l')
__device__ __forceinline__ void eo_to_points(double ss, double sa, double as, double aa, double (&r)[4]) {
    const double a_ = ss + sa, b_ = ss - sa, c_ = as + aa, d_ = as - aa;
    r[0] = a_ + c_;
    r[1] = a_ - c_;
    r[2] = b_ + d_;
    r[3] = b_ - d_;
}
// four mirror points -> 4 x parity blocks (the caller folds the 1/4)
__device__ __forceinline__ void eo_from_points4(const double (&r)[4], double (&q)[4]) {
    const double a_ = r[0] + r[1], b_ = r[0] - r[1], c_ = r[2] + r[3], d_ = r[2] - r[3];
    q[0] = a_ + c_;   // 4 SS
    q[1] = a_ - c_;   // 4 SA
    q[2] = b_ + d_;   // 4 AS
    q[3] = b_ - d_;   // 4 AA
}

// Boundary-line corrections + nonlinear assembly of one momentum equation at the thread's own elements.
// On return Rm = parity blocks of the momentum residual, dF = blocks of dF/dx (IS_U) or dF/dy (!IS_U).
template <int NT, int LD, int BLK, int HP, bool IS_U>
__device__ __forceinline__ void eo_assemble(const double* __restrict__ sVec, const double* __restrict__ sEdgeF,
                                            const double* __restrict__ sU, const double* __restrict__ sV, double nu,
                                            bool bc_lines, int r0, int c0, int g, int t, double (&a1)[4][NT][2],
                                            double (&a2)[4][NT][2], double (&aL)[4][NT][2]) {
    // sVec: d0_S d0_A dN_S dN_A dd0_S dd0_A ddN_S ddN_A ; sEdgeF: W_S W_A E_S E_A S_S S_A N_S N_A (parity forms)
    const int i = r0 + g;
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
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int j = c0 + nt * 8 + 2 * t + hh;
            double dx[4], dy[4];
            if (bc_lines) {
                // the boundary lines are the boundary conditions (always, after the first stage): W = E = S = 0, so only
                // the lid line N contributes, and only to the y-derivatives
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
            // Convective term u dF/dx + v dF/dy directly in parity form.  The four parity blocks are the coefficients of
            // the characters of Z2 x Z2 (mirror in x, mirror in y), so a pointwise product is the group convolution
            //   (f g)_q = sum_b f_{q xor b} g_b :  32 FMAs per element and no transform to the mirror points and back.
            const int o = i * LD + j;
            double uq[4], vq[4], cq[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uq[q] = sU[q * BLK + o];
                vq[q] = sV[q * BLK + o];
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                double c = uq[q] * dx[0];
#pragma unroll
                for (int b = 1; b < 4; ++b) c = fma(uq[q ^ b], dx[b], c);
#pragma unroll
                for (int b = 0; b < 4; ++b) c = fma(vq[q ^ b], dy[b], c);
                cq[q] = c;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                aL[q][nt][hh] -= cq[q];                              // R = lin - conv
                a1[q][nt][hh] = IS_U ? dx[q] : dy[q];                // the divergence part of this field
            }
        }
}

}  // namespace ldc
