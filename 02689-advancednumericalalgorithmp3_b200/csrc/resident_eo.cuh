// Even-odd (parity) resident path: layouts and pack / unpack kernels.
//
// Gauss-Lobatto nodes are symmetric about the centre, so D is centro-antisymmetric and Dxx, Ix centro-symmetric
// (to the rounding of the reference's numerically built matrices; the parity form keeps the symmetric part --
// measured effect on converged fields <= 3e-14, scripts/exp_symmetrised.py).  With m interior points,
// k' = m-1-k, h = ceil(m/2):
//     S[k] = (f[k] + f[k'])/2,  A[k] = (f[k] - f[k'])/2   (k < h;  the centre has A = 0, S = f)
// a field F becomes four h x h blocks F_ab (a: x parity, b: y parity; 0 = S, 1 = A) and every m x m contraction of
// the residual becomes two h x h ones per output block: half the DMMA work of the plain resident kernel
//     (Dx F)_ab = D_{a,1-a} F_{1-a,b}       (F Dy^T)_ab = F_{a,1-b} D_{b,1-b}^T
//     (Dxx F)_ab = DD_{aa} F_ab             (F Dyy^T)_ab = F_ab DD_{bb}^T
// The host-side statement of the same algebra (pinned by tests/test_parity.py) is tests/parity_model.py (test infrastructure, not shipped).
#pragma once
#include "resident.cuh"

namespace ldc {

__host__ __device__ inline int eo_half(int m) { return (m + 1) / 2; }
__host__ inline int eo_pick_hp(int m) {
    const int h = eo_half(m);
    if (h <= 8) return 8;
    if (h <= 16) return 16;
    if (h <= 24) return 24;
    if (h <= 32) return 32;
    return 0;
}
__host__ __device__ inline int eo_ld(int HP) { return HP + 4; }  // = 4 mod 8: conflict-free 64-bit fragment loads
// packed operators: 6 blocks HP x LD [D_SA, D_AS, DD_SS, DD_AA, G_SA, G_AS] then 8 vectors of HP
// [d0_S, d0_A, dN_S, dN_A, dd0_S, dd0_A, ddN_S, ddN_A] (columns 0 / N of D and DD at the interior rows)
__host__ __device__ inline size_t eo_ops_size(int HP) { return (size_t)6 * HP * eo_ld(HP) + 8 * HP; }
// packed state per cavity: 12 blocks [u_SS u_SA u_AS u_AA v.. p..], then the boundary lines in real space
// uW uE uS uN vW vE vS vN (2 HP each, interior-indexed, zero padded), then 8 corners
__host__ __device__ inline size_t eo_state_stride(int HP) { return (size_t)12 * HP * eo_ld(HP) + 16 * HP + 8; }

// parity block (a = row parity, b = column parity) of the m x m matrix M[i][k] = src[(i + ro) * ld + k + co]
__device__ __forceinline__ double eo_op_entry(const double* __restrict__ src, int ld, int ro, int co, int m, int i, int k,
                                              int a, int b) {
    const int ip = m - 1 - i, kp = m - 1 - k;
    const double sa = a ? -1.0 : 1.0, sb = b ? -1.0 : 1.0;
    // row transform: (M[i,:] + sa M[i',:]) / 2 ; column transform: X[:,k] + sb X[:,k'] (k' != k), X[:,k] or 0 (k' == k)
    const double x_k = 0.5 * (src[(i + ro) * ld + k + co] + sa * src[(ip + ro) * ld + k + co]);
    if (kp == k) return b ? 0.0 : x_k;
    const double x_kp = 0.5 * (src[(i + ro) * ld + kp + co] + sa * src[(ip + ro) * ld + kp + co]);
    return x_k + sb * x_kp;
}

// D (n1 x n1), DD (n1 x n1), DI = D I (n1 x m) -> parity blocks + boundary-column vectors
__global__ void k_pack_ops_eo(const double* __restrict__ D, const double* __restrict__ DD, const double* __restrict__ DI,
                              double* __restrict__ out, int n1, int HP) {
    const int LD = eo_ld(HP), m = n1 - 2, h = eo_half(m), BLK = HP * LD;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < BLK) {
        const int i = e / LD, k = e % LD;
        const bool in = (i < h && k < h);
        out[0 * BLK + e] = in ? eo_op_entry(D, n1, 1, 1, m, i, k, 0, 1) : 0.0;   // D_SA
        out[1 * BLK + e] = in ? eo_op_entry(D, n1, 1, 1, m, i, k, 1, 0) : 0.0;   // D_AS
        out[2 * BLK + e] = in ? eo_op_entry(DD, n1, 1, 1, m, i, k, 0, 0) : 0.0;  // DD_SS
        out[3 * BLK + e] = in ? eo_op_entry(DD, n1, 1, 1, m, i, k, 1, 1) : 0.0;  // DD_AA
        out[4 * BLK + e] = in ? eo_op_entry(DI, m, 1, 0, m, i, k, 0, 1) : 0.0;   // G_SA
        out[5 * BLK + e] = in ? eo_op_entry(DI, m, 1, 0, m, i, k, 1, 0) : 0.0;   // G_AS
    }
    if (e < HP) {
        double* vec = out + 6 * BLK;
        const bool in = e < h;
        const int i = e + 1, ip = m - 1 - e + 1, N = n1 - 1;
        const double a0 = in ? D[i * n1] : 0.0, b0 = in ? D[ip * n1] : 0.0;
        const double aN = in ? D[i * n1 + N] : 0.0, bN = in ? D[ip * n1 + N] : 0.0;
        const double c0 = in ? DD[i * n1] : 0.0, e0 = in ? DD[ip * n1] : 0.0;
        const double cN = in ? DD[i * n1 + N] : 0.0, eN = in ? DD[ip * n1 + N] : 0.0;
        vec[0 * HP + e] = 0.5 * (a0 + b0);
        vec[1 * HP + e] = 0.5 * (a0 - b0);
        vec[2 * HP + e] = 0.5 * (aN + bN);
        vec[3 * HP + e] = 0.5 * (aN - bN);
        vec[4 * HP + e] = 0.5 * (c0 + e0);
        vec[5 * HP + e] = 0.5 * (c0 - e0);
        vec[6 * HP + e] = 0.5 * (cN + eN);
        vec[7 * HP + e] = 0.5 * (cN - eN);
    }
}

// canonical (u, v, p) -> parity-packed state
__global__ void k_pack_state_eo(const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ p,
                                double* __restrict__ st, size_t stride, int n1, int HP) {
    const int b = blockIdx.y, LD = eo_ld(HP), m = n1 - 2, h = eo_half(m), BLK = HP * LD, N = n1 - 1;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    double* s = st + (size_t)b * stride;
    const double *ub = u + (size_t)b * n1 * n1, *vb = v + (size_t)b * n1 * n1, *pb = p + (size_t)b * m * m;
    if (e < BLK) {
        const int i = e / LD, k = e % LD;
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
            double* o = s + (size_t)4 * f * BLK + e;
            o[0 * BLK] = 0.25 * ((q0 + q1) + (q2 + q3));   // SS
            o[1 * BLK] = 0.25 * ((q0 + q1) - (q2 + q3));   // SA
            o[2 * BLK] = 0.25 * ((q0 - q1) + (q2 - q3));   // AS
            o[3 * BLK] = 0.25 * ((q0 - q1) - (q2 - q3));   // AA
        }
    }
    double* ed = s + (size_t)12 * BLK;
    const int ME = 2 * HP;
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

// parity-packed state -> canonical (u, v, p)
__global__ void k_unpack_state_eo(double* __restrict__ u, double* __restrict__ v, double* __restrict__ p,
                                  const double* __restrict__ st, size_t stride, int n1, int HP) {
    const int b = blockIdx.y, LD = eo_ld(HP), m = n1 - 2, BLK = HP * LD, N = n1 - 1, ME = 2 * HP;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n1 * n1) return;
    const double* s = st + (size_t)b * stride;
    const double* ed = s + (size_t)12 * BLK;
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
        const int o = r * LD + c;
        double f[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const double* B = s + (size_t)4 * q * BLK + o;
            f[q] = (B[0] + sb * B[BLK]) + sa * (B[2 * BLK] + sb * B[3 * BLK]);
        }
        uu = f[0];
        vv = f[1];
        p[(size_t)b * m * m + ii * m + jj] = f[2];
    }
    u[(size_t)b * n1 * n1 + e] = uu;
    v[(size_t)b * n1 * n1 + e] = vv;
}

template <int HP, int NT>
struct EoCfg {
    static constexpr int LD = HP + 4, BLK = HP * LD;
    static constexpr int WR = HP / 8, WC = HP / (8 * NT), WARPS = WR * WC, THREADS = WARPS * 32;
    static constexpr int ME = 2 * HP;
    static_assert(HP % (8 * NT) == 0, "warp tile is 8 x (8 NT)");
    __host__ __device__ static constexpr size_t smem_doubles() {
        return (size_t)6 * BLK + 8 * HP     // operators + boundary-column vectors
               + 12 * BLK                   // stage state u, v, p in parity blocks
               + 8 * ME + 8                 // boundary lines (real space) + corners
               + 16 * HP                    // boundary lines in parity form (u: W E S N x {S,A}; v likewise)
               + 2 * ME + 4                 // lid line (real space) + lid corners
               + 4 * HP                     // lid line in parity form (u_S u_A v_S v_A)
               + 8 * WARPS;                 // reduction scratch
    }
};

}  // namespace ldc
