// Shared declarations of the ldc_b200 library: handle layout, error plumbing, reduction helpers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/ldc_b200.h"

namespace ldc {

extern thread_local std::string g_last_error;

inline int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define LDC_CUDA(expr)                                                                              \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess)                                                                      \
            return ldc::fail(LDC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));     \
    } while (0)

// RK coefficients of the low-storage 4-stage scheme (spectral/sg.py:430, multigrid/fsg.py:973).
__host__ __device__ inline double rk_alpha(int s) {
    return s == 0 ? 0.25 : (s == 1 ? 1.0 / 3.0 : (s == 2 ? 0.5 : 1.0));
}

// Adaptive pseudo-time step (spectral/sg.py:387-408, multigrid/fsg.py:934-948).
__host__ __device__ inline double adaptive_dt(double umax_abs, double vmax_abs, double lid_velocity, double Re,
                                              double beta2, double CFL, double dx_min, double dy_min) {
    double u_max = umax_abs > lid_velocity ? umax_abs : lid_velocity;
    double v_max = vmax_abs > 1e-10 ? vmax_abs : 1e-10;
    double nu = 1.0 / Re;
    double lam_x = (u_max + sqrt(u_max * u_max + beta2)) / dx_min + nu / (dx_min * dx_min);
    double lam_y = (v_max + sqrt(v_max * v_max + beta2)) / dy_min + nu / (dy_min * dy_min);
    return CFL / (lam_x + lam_y);
}

// Per-cavity iteration status kept on the device.
struct CavityStatus {
    long long iters;   // RK4 steps taken in the current solve
    int active;        // still iterating
    int converged;
    int diverged;
    int pad;
    double u_res, v_res;  // last step's relative changes
    double umax, vmax;    // max |u|, max |v| of the current state (for the next dt)
};

}  // namespace ldc

struct ldc_handle {
    int nx = 0, ny = 0;    // polynomial orders
    int n1x = 0, n1y = 0;  // full-grid points per direction
    int mx = 0, my = 0;    // inner (pressure) points per direction
    int B = 0;
    int device = 0;
    cudaStream_t stream = nullptr;
    int path = LDC_PATH_AUTO;
    bool ops_set = false, cases_set = false;

    // 1-D operators (row-major) and their transposes where the tiled kernels want them
    double *Dx = nullptr, *Dxx = nullptr, *Dy = nullptr, *Dyy = nullptr, *Ix = nullptr, *Iy = nullptr;
    double *DyT = nullptr, *DyyT = nullptr, *IyT = nullptr;  // transposed copies (coalesced right-multiplications)
    double dx_min = 0, dy_min = 0;
    bool square_ops = false;  // Dx == Dy, Dxx == Dyy, Ix == Iy bit for bit

    // per-cavity physics
    double *Re = nullptr, *beta2 = nullptr, *CFL = nullptr, *lidvel = nullptr, *ulid = nullptr, *vlid = nullptr;

    // canonical state (reference layout) and work arrays of the tiled path
    double *u = nullptr, *v = nullptr, *p = nullptr;
    double *u0 = nullptr, *v0 = nullptr, *p0 = nullptr;
    double *Ru = nullptr, *Rv = nullptr, *Rp = nullptr;
    double *ptmp = nullptr, *pfull = nullptr;  // Ix p  and  Ix p Iy^T
    double* dt = nullptr;                      // [B]
    double* tol = nullptr;                     // [B]
    ldc::CavityStatus* status = nullptr;       // [B]
    int* n_active = nullptr;                   // device counter
    int* h_pinned = nullptr;                   // pinned host mirror of n_active

    // resident path: packed operators / state (see resident.cuh)
    int MP = 0, LD = 0;
    double* r_ops = nullptr;
    double* r_state = nullptr;
    size_t r_state_stride = 0;    // doubles per cavity
    int packed_valid = 0;         // which packed record holds the current state: 0 none, 1 r_state, 2 e_state, 3 s_state (buffer 0),
                                  // 4 t_state (buffer 0; stream_tma.cuh)
    // even-odd resident path (resident_eo.cuh)
    int HP = 0;
    double* e_ops = nullptr;
    double* e_state = nullptr;
    size_t e_state_stride = 0;
    // streamed even-odd path (stream_eo.cuh): three rotating copies of the packed state + per-tile partial sums
    double* s_state = nullptr;
    double* s_partials = nullptr;
    double* s_opsc = nullptr;     // per-cavity scaled operator copies
    bool canon_valid = true;      // u, v, p hold the current state
    int* q_data = nullptr;        // work queue of the resident kernel (resident.cuh: CavityQueue)
    unsigned* q_seq = nullptr;
    unsigned* q_ctl = nullptr;
    int num_sms = 0;
    int res_slots = 0;            // CTAs of the resident kernel the chip holds at once

    // streamed path: its own operator pack and record stride (ST_HP blocks), so that a grid the resident even-odd kernel
    // also takes (half size <= 32) can use either
    double* s_ops = nullptr;
    size_t s_stride = 0;
    // persistent TMA-fed streamed path (stream_tma.cuh): three rotating records per cavity in the k4-interleaved tile layout,
    // per-cavity operator panels, team barrier counters + mailboxes
    double* t_state = nullptr;
    double* t_opsc = nullptr;
    double* t_team = nullptr;     // [teams] unsigned counters, then [teams] 64-bit mailboxes
    bool eo_sym = false;          // operators pass the centro-(anti)symmetry test the even-odd kernels rely on
    // per-iteration history of ldc_solve_history: [B][hist_cap][7] records + the quadrature weights
    double* hist = nullptr;
    size_t hist_cap = 0;
    double* Wq = nullptr;
    double* fork = nullptr;       // ldc_solve_sampled: the status (and, on the launch-per-stage path, the state) kept across a record step

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev2 = nullptr, ev3 = nullptr;   // ldc_solve_sampled: the whole call (its segments use ev0 / ev1)
    double last_ms = 0.0;
    int64_t launches = 0;
};
