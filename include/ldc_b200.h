/*
 * ldc_b200.h -- C ABI of the B200-native batched spectral lid-driven-cavity stepper.
 *
 * The reference (PN-CourseWork/02689-AdvancedNumericalAlgorithmP3) is pure Python and has no
 * native boundary; this ABI is what a native binding for its spectral hot path binds to.  Each
 * entry point names the reference code it replaces (paths relative to src/solvers/ of the
 * reference).  All arrays are FP64, C-order.  Per cavity: u, v are (nx+1) x (ny+1) with x the slow
 * axis (meshgrid indexing="ij", spectral/sg.py:108); p is the (nx-1) x (ny-1) inner grid
 * (PN-PN-2, datastructures.py:282-353).  Batched arrays put the cavity index first.
 *
 * Conventions
 *   - every function returns 0 (LDC_OK) or a positive error code; ldc_last_error() gives the text
 *     of the last failure on the calling thread.  No C++ exception crosses the boundary.
 *   - data pointers may be host or device pointers (detected with cudaPointerGetAttributes);
 *     copies are enqueued on the handle's stream and host pointers are synchronised before return.
 *   - the caller owns every buffer it passes; the handle owns the device state.
 *   - one handle <-> one device + one stream + one polynomial order; a handle is not thread-safe,
 *     distinct handles are independent.  There is no CPU fallback: every call needs a CUDA device.
 */
#ifndef LDC_B200_H
#define LDC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ldc_handle ldc_t;

enum {
    LDC_OK = 0,
    LDC_ERR_INVALID = 1,     /* bad argument */
    LDC_ERR_CUDA = 2,        /* CUDA runtime failure (text in ldc_last_error) */
    LDC_ERR_STATE = 3,       /* call order: operators / cases not set yet */
    LDC_ERR_UNSUPPORTED = 4  /* requested kernel path cannot run this shape */
};

/* kernel path selection for ldc_step / ldc_solve */
enum {
    LDC_PATH_AUTO = 0,        /* even-odd resident kernel when the shape allows, else resident, else tiled */
    LDC_PATH_TILED = 1,       /* state in global memory (L2), one launch per RK stage; any nx, ny */
    LDC_PATH_RESIDENT = 2,    /* state + operators resident in shared memory, FP64 DMMA, many steps per launch */
    LDC_PATH_RESIDENT_EO = 3, /* resident + even-odd (parity) splitting of the operators: half the DMMA work;
                                 needs nx == ny, identical x / y operators and nx - 1 <= 64 */
    LDC_PATH_STREAM_EO = 4,   /* even-odd, one cavity spread over up to 16 CTAs, operands streamed from L2, one launch per
                                 RK stage (a programmatic-dependent-launch chain): 34 <= nx <= 129, nx == ny, identical
                                 x / y operators; AUTO picks it for nx >= 66 (e.g. N = 128) and, when 4 x batch <= #SM,
                                 for 50 <= nx <= 65 (four SMs per cavity instead of one) */
    LDC_PATH_RESIDENT_EO_SPLIT = 5, /* resident even-odd with the u and v equations of a tile on different warps and the
                                 base state in shared memory: lower latency per step for few cavities on small grids
                                 (nx - 1 <= 48); AUTO picks it when nx - 1 <= 32 and batch <= #SM */
    LDC_PATH_STREAM_TMA = 6   /* the streamed even-odd algebra as ONE persistent cooperative launch per batch of steps:
                                 teams of up to 16 co-resident CTAs per cavity fed from a cavity queue, operand panels by TMA
                                 bulk copies (cp.async.bulk + mbarrier) from a k4-interleaved tile layout, the RK stages
                                 separated by a team barrier in global memory instead of a kernel boundary.  Same shapes as
                                 LDC_PATH_STREAM_EO; AUTO picks it wherever it picked that path (LDC_STREAM_TMA=0 in the
                                 environment keeps the launch-per-stage kernels for A / B runs).  Its kernel comes in three
                                 shapes chosen per launch from the number of cavities still iterating: 16 x 16 tiles
                                 (k_stream_tma), 16 x 32 tiles (k_stream_tma_wide) and, for 50 <= nx <= 65 with at most two
                                 cavities per thread-block cluster the device holds (15 on B200), k_cluster_eo: a cavity over a
                                 cluster of 8 CTAs with operators and state resident in distributed shared memory */
};

/* Library / build information: "ldc_b200 <version> sm_100a". */
const char* ldc_version(void);
const char* ldc_last_error(void);

/* Allocate the device state of `batch` independent cavities of polynomial order (nx, ny).
 * Replaces SpectralSolverFields.allocate (datastructures.py:282-353) / SpectralLevel
 * (spectral/multigrid/fsg.py:262-354).  `cuda_stream` is a cudaStream_t (NULL = default stream). */
int ldc_create(ldc_t** out, int nx, int ny, int batch, int device, void* cuda_stream);
void ldc_destroy(ldc_t* h);
int ldc_set_path(ldc_t* h, int path);
/* 0 when the resident DMMA path can run this handle's shape, else LDC_ERR_UNSUPPORTED. */
int ldc_resident_supported(const ldc_t* h);
/* Name of the stepping kernel ldc_step / ldc_solve launch for this handle with its current path and operators:
 * "k_resident_eo", "k_resident_eo_split", "k_resident_rk4", "k_stream_stage", "tiled" (or "unsupported" when the forced path cannot run the shape). */
const char* ldc_kernel_name(const ldc_t* h);

/* 1-D operators shared by every cavity of the handle (host or device pointers, row-major):
 *   Dx, Dxx : (nx+1) x (nx+1)   Dy, Dyy : (ny+1) x (ny+1)
 *   Ix : (nx+1) x (nx-1)        Iy : (ny+1) x (ny-1)     inner -> full interpolation
 * Replaces SGSolver._build_diff_matrices (spectral/sg.py:181-210) as the kernels' operand source;
 * dx_min, dy_min are min(diff(nodes)) (spectral/sg.py:118-119). */
int ldc_set_operators(ldc_t* h, const double* Dx, const double* Dxx, const double* Dy, const double* Dyy,
                      const double* Ix, const double* Iy, double dx_min, double dy_min);

/* Per-cavity physics: Re, beta_squared, CFL, lid_velocity are [batch]; u_lid, v_lid are
 * [batch][nx+1] lid profiles on the north boundary (spectral/operators/corner.py:80-112). */
int ldc_set_cases(ldc_t* h, const double* Re, const double* beta2, const double* CFL, const double* lid_velocity,
                  const double* u_lid, const double* v_lid);

/* State transfer: u, v [batch][nx+1][ny+1]; p [batch][nx-1][ny-1].  NULL pointers are skipped. */
int ldc_set_state(ldc_t* h, const double* u, const double* v, const double* p);
int ldc_get_state(ldc_t* h, double* u, double* v, double* p);
/* u = v = p = 0 then the lid column (spectral/sg.py:98, 250-253; multigrid/fsg.py:1141-1161). */
int ldc_init_state(ldc_t* h);

/* One residual evaluation of the current state on ALL nodes, exactly the reference formula
 * (SGSolver._compute_residuals spectral/sg.py:278-346, MultigridSmoother._compute_residuals
 * multigrid/fsg.py:857-910): Ru, Rv [batch][nx+1][ny+1], Rp [batch][nx-1][ny-1]; norms (may be
 * NULL) receives [batch][3] = ||Ru||, ||Rv||, ||Rp|| (spectral/sg.py:463-473). */
int ldc_residual(ldc_t* h, double* Ru, double* Rv, double* Rp, double* norms);

/* n_steps RK4 pseudo-steps of every cavity, no convergence test (SGSolver.step spectral/sg.py:410-449
 * with freeze_pressure=1: all four stages take grad(p^n); MultigridSmoother.step
 * multigrid/fsg.py:956-995 with freeze_pressure=0).  rel (may be NULL) receives [batch][2] =
 * (u_res, v_res) of the last step, ||phi^{n+1}-phi^n|| / (||phi^n|| + 1e-12). */
int ldc_step(ldc_t* h, int64_t n_steps, int freeze_pressure, double* rel);

/* Iterate every cavity until max(u_res, v_res) < tol[b] (after `warmup` unconditional iterations:
 * 10 for base.solve base.py:283-286, 0 for solve_fsg multigrid/fsg.py:1174-1182), a non-finite
 * residual (nan_exit=1, multigrid/fsg.py:1184-1202) or max_iter steps.  Outputs are [batch]:
 * iters = steps taken, converged / diverged flags, rel_res = last max(u_res, v_res). */
int ldc_solve(ldc_t* h, const double* tol, int64_t max_iter, int warmup, int freeze_pressure, int nan_exit,
              int64_t* iters, int32_t* converged, int32_t* diverged, double* rel_res);

/* ldc_solve cut into pieces: iterations [start_iter, stop_iter) of the solve whose limit is max_iter (same loop, same
 * convergence rule: base.py:243-313 / multigrid/fsg.py:1168-1202).  start_iter = 0 starts a solve (status reset);
 * start_iter > 0 continues the previous call of this handle, whose active cavities have taken exactly start_iter
 * iterations.  A solve cut into segments takes bitwise the steps of the uncut ldc_solve; the caller may read the state
 * (ldc_get_state) between segments without disturbing it.  Outputs as ldc_solve (any may be NULL). */
int ldc_solve_segment(ldc_t* h, const double* tol, int64_t start_iter, int64_t stop_iter, int64_t max_iter, int warmup,
                      int freeze_pressure, int nan_exit, int64_t* iters, int32_t* converged, int32_t* diverged,
                      double* rel_res);

/* base.solve with its per-iteration records (base.py:243-313): ldc_solve on the launch-per-stage path, and after every
 * iteration i (0-based) of a cavity, hist[b][i - start_iter][7] = { rel_iter = max(u_res, v_res), ||R_u||, ||R_v||, ||R_p||
 * of the LAST stage's residual on all nodes (what arrays.R_* hold after step(), spectral/sg.py:463-473), energy,
 * enstrophy, palinstrophy of the new state with quadrature weights W[(nx+1)][(ny+1)] (spectral/sg.py:495-550) }.
 * The call takes at most hist_cap iterations: the caller streams a long solve in chunks, start_iter = 0 starts a solve
 * (status reset), start_iter > 0 continues the previous call of this handle (iters keeps counting; max_iter is absolute).
 * Rows of iterations a cavity did not take stay zero; the reference keeps rows i >= 10 only (base.py:262-276).
 * Grids whose operators, state and work arrays fit one SM's shared memory (N <= 36) run a whole call in ONE launch
 * (k_history_fused); larger grids, and handles forced to LDC_PATH_TILED, take one launch per RK stage. */
int ldc_solve_history(ldc_t* h, const double* tol, int64_t start_iter, int64_t max_iter, int warmup, int freeze_pressure,
                      int nan_exit, const double* W, double* hist, int64_t hist_cap, int64_t* iters, int32_t* converged,
                      int32_t* diverged, double* rel_res);

/* base.solve's records at chosen iterations only: the reference keeps every record and then thins the series to 1000 points
 * (_store_results base.py:136-141, indices np.linspace(0, L-1, 1000, dtype=int)), so a caller
 * that knows the iteration count L of the solve (from a first ldc_solve on the same start state: the stepping kernels are
 * deterministic) needs the records of those iterations alone.  The call is ldc_solve -- same stepping kernels, same
 * trajectory, same iters / flags / final state -- and, for each k < n_rec, hist[b][k][7] = the record (layout of
 * ldc_solve_history) of iteration rec_iter[k] (0-based, strictly increasing, < max_iter), formed by one iteration of the
 * history path on a copy of the state the cavity has after rec_iter[k] iterations.  Rows of iterations a cavity did not take
 * stay zero. */
int ldc_solve_sampled(ldc_t* h, const double* tol, int64_t max_iter, int warmup, int freeze_pressure, int nan_exit,
                      const double* W, int64_t n_rec, const int64_t* rec_iter, double* hist, int64_t* iters,
                      int32_t* converged, int32_t* diverged, double* rel_res);

/* Energy, enstrophy, palinstrophy with quadrature weights W[(nx+1)][(ny+1)]
 * (spectral/sg.py:495-550); out is [batch][3]. */
int ldc_diagnostics(ldc_t* h, const double* W, double* out);

/* Streamfunction + vortex metrics of every cavity of the handle (SGSolver._compute_streamfunction / compute_vortex_metrics,
 * spectral/sg.py:556-743; the reference assembles a sparse Kronecker Laplacian and calls SuperLU once per cavity).  Here the
 * interior system A psi + psi B^T = -omega (A = Dxx[1:-1,1:-1], B = Dyy[1:-1,1:-1]) is solved for the whole batch by
 * tensor-product fast diagonalisation with one step of iterative refinement; the caller supplies the eigen-decompositions
 * A = Va diag(la) Va^-1, B = Vb diag(lb) Vb^-1 (mx = nx - 1, my = ny - 1; all matrices row-major):
 *   Vai [mx][mx] = Va^-1, VbiT [my][my] = (Vb^-1)^T, Va [mx][mx], VbT [my][my] = Vb^T, inv_lam [mx][my] = 1 / (la_i + lb_j),
 *   x [nx+1], y [ny+1] the collocation nodes.
 * out [batch][19] = psi_min, psi_min_x, psi_min_y, omega_center, omega_max, omega_max_x, omega_max_y, then (psi, omega, x, y)
 * of the BR, BL, TL corner vortices (zeros where the quadrant's maximum of psi is not positive) -- the vortex fields of
 * `Metrics` (datastructures.py:73-95) in that order.  psi (optional) receives the streamfunction [batch][nx+1][ny+1]. */
int ldc_vortex_metrics(ldc_t* h, const double* Vai, const double* VbiT, const double* Va, const double* VbT,
                       const double* inv_lam, const double* x, const double* y, double* out, double* psi);

/* compute_validation_errors (base.py:970-1054) for every cavity of the handle: u, v are evaluated at nq scattered points by
 * bilinear interpolation on the collocation grid (_evaluate_at_points base.py:1122-1160: SciPy RegularGridInterpolator,
 * method="linear") and compared with a stored solution (ref_u, ref_v) [nq]:
 *   out[b] = { ||u - ref_u|| / (||ref_u|| + 1e-12), same for v, number of points used }
 * over the points with mask != 0 whose interpolated values are not NaN.  The caller supplies the interpolation plan, shared
 * by the batch: cell[q] = ix * (ny+1) + iy of the lower grid node of the cell holding point q, tx, ty = normalised
 * distances inside that cell; mask[q] = 0 drops a point (outside the domain, or within 1e-10 of the walls, base.py:1022-1027). */
int ldc_validation_errors(ldc_t* h, int64_t nq, const int32_t* cell, const double* tx, const double* ty, const int32_t* mask,
                          const double* ref_u, const double* ref_v, double* out);

/* Coarse -> fine transfer of (u, v, p): U_f = Px U_c Py^T with dense 1-D maps
 * Pxu (nfx+1)x(ncx+1), Pyu (nfy+1)x(ncy+1), Pxp (nfx-1)x(ncx-1), Pyp (nfy-1)x(ncy-1)
 * (prolongate_solution multigrid/fsg.py:551-614).  bc_quirk=1 reproduces the reference's
 * swapped-axes boundary re-enforcement (row 0 := 0, row -1 := lid_velocity, col 0 := 0, col -1 := 0)
 * followed by initialize_lid (multigrid/fsg.py:950-954); bc_quirk=0 applies the plain wall + lid
 * boundary conditions.  Both handles must share batch size, device and have their cases set. */
int ldc_prolongate(ldc_t* coarse, ldc_t* fine, const double* Pxu, const double* Pyu, const double* Pxp,
                   const double* Pyp, int bc_quirk);
/* Fine -> coarse transfer U_c = Rx U_f Ry^T (restrict_solution multigrid/fsg.py:622-664; no caller in
 * the reference's FSG driver) followed by the wall + lid boundary conditions on the coarse grid. */
int ldc_restrict(ldc_t* fine, ldc_t* coarse, const double* Rxu, const double* Ryu, const double* Rxp,
                 const double* Ryp);

/* Kernel launches issued by this handle since creation (bench.py's gpu_launches claim) and device
 * time in ms of the last ldc_step / ldc_solve call, measured with CUDA events on the handle's stream. */
int64_t ldc_launch_count(const ldc_t* h);
double ldc_last_kernel_ms(const ldc_t* h);

/* FP64 ceilings of the device, measured live (roofline denominators): which = 0 register-resident
 * DMMA (mma.sync m8n8k4 f64), 1 register-resident DFMA, 2 both pipes mixed.  TFLOP/s in *tflops. */
int ldc_fp64_peak(int device, int which, double* tflops);

/* Allocation-free variant of ldc_diagnostics.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It needs no handle and calls no cudaMalloc / cudaFree or allocating library routine: the vorticity field the two kernels
 * hand over lives in a file-scope static __device__ array of LDC_NOALLOC_MAX_DOUBLES doubles (4 MiB), which every GPU that
 * loads the library holds until the process exits.  All pointers are DEVICE pointers on `device`, row-major:
 *   d_u, d_v [batch][nx+1][ny+1]; d_Dx [nx+1][nx+1]; d_DyT [ny+1][ny+1] = Dy^T; d_W [nx+1][ny+1] quadrature weights;
 *   d_out [batch][3] = energy, enstrophy, palinstrophy (as ldc_diagnostics).
 * The launches go to `cuda_stream` (NULL: the default stream) and the call returns without synchronising.
 * batch * (nx+1) * (ny+1) > LDC_NOALLOC_MAX_DOUBLES (or batch > 65535) is refused with LDC_ERR_INVALID.
 * All calls share the one array per GPU, which keeps the last call's contents (the part in use is cleared first): two calls
 * that overlap on the same GPU (different streams or threads) RACE -- keep them on one stream or order them with events.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress) and should be made outside a capture. */
#define LDC_NOALLOC_MAX_DOUBLES 524288
int ldc_diagnostics_noalloc(int device, void* cuda_stream, int nx, int ny, int batch, const double* d_u, const double* d_v,
                            const double* d_Dx, const double* d_DyT, const double* d_W, double* d_out);

#ifdef __cplusplus
}
#endif
#endif /* LDC_B200_H */
