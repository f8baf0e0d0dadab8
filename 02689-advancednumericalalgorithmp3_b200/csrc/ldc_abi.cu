// C ABI of ldc_b200 (include/ldc_b200.h): handle management, host<->device plumbing and the launch
// sequences of the tiled and resident kernel paths.  No exception crosses the boundary.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"
#include "peak.cuh"
#include "resident_eo_main.cuh"
#include "resident_eo_split.cuh"
#include "resident_main.cuh"
#include "stream_eo.cuh"
#include "stream_tma.cuh"
#include "stream_tma_wide.cuh"
#include "cluster_eo.cuh"
#include "tiled.cuh"
#include "validate.cuh"
#include "vortex.cuh"
#include "history_fused.cuh"

namespace ldc {
thread_local std::string g_last_error;

static bool is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Copy in either direction; host pointers are complete when this returns.
static int copy_in(ldc_handle* h, double* dst_dev, const double* src, size_t n) {
    if (!src) return LDC_OK;
    const bool dev = is_device_ptr(src);
    LDC_CUDA(cudaMemcpyAsync(dst_dev, src, n * sizeof(double), dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                             h->stream));
    if (!dev) LDC_CUDA(cudaStreamSynchronize(h->stream));
    return LDC_OK;
}

static int copy_out(ldc_handle* h, void* dst, const void* src_dev, size_t bytes) {
    if (!dst) return LDC_OK;
    const bool dev = is_device_ptr(dst);
    LDC_CUDA(cudaMemcpyAsync(dst, src_dev, bytes, dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
    if (!dev) LDC_CUDA(cudaStreamSynchronize(h->stream));
    return LDC_OK;
}

// Zero-filled device allocation.  The fill is ordered on the handle's stream (a caller stream may be non-blocking, i.e.
// not ordered against the legacy default stream a plain cudaMemset would use).
template <typename T>
static int dalloc(ldc_handle* h, T** p, size_t n) {
    LDC_CUDA(cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(T)));
    LDC_CUDA(cudaMemsetAsync(*p, 0, std::max<size_t>(n, 1) * sizeof(T), h->stream));
    return LDC_OK;
}

static dim3 grid2d(int rows, int cols, int B) { return dim3((cols + 31) / 32, (rows + 7) / 8, B); }
static const dim3 kBlock2d(32, 8, 1);

// p_full = Ix p Iy^T of `psrc` into h->pfull (spectral/sg.py:255-276).
static void launch_pfull(ldc_handle* h, const double* psrc, const CavityStatus* mask) {
    k_left_mul<<<grid2d(h->n1x, h->my, h->B), kBlock2d, 0, h->stream>>>(h->Ix, psrc, h->ptmp, h->n1x, h->mx, h->my, mask);
    // Iy is (n1y x my) row-major, i.e. already "A^T" (my x n1y) of the map p -> p Iy^T when read as AT[k][j] = Iy[j][k];
    // k_right_mul wants AT (ca x ra) = Iy^T stored row-major, which set_operators prepared in h->IyT.
    k_right_mul<<<grid2d(h->n1x, h->n1y, h->B), kBlock2d, 0, h->stream>>>(h->ptmp, h->IyT, h->pfull, h->n1x, h->my, h->n1y,
                                                                          mask);
    h->launches += 2;
}

static void launch_residual(ldc_handle* h, const double* u, const double* v, const double* psrc, const CavityStatus* mask) {
    launch_pfull(h, psrc, mask);
    k_residual<<<grid2d(h->n1x, h->n1y, h->B), kBlock2d, 0, h->stream>>>(u, v, h->pfull, h->Dx, h->Dxx, h->DyT, h->DyyT, h->Re,
                                                                        h->beta2, h->Ru, h->Rv, h->Rp, h->n1x, h->n1y, mask);
    h->launches += 1;
}

// One RK4 step of every active cavity on the tiled path.
static void launch_tiled_step(ldc_handle* h, int freeze_pressure, int warmup, long long max_iter, int nan_exit,
                              const double* tol) {
    const int nfull = h->n1x * h->n1y, ninner = h->mx * h->my;
    k_step_begin<<<h->B, RED_THREADS, 0, h->stream>>>(h->u, h->v, h->p, h->u0, h->v0, h->p0, h->dt, h->Re, h->beta2, h->CFL,
                                                     h->lidvel, h->dx_min, h->dy_min, nfull, ninner, h->status);
    h->launches += 1;
    for (int s = 0; s < 4; ++s) {
        // freeze_pressure: SGSolver takes grad(p^n) in all four stages; p itself is still staged
        launch_residual(h, h->u, h->v, freeze_pressure ? h->p0 : h->p, h->status);
        k_stage_update<<<grid2d(h->n1x, h->n1y, h->B), kBlock2d, 0, h->stream>>>(
            h->u, h->v, h->p, h->u0, h->v0, h->p0, h->Ru, h->Rv, h->Rp, h->dt, h->ulid, h->vlid, rk_alpha(s), h->n1x, h->n1y, 1,
            h->status);
        h->launches += 1;
    }
    k_step_end<<<h->B, RED_THREADS, 0, h->stream>>>(h->u, h->v, h->u0, h->v0, nfull, tol, warmup, max_iter, nan_exit, h->status);
    h->launches += 1;
}

static int count_active(ldc_handle* h, int* out) {
    k_count_active<<<1, 256, 0, h->stream>>>(h->status, h->B, h->n_active);
    h->launches += 1;
    LDC_CUDA(cudaMemcpyAsync(h->h_pinned, h->n_active, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    LDC_CUDA(cudaStreamSynchronize(h->stream));
    *out = *h->h_pinned;
    return LDC_OK;
}

static int check_ready(const ldc_handle* h, const char* who) {
    if (!h) return fail(LDC_ERR_INVALID, std::string(who) + ": null handle");
    if (!h->ops_set) return fail(LDC_ERR_STATE, std::string(who) + ": ldc_set_operators has not been called");
    if (!h->cases_set) return fail(LDC_ERR_STATE, std::string(who) + ": ldc_set_cases has not been called");
    return LDC_OK;
}

}  // namespace ldc

using namespace ldc;

namespace ldc {

static bool resident_shape_ok(const ldc_handle* h) {
    return h->nx == h->ny && res_pick_mp(h->mx) != 0;
}

// Dynamic shared memory the resident kernel needs for this handle (bytes), 0 if unsupported.
static size_t resident_smem_bytes(const ldc_handle* h) {
    if (!resident_shape_ok(h)) return 0;
    const int MP = res_pick_mp(h->mx);
    const int warps = MP == 64 ? 8 : (MP == 48 ? 6 : 4);
    const size_t MAT = (size_t)MP * res_ld(MP);
    const size_t d = (h->square_ops ? 3 : 6) * MAT + 8 * MP + 3 * MAT + 8 * MP + 8 + 2 * MP + 4 + 8 * warps;
    return d * sizeof(double);
}

static bool resident_ok(const ldc_handle* h) {
    const size_t bytes = resident_smem_bytes(h);
    if (bytes == 0) return false;
    int max_optin = 0;
    if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device) != cudaSuccess) return false;
    return bytes <= (size_t)max_optin;
}

static bool eo_shape_ok(const ldc_handle* h) { return h->nx == h->ny && eo_pick_hp(h->mx) != 0; }

static size_t eo_smem_bytes(int HP) {
    switch (HP) {
        case 8: return EoCfg<8, 1>::smem_doubles() * sizeof(double);
        case 16: return EoCfg<16, 1>::smem_doubles() * sizeof(double);
        case 24: return EoCfg<24, 1>::smem_doubles() * sizeof(double);
        case 32: return EoCfg<32, 2>::smem_doubles() * sizeof(double);
    }
    return 0;
}

// the even-odd kernel needs one set of operators for both directions (checked bit for bit in ldc_set_operators)
static bool eo_ok(const ldc_handle* h) {
    if (!eo_shape_ok(h) || !h->square_ops || !h->eo_sym) return false;
    int max_optin = 0;
    if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device) != cudaSuccess) return false;
    return eo_smem_bytes(h->HP) <= (size_t)max_optin;
}

// equation-split variant (resident_eo_split.cuh): small grids only, the base state stays in shared memory too
static size_t eo_split_smem_bytes(int HP) {
    switch (HP) {
        case 8: return EoSplitCfg<8>::smem_doubles() * sizeof(double);
        case 16: return EoSplitCfg<16>::smem_doubles() * sizeof(double);
        case 24: return EoSplitCfg<24>::smem_doubles() * sizeof(double);
    }
    return 0;
}
static bool eo_split_ok(const ldc_handle* h) {
    if (!eo_ok(h) || h->HP > 24) return false;
    int max_optin = 0;
    if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device) != cudaSuccess) return false;
    return eo_split_smem_bytes(h->HP) <= (size_t)max_optin;
}

static bool stream_shape_ok(const ldc_handle* h) {
    const int hh = eo_half(h->mx);
    return h->nx == h->ny && hh > 16 && hh <= ST_HP;   // below: one 16 x 16 tile per block, nothing to spread
}
static bool stream_ok(const ldc_handle* h) { return stream_shape_ok(h) && h->square_ops && h->eo_sym; }
// persistent TMA-fed variant: needs a cooperative launch (co-resident teams) and its shared memory
static bool stream_tma_ok(const ldc_handle* h) {
    if (!stream_ok(h)) return false;
    int coop = 0, max_optin = 0;
    if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device) != cudaSuccess || !coop) return false;
    if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device) != cudaSuccess) return false;
    return S2_SMEM_DOUBLES * sizeof(double) <= (size_t)max_optin;
}
// which streamed kernel AUTO takes: the persistent one unless LDC_STREAM_TMA=0 (A / B runs)
static int stream_auto(const ldc_handle* h) {
    const char* e = getenv("LDC_STREAM_TMA");
    return (stream_tma_ok(h) && !(e && atoi(e) == 0)) ? 5 : 3;
}

// Clusters of 8 CTAs of k_cluster_eo the device holds at once (0: the kernel cannot run here).
static int cluster_eo_clusters(const ldc_handle* h) {
    const size_t smem = CE_SMEM_DOUBLES * sizeof(double);
    if (cudaFuncSetAttribute(k_cluster_eo, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CE_CLUSTER;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3(CE_CLUSTER * std::max(1, h->num_sms / CE_CLUSTER));
    cfg.blockDim = dim3(CE_THREADS);
    cfg.dynamicSmemBytes = smem;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, k_cluster_eo, &cfg) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

// Shape of the persistent streamed path for `nact` cavities still iterating: 0 = 16 x 16 tiles (k_stream_tma), 1 = 16 x 32
// tiles (k_stream_tma_wide), 2 = a cluster of 8 CTAs per cavity with the state in distributed shared memory (k_cluster_eo:
// parity blocks of 32 with all four 8-row tiles in use, i.e. 50 <= N <= 65, and at most two cavities per cluster the device holds).
// LDC_STREAM_CW = 1 | 2 | 3 forces a shape (3 = cluster, where the grid allows it).
static int stream_tma_shape(const ldc_handle* h, int nact, int* clusters_out = nullptr) {
    const int hh = eo_half(h->mx), ntile = (hh + ST_T - 1) / ST_T;
    const char* e = getenv("LDC_STREAM_CW");
    const int forced = e ? atoi(e) : 0;
    if (ntile == 2 && hh > 24 && (forced == 0 || forced == 3)) {
        const int nc = cluster_eo_clusters(h);
        if (clusters_out) *clusters_out = nc;
        // up to two rounds of clusters it also wins on throughput: the queue keeps every cluster busy (N = 64, B200: 29 cavities
        // 909 k against 810 k steps/s on the 16 x 16 tiles, 33 cavities 914 k against 922 k)
        if (nc > 0 && (forced == 3 || nact <= 2 * nc + 1)) return 2;
    }
    if (forced == 1) return 0;
    if (forced == 2) return ntile % 2 == 0 ? 1 : 0;
    return (ntile % 2 == 0 && (long long)nact * 10 >= 16LL * (h->num_sms / (ntile * ntile))) ? 1 : 0;
}

// 0 tiled, 1 resident, 2 even-odd resident, 3 streamed even-odd, 4 even-odd resident split by equation,
// 5 streamed even-odd, persistent + TMA
static int pick_kernel(const ldc_handle* h) {
    if (h->path == LDC_PATH_TILED) return 0;
    if (h->path == LDC_PATH_STREAM_TMA) return stream_tma_ok(h) ? 5 : -1;
    if (h->path == LDC_PATH_RESIDENT_EO_SPLIT) return eo_split_ok(h) ? 4 : -1;
    if (h->path == LDC_PATH_RESIDENT) return resident_ok(h) ? 1 : -1;
    if (h->path == LDC_PATH_RESIDENT_EO) return eo_ok(h) ? 2 : -1;
    if (h->path == LDC_PATH_STREAM_EO) return stream_ok(h) ? 3 : -1;
    // few cavities on a small grid: latency matters, spread each cavity over twice the warps
    if (eo_split_ok(h) && h->HP <= 16 && h->B <= h->num_sms) return 4;   // measured: no gain at HP = 24
    // few cavities on the largest resident grids (50 <= N <= 65): four SMs per cavity on the streamed path beat one
    // (N = 64: 28.0 k against 16.7 k steps/s for one cavity, 822 k against 614 k for 37; 74 cavities: 1.02 M against 1.23 M)
    if (stream_ok(h) && eo_half(h->mx) > 24 && eo_half(h->mx) <= 32 &&
        (stream_auto(h) == 5 ? 2LL : 4LL) * h->B <= h->num_sms) return stream_auto(h);
    if (eo_ok(h)) return 2;
    if (resident_ok(h)) return 1;
    return stream_ok(h) ? stream_auto(h) : 0;
}

// canonical <-> packed state synchronisation
static int ensure_canonical(ldc_handle* h) {
    if (h->canon_valid) return LDC_OK;
    const int n1 = h->n1x;
    const dim3 grid((n1 * n1 + 255) / 256, h->B);
    if (h->packed_valid == 4)
        k_s2_unpack_state<<<grid, 256, 0, h->stream>>>(h->u, h->v, h->p, h->t_state, (size_t)h->B * S2_STRIDE, h->status, n1);
    else if (h->packed_valid == 3)
        k_unpack_state_eo<<<grid, 256, 0, h->stream>>>(h->u, h->v, h->p, h->s_state, h->s_stride, n1, ST_HP);
    else if (h->packed_valid == 2)
        k_unpack_state_eo<<<grid, 256, 0, h->stream>>>(h->u, h->v, h->p, h->e_state, h->e_state_stride, n1, h->HP);
    else
        k_unpack_state<<<grid, 256, 0, h->stream>>>(h->u, h->v, h->p, h->r_state, h->r_state_stride, n1, h->MP);
    h->launches += 1;
    LDC_CUDA(cudaGetLastError());
    h->canon_valid = true;
    return LDC_OK;
}

static int ensure_packed(ldc_handle* h, int kind) {
    if (h->packed_valid == kind) return LDC_OK;
    int rc = ensure_canonical(h);
    if (rc) return rc;
    if (kind == 4) {
        k_s2_pack_state<<<dim3((S2_HP * S2_HP + 255) / 256, h->B), 256, 0, h->stream>>>(h->u, h->v, h->p, h->t_state, h->n1x);
    } else if (kind == 2 || kind == 3) {
        const int HP = kind == 3 ? ST_HP : h->HP, LD = eo_ld(HP);
        k_pack_state_eo<<<dim3((std::max(HP * LD, 2 * HP) + 255) / 256, h->B), 256, 0, h->stream>>>(
            h->u, h->v, h->p, kind == 3 ? h->s_state : h->e_state, kind == 3 ? h->s_stride : h->e_state_stride, h->n1x, HP);
    } else {
        const int MP = h->MP, LD = res_ld(MP);
        k_pack_state<<<dim3((MP * LD + 255) / 256, h->B), 256, 0, h->stream>>>(h->u, h->v, h->p, h->r_state, h->r_state_stride,
                                                                               h->n1x, MP);
    }
    h->launches += 1;
    LDC_CUDA(cudaGetLastError());
    h->packed_valid = kind;
    return LDC_OK;
}

}  // namespace ldc

extern "C" {

const char* ldc_version(void) { return "ldc_b200 0.1.0 sm_100a"; }
const char* ldc_last_error(void) { return g_last_error.c_str(); }

int ldc_create(ldc_t** out, int nx, int ny, int batch, int device, void* cuda_stream) {
    if (!out) return fail(LDC_ERR_INVALID, "ldc_create: out is null");
    *out = nullptr;
    if (nx < 2 || ny < 2 || batch < 1) return fail(LDC_ERR_INVALID, "ldc_create: need nx, ny >= 2 and batch >= 1");
    if (batch > 65535)   // the batch index is gridDim.y / gridDim.z of the per-cavity kernels
        return fail(LDC_ERR_INVALID, "ldc_create: batch must be <= 65535 (split larger sweeps over several handles)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(LDC_ERR_CUDA, "ldc_create: no CUDA device is visible (this library has no CPU fallback)");
    }
    if (device < 0 || device >= ndev) return fail(LDC_ERR_INVALID, "ldc_create: device index out of range");
    LDC_CUDA(cudaSetDevice(device));
    ldc_handle* h = new (std::nothrow) ldc_handle();
    if (!h) return fail(LDC_ERR_INVALID, "ldc_create: out of host memory");
    h->nx = nx; h->ny = ny; h->n1x = nx + 1; h->n1y = ny + 1; h->mx = nx - 1; h->my = ny - 1;
    h->B = batch; h->device = device; h->stream = (cudaStream_t)cuda_stream;
    const size_t nf = (size_t)h->n1x * h->n1y, ni = (size_t)h->mx * h->my, B = batch;
    int rc = LDC_OK;
#define A_(p, n) if (rc == LDC_OK) rc = dalloc(h, &h->p, (n))
    A_(Dx, (size_t)h->n1x * h->n1x); A_(Dxx, (size_t)h->n1x * h->n1x);
    A_(Dy, (size_t)h->n1y * h->n1y); A_(Dyy, (size_t)h->n1y * h->n1y);
    A_(DyT, (size_t)h->n1y * h->n1y); A_(DyyT, (size_t)h->n1y * h->n1y);
    A_(Ix, (size_t)h->n1x * h->mx); A_(Iy, (size_t)h->n1y * h->my); A_(IyT, (size_t)h->n1y * h->my);
    A_(Re, B); A_(beta2, B); A_(CFL, B); A_(lidvel, B); A_(ulid, B * h->n1x); A_(vlid, B * h->n1x);
    A_(u, B * nf); A_(v, B * nf); A_(p, B * ni); A_(u0, B * nf); A_(v0, B * nf); A_(p0, B * ni);
    A_(Ru, B * nf); A_(Rv, B * nf); A_(Rp, B * ni); A_(ptmp, B * h->n1x * h->my); A_(pfull, B * nf);
    A_(dt, B); A_(tol, B); A_(status, B); A_(n_active, 1);
    if (rc == LDC_OK && resident_shape_ok(h)) {
        h->MP = res_pick_mp(h->mx);
        h->LD = res_ld(h->MP);
        h->r_state_stride = res_state_stride(h->MP);
        A_(r_ops, res_ops_size(h->MP, false));
        A_(r_state, B * h->r_state_stride);
        A_(q_data, B); A_(q_seq, B); A_(q_ctl, 4);
    }
    if (rc == LDC_OK && eo_shape_ok(h)) {
        h->HP = eo_pick_hp(h->mx);
        h->e_state_stride = eo_state_stride(h->HP);
        A_(e_ops, eo_ops_size(h->HP));
        A_(e_state, B * h->e_state_stride);
    }
    if (rc == LDC_OK && stream_shape_ok(h)) {   // the state buffers of the streamed path come with its first run
        if (!eo_shape_ok(h)) h->HP = ST_HP;
        h->s_stride = eo_state_stride(ST_HP);
        A_(s_ops, eo_ops_size(ST_HP));
    }
    if (rc == LDC_OK && cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess)
        rc = fail(LDC_ERR_CUDA, "ldc_create: cannot query the SM count");
#undef A_
    if (rc == LDC_OK && cudaMallocHost((void**)&h->h_pinned, sizeof(int)) != cudaSuccess)
        rc = fail(LDC_ERR_CUDA, "ldc_create: cudaMallocHost failed");
    if (rc == LDC_OK && (cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess))
        rc = fail(LDC_ERR_CUDA, "ldc_create: cudaEventCreate failed");
    if (rc != LDC_OK) {
        const std::string keep = g_last_error;
        ldc_destroy(h);
        g_last_error = keep;
        return rc;
    }
    *out = h;
    return LDC_OK;
}

void ldc_destroy(ldc_t* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    double* ptrs[] = {h->Dx, h->Dxx, h->Dy, h->Dyy, h->DyT, h->DyyT, h->Ix, h->Iy, h->IyT, h->Re, h->beta2, h->CFL, h->lidvel,
                      h->ulid, h->vlid, h->u, h->v, h->p, h->u0, h->v0, h->p0, h->Ru, h->Rv, h->Rp, h->ptmp, h->pfull, h->dt,
                      h->tol, h->r_ops, h->r_state, h->e_ops, h->e_state, h->s_state, h->s_partials, h->s_opsc, h->s_ops, h->hist, h->Wq, h->fork, h->t_state,
                      h->t_opsc, h->t_team};
    for (double* q : ptrs) if (q) cudaFree(q);
    if (h->status) cudaFree(h->status);
    if (h->n_active) cudaFree(h->n_active);
    if (h->q_data) cudaFree(h->q_data);
    if (h->q_seq) cudaFree(h->q_seq);
    if (h->q_ctl) cudaFree(h->q_ctl);
    if (h->h_pinned) cudaFreeHost(h->h_pinned);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->ev2) cudaEventDestroy(h->ev2);
    if (h->ev3) cudaEventDestroy(h->ev3);
    delete h;
}

int ldc_set_path(ldc_t* h, int path) {
    if (!h || path < LDC_PATH_AUTO || path > LDC_PATH_STREAM_TMA) return fail(LDC_ERR_INVALID, "ldc_set_path: bad argument");
    if (path == LDC_PATH_RESIDENT_EO_SPLIT && !(eo_shape_ok(h) && eo_pick_hp(h->mx) <= 24))
        return fail(LDC_ERR_UNSUPPORTED, "ldc_set_path: the equation-split even-odd path needs nx == ny and nx - 1 <= 48");
    if ((path == LDC_PATH_STREAM_EO || path == LDC_PATH_STREAM_TMA) && !stream_shape_ok(h))
        return fail(LDC_ERR_UNSUPPORTED, "ldc_set_path: the streamed even-odd path needs nx == ny and 34 <= nx <= 129");
    if (path == LDC_PATH_RESIDENT_EO && !eo_shape_ok(h))
        return fail(LDC_ERR_UNSUPPORTED, "ldc_set_path: the even-odd resident path needs nx == ny and nx - 1 <= 64");
    if (path == LDC_PATH_RESIDENT && !resident_shape_ok(h))
        return fail(LDC_ERR_UNSUPPORTED, "ldc_set_path: the resident path needs nx == ny and nx - 1 <= 64");
    h->path = path;
    return LDC_OK;
}

int ldc_resident_supported(const ldc_t* h) {
    if (!h) return fail(LDC_ERR_INVALID, "ldc_resident_supported: null handle");
    return resident_ok(h) ? LDC_OK : LDC_ERR_UNSUPPORTED;
}

const char* ldc_kernel_name(const ldc_t* h) {
    if (!h) return "unsupported";
    switch (pick_kernel(h)) {
        case 0: return "tiled";
        case 1: return "k_resident_rk4";
        case 2: return "k_resident_eo";
        case 3: return "k_stream_stage";
        case 4: return "k_resident_eo_split";
        case 5: {
            const int shape = stream_tma_shape(h, h->B);
            return shape == 2 ? "k_cluster_eo" : (shape == 1 ? "k_stream_tma_wide" : "k_stream_tma");
        }
    }
    return "unsupported";
}

int ldc_set_operators(ldc_t* h, const double* Dx, const double* Dxx, const double* Dy, const double* Dyy, const double* Ix,
                      const double* Iy, double dx_min, double dy_min) {
    if (!h || !Dx || !Dxx || !Dy || !Dyy || !Ix || !Iy) return fail(LDC_ERR_INVALID, "ldc_set_operators: null argument");
    if (!(dx_min > 0) || !(dy_min > 0)) return fail(LDC_ERR_INVALID, "ldc_set_operators: dx_min, dy_min must be positive");
    LDC_CUDA(cudaSetDevice(h->device));
    const size_t sx = (size_t)h->n1x * h->n1x, sy = (size_t)h->n1y * h->n1y;
    int rc;
    if ((rc = copy_in(h, h->Dx, Dx, sx))) return rc;
    if ((rc = copy_in(h, h->Dxx, Dxx, sx))) return rc;
    if ((rc = copy_in(h, h->Dy, Dy, sy))) return rc;
    if ((rc = copy_in(h, h->Dyy, Dyy, sy))) return rc;
    if ((rc = copy_in(h, h->Ix, Ix, (size_t)h->n1x * h->mx))) return rc;
    if ((rc = copy_in(h, h->Iy, Iy, (size_t)h->n1y * h->my))) return rc;
    h->dx_min = dx_min;
    h->dy_min = dy_min;
    const int T = 256;
    k_transpose<<<(int)((sy + T - 1) / T), T, 0, h->stream>>>(h->Dy, h->DyT, h->n1y, h->n1y);
    k_transpose<<<(int)((sy + T - 1) / T), T, 0, h->stream>>>(h->Dyy, h->DyyT, h->n1y, h->n1y);
    k_transpose<<<(h->n1y * h->my + T - 1) / T, T, 0, h->stream>>>(h->Iy, h->IyT, h->n1y, h->my);
    h->launches += 3;
    // square_ops: x and y operators identical bit for bit (lets the resident kernel keep one copy in shared memory)
    h->square_ops = false;
    if (h->nx == h->ny) {
        std::vector<double> a(sx), b(sx);
        bool same = (dx_min == dy_min);
        const double* px[3] = {h->Dx, h->Dxx, h->Ix};
        const double* py[3] = {h->Dy, h->Dyy, h->Iy};
        const size_t sz[3] = {sx, sx, (size_t)h->n1x * h->mx};
        for (int q = 0; q < 3 && same; ++q) {
            LDC_CUDA(cudaMemcpyAsync(a.data(), px[q], sz[q] * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
            LDC_CUDA(cudaMemcpyAsync(b.data(), py[q], sz[q] * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
            LDC_CUDA(cudaStreamSynchronize(h->stream));
            same = std::memcmp(a.data(), b.data(), sz[q] * sizeof(double)) == 0;
        }
        h->square_ops = same;
        // The even-odd kernels keep only the D_SA / D_AS, DD_SS / DD_AA and G_SA / G_AS parity blocks: they need D centro-
        // antisymmetric and DD, I centro-symmetric (nodes symmetric about the mid-point, as the Chebyshev and Legendre
        // Gauss-Lobatto grids are).  Checked here to 1e-10 of the largest entry; operators that fail it (a mapped or
        // stretched grid) run on the plain resident / tiled kernels, and a forced even-odd path reports LDC_ERR_UNSUPPORTED.
        h->eo_sym = false;
        if (same) {
            const int n1 = h->n1x, m = h->mx, N = n1 - 1;
            std::vector<double> d(sx), dd(sx), ix((size_t)n1 * m);
            LDC_CUDA(cudaMemcpyAsync(d.data(), h->Dx, sx * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
            LDC_CUDA(cudaMemcpyAsync(dd.data(), h->Dxx, sx * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
            LDC_CUDA(cudaMemcpyAsync(ix.data(), h->Ix, ix.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
            LDC_CUDA(cudaStreamSynchronize(h->stream));
            double md = 0, mdd = 0, mi = 0, ed = 0, edd = 0, ei = 0;
            for (int i = 0; i < n1; ++i)
                for (int k = 0; k < n1; ++k) {
                    const double a = d[(size_t)i * n1 + k], a2 = d[(size_t)(N - i) * n1 + (N - k)];
                    const double c = dd[(size_t)i * n1 + k], c2 = dd[(size_t)(N - i) * n1 + (N - k)];
                    md = std::max(md, std::fabs(a));
                    ed = std::max(ed, std::fabs(a + a2));
                    mdd = std::max(mdd, std::fabs(c));
                    edd = std::max(edd, std::fabs(c - c2));
                }
            for (int i = 0; i < n1; ++i)
                for (int k = 0; k < m; ++k) {
                    const double a = ix[(size_t)i * m + k], a2 = ix[(size_t)(N - i) * m + (m - 1 - k)];
                    mi = std::max(mi, std::fabs(a));
                    ei = std::max(ei, std::fabs(a - a2));
                }
            const double tol = 1e-10;
            h->eo_sym = ed <= tol * md && edd <= tol * mdd && ei <= tol * mi;   // (false when an entry is NaN)
        }
    }
    if (h->r_ops) {
        // G = (D I) rows 1..n-2: the pressure interpolation folded into the derivative (resident.cuh)
        const int MP = h->MP, LD = res_ld(MP), n1 = h->n1x, m = h->mx;
        const size_t MAT = (size_t)MP * LD;
        double* DI = h->pfull;  // scratch: n1 x m doubles fit in one cavity's full grid
        LDC_CUDA(cudaMemsetAsync(h->r_ops, 0, res_ops_size(MP, false) * sizeof(double), h->stream));
        double* vecs = h->r_ops + (h->square_ops ? 3 : 6) * MAT;
        k_left_mul<<<grid2d(n1, m, 1), kBlock2d, 0, h->stream>>>(h->Dx, h->Ix, DI, n1, n1, m, nullptr);
        k_pack_ops<<<(int)((MAT + T - 1) / T), T, 0, h->stream>>>(h->Dx, h->Dxx, DI, h->r_ops, vecs, n1, MP);
        k_left_mul<<<grid2d(n1, m, 1), kBlock2d, 0, h->stream>>>(h->Dy, h->Iy, DI, n1, n1, m, nullptr);
        k_pack_ops<<<(int)((MAT + T - 1) / T), T, 0, h->stream>>>(h->Dy, h->Dyy, DI, h->square_ops ? h->r_ops : h->r_ops + 3 * MAT,
                                                                 vecs + 4 * MP, n1, MP);
        h->launches += 4;
    }
    if (h->e_ops && h->square_ops) {
        const int HP = h->HP, n1 = h->n1x, m = h->mx;
        double* DI = h->pfull;  // scratch: D I (n1 x m)
        k_left_mul<<<grid2d(n1, m, 1), kBlock2d, 0, h->stream>>>(h->Dx, h->Ix, DI, n1, n1, m, nullptr);
        k_pack_ops_eo<<<(HP * eo_ld(HP) + T - 1) / T, T, 0, h->stream>>>(h->Dx, h->Dxx, DI, h->e_ops, n1, HP);
        h->launches += 2;
    }
    if (h->s_ops && h->square_ops) {
        const int n1 = h->n1x, m = h->mx;
        double* DI = h->pfull;
        k_left_mul<<<grid2d(n1, m, 1), kBlock2d, 0, h->stream>>>(h->Dx, h->Ix, DI, n1, n1, m, nullptr);
        k_pack_ops_eo<<<(ST_HP * eo_ld(ST_HP) + T - 1) / T, T, 0, h->stream>>>(h->Dx, h->Dxx, DI, h->s_ops, n1, ST_HP);
        h->launches += 2;
    }
    LDC_CUDA(cudaGetLastError());
    LDC_CUDA(cudaStreamSynchronize(h->stream));
    h->ops_set = true;
    return LDC_OK;
}

int ldc_set_cases(ldc_t* h, const double* Re, const double* beta2, const double* CFL, const double* lid_velocity,
                  const double* u_lid, const double* v_lid) {
    if (!h || !Re || !beta2 || !CFL || !lid_velocity || !u_lid || !v_lid)
        return fail(LDC_ERR_INVALID, "ldc_set_cases: null argument");
    LDC_CUDA(cudaSetDevice(h->device));
    int rc;
    if ((rc = copy_in(h, h->Re, Re, h->B))) return rc;
    if ((rc = copy_in(h, h->beta2, beta2, h->B))) return rc;
    if ((rc = copy_in(h, h->CFL, CFL, h->B))) return rc;
    if ((rc = copy_in(h, h->lidvel, lid_velocity, h->B))) return rc;
    if ((rc = copy_in(h, h->ulid, u_lid, (size_t)h->B * h->n1x))) return rc;
    if ((rc = copy_in(h, h->vlid, v_lid, (size_t)h->B * h->n1x))) return rc;
    h->cases_set = true;
    return LDC_OK;
}

int ldc_set_state(ldc_t* h, const double* u, const double* v, const double* p) {
    if (!h) return fail(LDC_ERR_INVALID, "ldc_set_state: null handle");
    LDC_CUDA(cudaSetDevice(h->device));
    int rc;
    if ((rc = ensure_canonical(h))) return rc;  // partial updates keep the other fields
    const size_t nf = (size_t)h->n1x * h->n1y, ni = (size_t)h->mx * h->my;
    if ((rc = copy_in(h, h->u, u, h->B * nf))) return rc;
    if ((rc = copy_in(h, h->v, v, h->B * nf))) return rc;
    if ((rc = copy_in(h, h->p, p, h->B * ni))) return rc;
    h->packed_valid = 0;
    return LDC_OK;
}

int ldc_get_state(ldc_t* h, double* u, double* v, double* p) {
    if (!h) return fail(LDC_ERR_INVALID, "ldc_get_state: null handle");
    LDC_CUDA(cudaSetDevice(h->device));
    int rc;
    if ((rc = ensure_canonical(h))) return rc;
    const size_t nf = (size_t)h->n1x * h->n1y, ni = (size_t)h->mx * h->my;
    if ((rc = copy_out(h, u, h->u, h->B * nf * sizeof(double)))) return rc;
    if ((rc = copy_out(h, v, h->v, h->B * nf * sizeof(double)))) return rc;
    if ((rc = copy_out(h, p, h->p, h->B * ni * sizeof(double)))) return rc;
    return LDC_OK;
}

int ldc_init_state(ldc_t* h) {
    if (!h) return fail(LDC_ERR_INVALID, "ldc_init_state: null handle");
    if (!h->cases_set) return fail(LDC_ERR_STATE, "ldc_init_state: ldc_set_cases has not been called");
    LDC_CUDA(cudaSetDevice(h->device));
    const int nfull = h->n1x * h->n1y;
    k_init_state<<<dim3((nfull + 255) / 256, h->B), 256, 0, h->stream>>>(h->u, h->v, h->p, h->ulid, h->vlid, h->n1x, h->n1y);
    h->launches += 1;
    LDC_CUDA(cudaGetLastError());
    h->canon_valid = true;
    h->packed_valid = 0;
    return LDC_OK;
}

int ldc_residual(ldc_t* h, double* Ru, double* Rv, double* Rp, double* norms) {
    int rc = check_ready(h, "ldc_residual");
    if (rc) return rc;
    LDC_CUDA(cudaSetDevice(h->device));
    if ((rc = ensure_canonical(h))) return rc;
    launch_residual(h, h->u, h->v, h->p, nullptr);
    LDC_CUDA(cudaGetLastError());
    const size_t nf = (size_t)h->n1x * h->n1y, ni = (size_t)h->mx * h->my;
    if ((rc = copy_out(h, Ru, h->Ru, h->B * nf * sizeof(double)))) return rc;
    if ((rc = copy_out(h, Rv, h->Rv, h->B * nf * sizeof(double)))) return rc;
    if ((rc = copy_out(h, Rp, h->Rp, h->B * ni * sizeof(double)))) return rc;
    if (norms) {
        double* d = h->ptmp;  // scratch [B][3] (B * n1x * my >= 3 B)
        k_residual_norms<<<h->B, RED_THREADS, 0, h->stream>>>(h->Ru, h->Rv, h->Rp, (int)nf, (int)ni, d);
        h->launches += 1;
        LDC_CUDA(cudaGetLastError());
        if ((rc = copy_out(h, norms, d, (size_t)h->B * 3 * sizeof(double)))) return rc;
    }
    return LDC_OK;
}

}  // extern "C"

namespace ldc {

__global__ void k_export_status(const CavityStatus* __restrict__ st, int B, long long* iters, int* conv, int* divg,
                                double* rel, double* uv) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const CavityStatus s = st[b];
    iters[b] = s.iters;
    conv[b] = s.converged;
    divg[b] = s.diverged;
    rel[b] = (s.v_res > s.u_res) ? s.v_res : s.u_res;
    uv[2 * b] = s.u_res;
    uv[2 * b + 1] = s.v_res;
}

// CTAs of the resident kernel that are resident on the whole chip at once for this handle's shape.
template <int MP, int WM, int WN>
static int resident_slots_t(ldc_handle* h) {
    using C = ResCfg<MP, WM, WN>;
    const size_t bytes = C::smem_doubles(h->square_ops) * sizeof(double);
    LDC_CUDA(cudaFuncSetAttribute(k_resident_rk4<MP, WM, WN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int per_sm = 0;
    LDC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_resident_rk4<MP, WM, WN>, C::THREADS, bytes));
    h->res_slots = std::max(1, per_sm) * h->num_sms;
    return LDC_OK;
}

static int resident_slots(ldc_handle* h) {
    switch (h->MP) {
        case 16: return resident_slots_t<16, 2, 2>(h);
        case 32: return resident_slots_t<32, 2, 2>(h);
        case 48: return resident_slots_t<48, 2, 3>(h);
        case 64: return resident_slots_t<64, 2, 4>(h);
    }
    return fail(LDC_ERR_UNSUPPORTED, "resident path: unsupported padded size");
}

template <int MP, int WM, int WN>
static int launch_resident_t(ldc_handle* h, const ResParams& P) {
    using C = ResCfg<MP, WM, WN>;
    const size_t bytes = C::smem_doubles(h->square_ops) * sizeof(double);
    LDC_CUDA(cudaFuncSetAttribute(k_resident_rk4<MP, WM, WN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    // persistent CTAs, as many as fit the chip at once, fed by the cavity queue
    k_queue_init<<<1, 1024, 0, h->stream>>>(h->status, h->B, P.launch_end, h->q_data, h->q_seq, h->q_ctl);
    k_resident_rk4<MP, WM, WN><<<std::min(h->B, h->res_slots), C::THREADS, bytes, h->stream>>>(P);
    h->launches += 2;
    LDC_CUDA(cudaGetLastError());
    return LDC_OK;
}

template <int HP, int NT>
static int eo_slots_t(ldc_handle* h) {
    using C = EoCfg<HP, NT>;
    const size_t bytes = C::smem_doubles() * sizeof(double);
    LDC_CUDA(cudaFuncSetAttribute(k_resident_eo<HP, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int per_sm = 0;
    LDC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_resident_eo<HP, NT>, C::THREADS, bytes));
    h->res_slots = std::max(1, per_sm) * h->num_sms;
    return LDC_OK;
}

static int eo_slots(ldc_handle* h) {
    switch (h->HP) {
        case 8: return eo_slots_t<8, 1>(h);
        case 16: return eo_slots_t<16, 1>(h);
        case 24: return eo_slots_t<24, 1>(h);
        case 32: return eo_slots_t<32, 2>(h);
    }
    return fail(LDC_ERR_UNSUPPORTED, "even-odd resident path: unsupported padded size");
}

template <int HP, int NT>
static int launch_eo_t(ldc_handle* h, const ResParams& P) {
    using C = EoCfg<HP, NT>;
    const size_t bytes = C::smem_doubles() * sizeof(double);
    k_queue_init<<<1, 1024, 0, h->stream>>>(h->status, h->B, P.launch_end, h->q_data, h->q_seq, h->q_ctl);
    k_resident_eo<HP, NT><<<std::min(h->B, h->res_slots), C::THREADS, bytes, h->stream>>>(P);
    h->launches += 2;
    LDC_CUDA(cudaGetLastError());
    return LDC_OK;
}

static int launch_eo(ldc_handle* h, const ResParams& P) {
    switch (h->HP) {
        case 8: return launch_eo_t<8, 1>(h, P);
        case 16: return launch_eo_t<16, 1>(h, P);
        case 24: return launch_eo_t<24, 1>(h, P);
        case 32: return launch_eo_t<32, 2>(h, P);
    }
    return fail(LDC_ERR_UNSUPPORTED, "even-odd resident path: unsupported padded size");
}

template <int HP>
static int eo_split_slots_t(ldc_handle* h) {
    using C = EoSplitCfg<HP>;
    const size_t bytes = C::smem_doubles() * sizeof(double);
    LDC_CUDA(cudaFuncSetAttribute(k_resident_eo_split<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int per_sm = 0;
    LDC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_resident_eo_split<HP>, C::THREADS, bytes));
    h->res_slots = std::max(1, per_sm) * h->num_sms;
    return LDC_OK;
}
static int eo_split_slots(ldc_handle* h) {
    switch (h->HP) {
        case 8: return eo_split_slots_t<8>(h);
        case 16: return eo_split_slots_t<16>(h);
        case 24: return eo_split_slots_t<24>(h);
    }
    return fail(LDC_ERR_UNSUPPORTED, "equation-split even-odd path: unsupported padded size");
}
template <int HP>
static int launch_eo_split_t(ldc_handle* h, const ResParams& P) {
    using C = EoSplitCfg<HP>;
    const size_t bytes = C::smem_doubles() * sizeof(double);
    k_queue_init<<<1, 1024, 0, h->stream>>>(h->status, h->B, P.launch_end, h->q_data, h->q_seq, h->q_ctl);
    k_resident_eo_split<HP><<<std::min(h->B, h->res_slots), C::THREADS, bytes, h->stream>>>(P);
    h->launches += 2;
    LDC_CUDA(cudaGetLastError());
    return LDC_OK;
}
static int launch_eo_split(ldc_handle* h, const ResParams& P) {
    switch (h->HP) {
        case 8: return launch_eo_split_t<8>(h, P);
        case 16: return launch_eo_split_t<16>(h, P);
        case 24: return launch_eo_split_t<24>(h, P);
    }
    return fail(LDC_ERR_UNSUPPORTED, "equation-split even-odd path: unsupported padded size");
}

static int launch_resident(ldc_handle* h, const ResParams& P) {
    switch (h->MP) {
        case 16: return launch_resident_t<16, 2, 2>(h, P);
        case 32: return launch_resident_t<32, 2, 2>(h, P);
        case 48: return launch_resident_t<48, 2, 3>(h, P);
        case 64: return launch_resident_t<64, 2, 4>(h, P);
    }
    return fail(LDC_ERR_UNSUPPORTED, "resident path: unsupported padded size");
}

// A launch of the streamed path's kernel chain: with pdl, the kernel may start while its predecessor in the stream drains
// (programmatic dependent launch; the kernels order themselves with griddepcontrol.wait, stream_eo.cuh).
template <typename K>
static cudaError_t launch_chain(bool pdl, K kernel, dim3 grid, int threads, size_t smem, cudaStream_t stream,
                                const StreamParams& P) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, P);
}

// Shared driver of ldc_step / ldc_solve / ldc_solve_sampled: iterations [start, stop) of a solve whose limit is max_iter,
// until every cavity is inactive.  start = 0 starts a solve (status reset); start > 0 continues from the status the
// previous call left (every active cavity has taken exactly `start` iterations).
// n_active: cavities known to be still iterating (a hand-over from another kernel in mid-solve), else -1 = the whole batch.
static int run_steps(ldc_handle* h, const double* tol_dev, long long start, long long stop, long long max_iter, int warmup,
                     int freeze, int nan_exit, int n_active = -1) {
    if (start == 0) {
        k_reset_status<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B);
        h->launches += 1;
    }
    int rc;
    const int kernel = pick_kernel(h);
    if (kernel < 0)
        return fail(LDC_ERR_UNSUPPORTED, "the requested resident path cannot run this shape (shared memory, or x / y "
                                         "operators differ for the even-odd kernel)");
    const bool resident = kernel == 1 || kernel == 2 || kernel == 4, eo = kernel == 2 || kernel == 4, split = kernel == 4;
    if (kernel == 5) {
        // persistent TMA-fed streamed path (stream_tma.cuh): teams of ntile^2 co-resident CTAs, one cooperative launch per
        // batch of steps
        const int ntile = (eo_half(h->mx) + ST_T - 1) / ST_T;
        // Tile shape: 16 x 16 per CTA (k_stream_tma: a cavity over ntile^2 SMs, lowest latency per step) for few cavities;
        // 16 x 32 per CTA (k_stream_tma_wide: half the CTAs per cavity, the per-stage synchronisation paid once per twice
        // the DMMA work) once the batch would queue up behind the larger teams.  LDC_STREAM_CW=1 / 2 forces a shape.
        // (measured, B200: N = 128 B = 64 196 k -> 247 k steps/s; with an odd number of 16-wide tiles per block edge the wide
        // tiles of the last column are half padding: N = 96 B = 64 399 k -> 384 k, so those grids keep the 16 x 16 shape)
        // The shape is chosen per launch from the number of cavities still iterating (both shapes work on the same records).
        // Third shape (k_cluster_eo, cluster_eo.cuh): on the 32-wide parity blocks, once every cavity still iterating can
        // have a cluster of 8 CTAs, the state stays in the cluster's shared memory and the stages exchange it there.
        bool wide = false, cluster = false;
        int team_size = 0, teams = 0;
        size_t smem = 0;
        const void* kfun = nullptr;
        auto pick_shape = [&](int nact) -> int {
            int nclusters = 0;
            const int shape = stream_tma_shape(h, nact, &nclusters);
            cluster = shape == 2;
            if (cluster) {
                wide = false;
                team_size = CE_CLUSTER;
                smem = CE_SMEM_DOUBLES * sizeof(double);
                kfun = (const void*)k_cluster_eo;
                teams = std::min(nact, nclusters);
                if (const char* e = getenv("LDC_STREAM_TEAMS")) teams = std::max(1, std::min(teams, atoi(e)));
                return LDC_OK;
            }
            wide = shape == 1;
            team_size = wide ? ntile * ((ntile + 1) / 2) : ntile * ntile;
            smem = (wide ? S2W_SMEM_DOUBLES : S2_SMEM_DOUBLES) * sizeof(double);
            kfun = wide ? (const void*)k_stream_tma_wide : (const void*)k_stream_tma;
            LDC_CUDA(cudaFuncSetAttribute(kfun, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            int per_sm = 0;
            if (wide) LDC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_stream_tma_wide, S2_THREADS, smem));
            else LDC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_stream_tma, S2_THREADS, smem));
            teams = std::min(nact, per_sm * h->num_sms / team_size);
            if (const char* e = getenv("LDC_STREAM_TEAMS")) teams = std::max(1, std::min(teams, atoi(e)));   // A / B measurements
            if (teams < 1) return fail(LDC_ERR_UNSUPPORTED, "streamed TMA path: a team of CTAs does not fit the device");
            return LDC_OK;
        };
        int nact = (n_active > 0 && n_active <= h->B) ? n_active : h->B;
        if ((rc = pick_shape(nact))) return rc;
        const size_t bs = (size_t)h->B * S2_STRIDE;
        if (!h->t_state) {
            if ((rc = dalloc(h, &h->t_state, 3 * bs))) return rc;
            if ((rc = dalloc(h, &h->t_team, 2 * (size_t)h->num_sms))) return rc;
            if (!h->s_partials && (rc = dalloc(h, &h->s_partials, (size_t)h->B * ST_TILES * 8))) return rc;
            if (!h->q_data) {
                if ((rc = dalloc(h, &h->q_data, h->B))) return rc;
                if ((rc = dalloc(h, &h->q_seq, h->B))) return rc;
                if ((rc = dalloc(h, &h->q_ctl, 4))) return rc;
            }
        }
        if ((rc = ensure_packed(h, 4))) return rc;
        LDC_CUDA(cudaEventRecord(h->ev0, h->stream));
        k_s2_pack_ops<<<dim3(48, h->B), 256, 0, h->stream>>>(h->s_ops, h->Re, h->t_state, bs);
        h->launches += 1;
        k_s2_begin<<<h->B, 256, 0, h->stream>>>(h->t_state, h->mx, h->status);   // max|u|, max|v| of the start state
        h->launches += 1;
        S2Params P;
        P.ops = h->s_ops; P.rec = h->t_state; P.buf_stride = bs;
        P.Re = h->Re; P.beta2 = h->beta2; P.CFL = h->CFL; P.lidvel = h->lidvel; P.ulid = h->ulid; P.vlid = h->vlid;
        P.n1 = h->n1x; P.m = h->mx; P.K4 = (eo_half(h->mx) + 3) / 4 * 4; P.ntile = ntile;
        P.status = h->status; P.partials = h->s_partials; P.tol = tol_dev;
        P.warmup = warmup; P.nan_exit = nan_exit; P.freeze = freeze; P.max_iter = max_iter;
        P.dx_min = h->dx_min; P.dy_min = h->dy_min;
        P.q_data = h->q_data; P.q_seq = h->q_seq; P.q_ctl = h->q_ctl; P.B = h->B;
        P.team_bar = reinterpret_cast<unsigned*>(h->t_team);
        P.team_mail = reinterpret_cast<unsigned long long*>(h->t_team + h->num_sms);
        // LDC_S2_DEBUG=<file>: clock samples of CTA 0 for the first stages of the first launch (stream_tma.cuh)
        const char* dbg_path = getenv("LDC_S2_DEBUG");
        long long* dbg = nullptr;
        if (dbg_path) {
            LDC_CUDA(cudaMalloc((void**)&dbg, S2_DBG_STAGES * 16 * sizeof(long long)));
            LDC_CUDA(cudaMemsetAsync(dbg, 0, S2_DBG_STAGES * 16 * sizeof(long long), h->stream));
        }
        P.dbg = dbg;
        h->canon_valid = false;
        const long long per_launch = 512;   // steps per launch (the host looks at the active count between launches)
        for (long long done = start; done < stop; done += per_launch) {
            const long long L = std::min(per_launch, stop - done);
            P.launch_end = done + L;
            // steps per queue visit: the whole launch when every cavity has a team, else ~32 visits per team so that the
            // tail of the launch (teams idling while the last visits finish) stays small
            P.chunk = L;
            if (nact > teams) {
                const long long c = (L * nact + 32LL * teams - 1) / (32LL * teams);
                P.chunk = std::min(L, std::max(4LL, c));
            }
            k_queue_init<<<1, 1024, 0, h->stream>>>(h->status, h->B, P.launch_end, h->q_data, h->q_seq, h->q_ctl);
            LDC_CUDA(cudaMemsetAsync(h->t_team, 0, 2 * (size_t)h->num_sms * sizeof(double), h->stream));
            void* args[] = {(void*)&P};
            if (cluster) {
                // clusters need no co-residency with each other (a cluster waiting at the queue only waits for running ones)
                cudaLaunchConfig_t cfg = {};
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = CE_CLUSTER;
                at[0].val.clusterDim.y = 1;
                at[0].val.clusterDim.z = 1;
                cfg.attrs = at;
                cfg.numAttrs = 1;
                cfg.gridDim = dim3(teams * CE_CLUSTER);
                cfg.blockDim = dim3(CE_THREADS);
                cfg.dynamicSmemBytes = smem;
                cfg.stream = h->stream;
                LDC_CUDA(cudaLaunchKernelExC(&cfg, kfun, args));
            } else {
                LDC_CUDA(cudaLaunchCooperativeKernel(kfun, dim3(teams * team_size), dim3(S2_THREADS), args, smem,
                                                     h->stream));
            }
            h->launches += 2;
            if (dbg) {
                std::vector<long long> hd(S2_DBG_STAGES * 16);
                LDC_CUDA(cudaMemcpyAsync(hd.data(), dbg, hd.size() * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
                LDC_CUDA(cudaStreamSynchronize(h->stream));
                if (FILE* f = fopen(dbg_path, "w")) {
                    for (int st = 0; st < S2_DBG_STAGES; ++st) {
                        for (int q = 0; q < 16; ++q) fprintf(f, "%lld ", hd[st * 16 + q]);
                        fprintf(f, "\n");
                    }
                    fclose(f);
                }
                cudaFree(dbg);
                dbg = nullptr;
                P.dbg = nullptr;
            }
            if (done + per_launch < stop) {
                int na = 0;
                if ((rc = count_active(h, &na))) return rc;
                if (na == 0) break;
                if (na != nact) {
                    nact = na;
                    if ((rc = pick_shape(nact))) return rc;
                }
            }
        }
        if (dbg) cudaFree(dbg);
        // every cavity's record back into buffer 0
        k_s2_normalise<<<dim3(64, h->B), 256, 0, h->stream>>>(h->t_state, bs, h->status);
        k_stream_clear_pad<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B);
        h->launches += 2;
    } else if (kernel == 3) {
        if (!h->s_state) {
            if ((rc = dalloc(h, &h->s_state, 3 * (size_t)h->B * h->s_stride))) return rc;
            if ((rc = dalloc(h, &h->s_partials, (size_t)h->B * ST_TILES * 8))) return rc;
            if ((rc = dalloc(h, &h->s_opsc, (size_t)h->B * 6 * ST_BLK))) return rc;
        }
        if ((rc = ensure_packed(h, 3))) return rc;
        // few CTAs (one per SM): double-buffer the chunk panels inside the CTA; many: two CTAs per SM overlap each other
        // one CTA per SM, double-buffered panels + split-K warps: for few CTAs, and (measured, N = 128 B = 64: 153 k ->
        // 171 k steps/s; N = 96: 355 k -> 341 k) whenever all 16 tiles of the padded blocks are in use
        const int trow = (eo_half(h->mx) + ST_T - 1) / ST_T;   // tiles of a block row that hold data
        bool db = (long long)h->B * trow * trow <= h->num_sms || eo_half(h->mx) > 48;
        if (const char* e = getenv("LDC_STREAM_DB")) db = atoi(e) != 0;   // A / B measurements of the two shapes
        const size_t smem = st_smem_doubles(db) * sizeof(double);
        if (db) LDC_CUDA(cudaFuncSetAttribute(k_stream_stage<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        else LDC_CUDA(cudaFuncSetAttribute(k_stream_stage<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        LDC_CUDA(cudaEventRecord(h->ev0, h->stream));
        StreamParams P;
        P.ops = h->s_ops;
        P.opsc = h->s_opsc;
        k_stream_scale_ops<<<dim3(32, h->B), 256, 0, h->stream>>>(h->s_ops, h->Re, h->s_opsc);
        h->launches += 1;
        const size_t bs = (size_t)h->B * h->s_stride;
        P.buf[0] = h->s_state; P.buf[1] = h->s_state + bs; P.buf[2] = h->s_state + 2 * bs;
        P.stride = h->s_stride;
        P.Re = h->Re; P.beta2 = h->beta2; P.CFL = h->CFL; P.lidvel = h->lidvel; P.ulid = h->ulid; P.vlid = h->vlid;
        P.dt = h->dt; P.n1 = h->n1x; P.m = h->mx; P.K4 = (eo_half(h->mx) + 3) / 4 * 4;
        P.status = h->status; P.partials = h->s_partials; P.freeze = freeze;
        P.tol = tol_dev; P.warmup = warmup; P.nan_exit = nan_exit; P.max_iter = max_iter;
        P.dx_min = h->dx_min; P.dy_min = h->dy_min;
        k_stream_begin<<<h->B, 256, 0, h->stream>>>(h->s_state, h->s_stride, h->mx, h->Re, h->beta2, h->CFL, h->lidvel,
                                                    h->dx_min, h->dy_min, h->dt, h->status);
        h->launches += 1;
        h->canon_valid = false;
        const long long chunk = 256;
        const char* env_pdl = getenv("LDC_STREAM_PDL");   // 0: plain stream-ordered launches (A / B measurements)
        const bool pdl = !(env_pdl && atoi(env_pdl) == 0);
        // a continued solve (start > 0) finds the boundary lines as its own last step left them: no "first step after a
        // load" there, so that a solve cut into segments takes bitwise the steps of an uncut one
        for (long long done = start; done < stop; ++done) {
            P.first = done == 0 ? 1 : 0;
            for (int stg = 0; stg < 4; ++stg) {
                P.stage = stg;
                P.bc_edges = (done == 0 && stg == 0) ? 0 : 1;
                if (db) LDC_CUDA(launch_chain(pdl, k_stream_stage<true>, dim3(ST_TILES, h->B), 2 * ST_THREADS, smem, h->stream, P));
                else LDC_CUDA(launch_chain(pdl, k_stream_stage<false>, dim3(ST_TILES, h->B), ST_THREADS, smem, h->stream, P));
            }
            h->launches += 4;
            LDC_CUDA(launch_chain(pdl, k_stream_end, dim3(h->B), ST_THREADS, 0, h->stream, P));
            h->launches += 1;
            if ((done + 1 - start) % chunk == 0 && done + 1 < stop) {
                LDC_CUDA(cudaGetLastError());
                int na = 0;
                if ((rc = count_active(h, &na))) return rc;
                if (na == 0) break;
            }
        }
        // every cavity's record back into buffer 0
        k_stream_normalise<<<dim3(32, h->B), 256, 0, h->stream>>>(h->s_state, bs, h->s_stride, h->status);
        k_stream_clear_pad<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B);
        h->launches += 2;
    } else if (resident) {
        if ((rc = ensure_packed(h, eo ? 2 : 1))) return rc;
        LDC_CUDA(cudaEventRecord(h->ev0, h->stream));  // device time of the stepping kernels only
        ResParams P;
        P.ops = eo ? h->e_ops : h->r_ops; P.square = h->square_ops ? 1 : 0;
        P.state = eo ? h->e_state : h->r_state; P.state_stride = eo ? h->e_state_stride : h->r_state_stride;
        P.Re = h->Re; P.beta2 = h->beta2; P.CFL = h->CFL; P.lidvel = h->lidvel; P.ulid = h->ulid; P.vlid = h->vlid;
        P.n1 = h->n1x; P.m = h->mx; P.K4 = ((eo ? eo_half(h->mx) : h->mx) + 3) / 4 * 4;
        P.dx_min = h->dx_min; P.dy_min = h->dy_min;
        P.status = h->status; P.tol = tol_dev; P.warmup = warmup; P.nan_exit = nan_exit; P.freeze = freeze;
        P.max_iter = max_iter;
        P.q_data = h->q_data; P.q_seq = h->q_seq; P.q_ctl = h->q_ctl; P.B = h->B;
        if ((rc = split ? eo_split_slots(h) : (eo ? eo_slots(h) : resident_slots(h)))) return rc;
        // steps per launch (the host looks at the active count between launches)
        const long long chunk = h->mx >= 49 ? 1024 : (h->mx >= 33 ? 2048 : (h->mx >= 17 ? 4096 : 16384));
        const char* env_tail = getenv("LDC_TAIL_SWITCH");
        const bool tail_switch = kernel == 2 && h->path == LDC_PATH_AUTO && h->HP == 32 && eo_half(h->mx) > 24 && stream_tma_ok(h) &&
                                 !(env_tail && atoi(env_tail) == 0) && stream_auto(h) == 5;
        h->canon_valid = false;
        for (long long done = start; done < stop; done += chunk) {
            const long long L = std::min(chunk, stop - done);
            P.launch_end = done + L;
            // steps per queue visit: the whole launch when every cavity has its own CTA slot, else ~64 visits per slot so
            // that the tail of the launch (slots idling while the last visits finish) stays near 1.5 %
            P.chunk = L;
            if (h->B > h->res_slots) {
                const long long c = (L * h->B + 64LL * h->res_slots - 1) / (64LL * h->res_slots);
                P.chunk = std::min(L, std::max(4LL, c));
            }
            if ((rc = split ? launch_eo_split(h, P) : (eo ? launch_eo(h, P) : launch_resident(h, P)))) return rc;
            if (done + chunk < stop) {
                int na = 0;
                if ((rc = count_active(h, &na))) return rc;
                if (na == 0) break;
                // The tail of a to-tolerance solve: once every cavity still iterating can have two (then four) SMs, the
                // streamed kernel takes over -- a cavity's chain of dependent steps is what is left, and on the largest
                // resident grids it runs 1.2 x (two SMs) to 1.9 x (four) faster there.  AUTO only (a forced path stays put,
                // so that runs which must agree bit for bit can); LDC_TAIL_SWITCH=0 turns it off.
                if (tail_switch && 2 * na <= h->num_sms) {
                    LDC_CUDA(cudaEventRecord(h->ev1, h->stream));
                    LDC_CUDA(cudaEventSynchronize(h->ev1));
                    float ms1 = 0;
                    LDC_CUDA(cudaEventElapsedTime(&ms1, h->ev0, h->ev1));
                    const int saved = h->path;
                    h->path = LDC_PATH_STREAM_TMA;
                    rc = run_steps(h, tol_dev, done + L, stop, max_iter, warmup, freeze, nan_exit, na);
                    h->path = saved;
                    h->last_ms += ms1;
                    return rc;
                }
            }
        }
    } else {
        if ((rc = ensure_canonical(h))) return rc;
        LDC_CUDA(cudaEventRecord(h->ev0, h->stream));
        h->packed_valid = 0;
        const long long chunk = 64;
        for (long long done = start; done < stop; done += chunk) {
            const long long n = std::min(chunk, stop - done);
            for (long long s = 0; s < n; ++s) launch_tiled_step(h, freeze, warmup, max_iter, nan_exit, tol_dev);
            LDC_CUDA(cudaGetLastError());
            if (done + chunk < stop) {
                int na = 0;
                if ((rc = count_active(h, &na))) return rc;
                if (na == 0) break;
            }
        }
    }
    LDC_CUDA(cudaEventRecord(h->ev1, h->stream));
    LDC_CUDA(cudaEventSynchronize(h->ev1));
    float ms = 0;
    LDC_CUDA(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    h->last_ms = ms;
    LDC_CUDA(cudaGetLastError());
    return LDC_OK;
}

// Iterations [start_iter, end) on the canonical state with their records into h->hist rows slot0, slot0 + 1, ...:
// one launch for the lot where the grid fits one SM's shared memory (history_fused.cuh), else a launch per RK stage.
static int hist_iterations(ldc_handle* h, long long start_iter, long long end, long long max_iter, int warmup, int freeze,
                           int nan_exit, long long slot0) {
    int rc;
    const int nfull = h->n1x * h->n1y, ninner = h->mx * h->my;
    const size_t fused_bytes = hist_fused_smem_doubles(h->n1x, h->n1y) * sizeof(double);
    int max_optin = 0;
    LDC_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
    const char* env_fused = getenv("LDC_HISTORY_FUSED");
    const bool fused = fused_bytes <= (size_t)max_optin && !(env_fused && atoi(env_fused) == 0) && h->path != LDC_PATH_TILED;
    if (fused) {
        LDC_CUDA(cudaFuncSetAttribute(k_history_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fused_bytes));
        HistFusedParams F;
        F.u = h->u; F.v = h->v; F.p = h->p;
        F.Dx = h->Dx; F.Dxx = h->Dxx; F.DyT = h->DyT; F.DyyT = h->DyyT; F.Ix = h->Ix; F.IyT = h->IyT;
        F.Re = h->Re; F.beta2 = h->beta2; F.CFL = h->CFL; F.lidvel = h->lidvel; F.ulid = h->ulid; F.vlid = h->vlid;
        F.W = h->Wq; F.tol = h->tol; F.status = h->status; F.hist = h->hist + (size_t)slot0 * HIST_W; F.cap = h->hist_cap;
        F.start_iter = start_iter; F.end_iter = end; F.max_iter = max_iter;
        F.warmup = warmup; F.nan_exit = nan_exit; F.freeze = freeze;
        F.n1x = h->n1x; F.n1y = h->n1y; F.dx_min = h->dx_min; F.dy_min = h->dy_min;
        k_history_fused<<<h->B, HF_THREADS, fused_bytes, h->stream>>>(F);
        h->launches += 1;
        LDC_CUDA(cudaGetLastError());
        return LDC_OK;
    }
    for (long long it = start_iter; it < end; ++it) {
        launch_tiled_step(h, freeze, warmup, max_iter, nan_exit, h->tol);
        const long long slot = slot0 + it - start_iter;
        // h->Ru, Rv, Rp still hold the last stage's residual (what the reference's arrays.R_* hold after step())
        k_hist_norms<<<h->B, RED_THREADS, 0, h->stream>>>(h->Ru, h->Rv, h->Rp, nfull, ninner, h->status, it, slot, h->hist_cap,
                                                         h->hist);
        k_vorticity<<<grid2d(h->n1x, h->n1y, h->B), kBlock2d, 0, h->stream>>>(h->u, h->v, h->Dx, h->DyT, h->pfull, h->n1x,
                                                                              h->n1y);
        k_hist_diag<<<h->B, RED_THREADS, 0, h->stream>>>(h->u, h->v, h->pfull, h->Dx, h->DyT, h->Wq, h->n1x, h->n1y, h->status,
                                                        it, slot, h->hist_cap, h->hist);
        h->launches += 3;
        if ((it + 1 - start_iter) % 256 == 0 && it + 1 < end) {
            LDC_CUDA(cudaGetLastError());
            int na = 0;
            if ((rc = count_active(h, &na))) return rc;
            if (na == 0) break;
        }
    }
    return LDC_OK;
}

// Export per-cavity status through scratch device buffers carved out of h->Ru.
static int export_status(ldc_handle* h, int64_t* iters, int32_t* conv, int32_t* divg, double* rel, double* uv) {
    char* base = reinterpret_cast<char*>(h->Ru);
    const size_t B = h->B;
    long long* d_it = reinterpret_cast<long long*>(base);
    double* d_rel = reinterpret_cast<double*>(base + 8 * B);
    double* d_uv = reinterpret_cast<double*>(base + 16 * B);
    int* d_cv = reinterpret_cast<int*>(base + 32 * B);
    int* d_dv = reinterpret_cast<int*>(base + 36 * B);
    k_export_status<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B, d_it, d_cv, d_dv, d_rel, d_uv);
    h->launches += 1;
    LDC_CUDA(cudaGetLastError());
    int rc;
    if ((rc = copy_out(h, iters, d_it, B * 8))) return rc;
    if ((rc = copy_out(h, conv, d_cv, B * 4))) return rc;
    if ((rc = copy_out(h, divg, d_dv, B * 4))) return rc;
    if ((rc = copy_out(h, rel, d_rel, B * 8))) return rc;
    if ((rc = copy_out(h, uv, d_uv, B * 16))) return rc;
    return LDC_OK;
}

}  // namespace ldc

extern "C" {

int ldc_step(ldc_t* h, int64_t n_steps, int freeze_pressure, double* rel) {
    int rc = check_ready(h, "ldc_step");
    if (rc) return rc;
    if (n_steps < 0) return fail(LDC_ERR_INVALID, "ldc_step: n_steps must be >= 0");
    LDC_CUDA(cudaSetDevice(h->device));
    if (n_steps == 0) return LDC_OK;
    if ((rc = run_steps(h, nullptr, 0, n_steps, n_steps, 0, freeze_pressure ? 1 : 0, 0))) return rc;
    if (rel) return export_status(h, nullptr, nullptr, nullptr, nullptr, rel);
    return LDC_OK;
}

int ldc_solve(ldc_t* h, const double* tol, int64_t max_iter, int warmup, int freeze_pressure, int nan_exit, int64_t* iters,
              int32_t* converged, int32_t* diverged, double* rel_res) {
    int rc = check_ready(h, "ldc_solve");
    if (rc) return rc;
    if (!tol || max_iter < 0 || warmup < 0) return fail(LDC_ERR_INVALID, "ldc_solve: bad argument");
    LDC_CUDA(cudaSetDevice(h->device));
    if ((rc = copy_in(h, h->tol, tol, h->B))) return rc;
    if (max_iter > 0) {
        if ((rc = run_steps(h, h->tol, 0, max_iter, max_iter, warmup, freeze_pressure ? 1 : 0, nan_exit ? 1 : 0))) return rc;
    } else {
        k_reset_status<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B);
        h->launches += 1;
    }
    return export_status(h, iters, converged, diverged, rel_res, nullptr);
}

int ldc_solve_segment(ldc_t* h, const double* tol, int64_t start_iter, int64_t stop_iter, int64_t max_iter, int warmup,
                      int freeze_pressure, int nan_exit, int64_t* iters, int32_t* converged, int32_t* diverged,
                      double* rel_res) {
    int rc = check_ready(h, "ldc_solve_segment");
    if (rc) return rc;
    if (!tol || start_iter < 0 || stop_iter < start_iter || max_iter < stop_iter || warmup < 0)
        return fail(LDC_ERR_INVALID, "ldc_solve_segment: need tol and 0 <= start_iter <= stop_iter <= max_iter");
    LDC_CUDA(cudaSetDevice(h->device));
    if ((rc = copy_in(h, h->tol, tol, h->B))) return rc;
    if (stop_iter > start_iter) {
        if ((rc = run_steps(h, h->tol, start_iter, stop_iter, max_iter, warmup, freeze_pressure ? 1 : 0, nan_exit ? 1 : 0)))
            return rc;
    } else if (start_iter == 0) {
        k_reset_status<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B);
        h->launches += 1;
    }
    return export_status(h, iters, converged, diverged, rel_res, nullptr);
}

int ldc_solve_history(ldc_t* h, const double* tol, int64_t start_iter, int64_t max_iter, int warmup, int freeze_pressure,
                      int nan_exit, const double* W, double* hist, int64_t hist_cap, int64_t* iters, int32_t* converged,
                      int32_t* diverged, double* rel_res) {
    int rc = check_ready(h, "ldc_solve_history");
    if (rc) return rc;
    if (!tol || !W || !hist || hist_cap < 1 || start_iter < 0 || max_iter < start_iter || warmup < 0)
        return fail(LDC_ERR_INVALID, "ldc_solve_history: bad argument");
    LDC_CUDA(cudaSetDevice(h->device));
    const size_t nf = (size_t)h->n1x * h->n1y, B = h->B;
    if (!h->Wq && (rc = dalloc(h, &h->Wq, nf))) return rc;
    if (h->hist_cap < (size_t)hist_cap) {
        if (h->hist) cudaFree(h->hist);
        h->hist = nullptr;
        h->hist_cap = 0;
        if ((rc = dalloc(h, &h->hist, B * (size_t)hist_cap * HIST_W))) return rc;
        h->hist_cap = (size_t)hist_cap;
    }
    if ((rc = copy_in(h, h->tol, tol, B))) return rc;
    if ((rc = copy_in(h, h->Wq, W, nf))) return rc;
    if ((rc = ensure_canonical(h))) return rc;
    h->packed_valid = 0;
    LDC_CUDA(cudaMemsetAsync(h->hist, 0, B * h->hist_cap * HIST_W * sizeof(double), h->stream));
    if (start_iter == 0) {
        k_reset_status<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B);
        h->launches += 1;
    }
    LDC_CUDA(cudaEventRecord(h->ev0, h->stream));
    const long long end = std::min<long long>(max_iter, start_iter + hist_cap);
    if ((rc = hist_iterations(h, start_iter, end, max_iter, warmup, freeze_pressure ? 1 : 0, nan_exit ? 1 : 0, 0))) return rc;
    LDC_CUDA(cudaEventRecord(h->ev1, h->stream));
    LDC_CUDA(cudaEventSynchronize(h->ev1));
    float ms = 0;
    LDC_CUDA(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    h->last_ms = ms;
    LDC_CUDA(cudaGetLastError());
    if (h->hist_cap == (size_t)hist_cap) {
        if ((rc = copy_out(h, hist, h->hist, B * h->hist_cap * HIST_W * sizeof(double)))) return rc;
    } else {
        for (size_t b = 0; b < B; ++b)
            if ((rc = copy_out(h, hist + b * (size_t)hist_cap * HIST_W, h->hist + b * h->hist_cap * HIST_W,
                               (size_t)hist_cap * HIST_W * sizeof(double))))
                return rc;
    }
    return export_status(h, iters, converged, diverged, rel_res, nullptr);
}

int ldc_solve_sampled(ldc_t* h, const double* tol, int64_t max_iter, int warmup, int freeze_pressure, int nan_exit,
                      const double* W, int64_t n_rec, const int64_t* rec_iter, double* hist, int64_t* iters, int32_t* converged,
                      int32_t* diverged, double* rel_res) {
    int rc = check_ready(h, "ldc_solve_sampled");
    if (rc) return rc;
    if (!tol || !W || !hist || !rec_iter || n_rec < 1 || max_iter < 1 || warmup < 0)
        return fail(LDC_ERR_INVALID, "ldc_solve_sampled: bad argument");
    for (int64_t k = 0; k < n_rec; ++k)
        if (rec_iter[k] < 0 || rec_iter[k] >= max_iter || (k && rec_iter[k] <= rec_iter[k - 1]))
            return fail(LDC_ERR_INVALID, "ldc_solve_sampled: rec_iter must be strictly increasing inside [0, max_iter)");
    LDC_CUDA(cudaSetDevice(h->device));
    const size_t nf = (size_t)h->n1x * h->n1y, ni = (size_t)h->mx * h->my, B = h->B;
    const size_t st_doubles = (B * sizeof(CavityStatus) + 7) / 8;
    if (!h->Wq && (rc = dalloc(h, &h->Wq, nf))) return rc;
    if (!h->fork && (rc = dalloc(h, &h->fork, st_doubles + B * (2 * nf + ni)))) return rc;
    if (h->hist_cap < (size_t)n_rec) {
        if (h->hist) cudaFree(h->hist);
        h->hist = nullptr;
        h->hist_cap = 0;
        if ((rc = dalloc(h, &h->hist, B * (size_t)n_rec * HIST_W))) return rc;
        h->hist_cap = (size_t)n_rec;
    }
    if ((rc = copy_in(h, h->tol, tol, B))) return rc;
    if ((rc = copy_in(h, h->Wq, W, nf))) return rc;
    LDC_CUDA(cudaMemsetAsync(h->hist, 0, B * h->hist_cap * HIST_W * sizeof(double), h->stream));
    if (!h->ev2) LDC_CUDA(cudaEventCreate(&h->ev2));
    if (!h->ev3) LDC_CUDA(cudaEventCreate(&h->ev3));
    LDC_CUDA(cudaEventRecord(h->ev2, h->stream));
    const int fz = freeze_pressure ? 1 : 0, ne = nan_exit ? 1 : 0;
    double *fu = h->fork + st_doubles, *fv = fu + B * nf, *fp = fv + B * nf;
    long long cur = 0;
    bool started = false;
    for (int64_t k = 0; k < n_rec; ++k) {
        const long long K = rec_iter[k];
        if (K > cur) {
            if ((rc = run_steps(h, h->tol, cur, K, max_iter, warmup, fz, ne))) return rc;
            cur = K;
            started = true;
        } else if (!started) {   // a record of iteration 0: the solve starts here
            k_reset_status<<<(h->B + 127) / 128, 128, 0, h->stream>>>(h->status, h->B);
            h->launches += 1;
            started = true;
        }
        if (k % 16 == 15) {   // everything converged already: the remaining records stay zero
            int na = 0;
            if ((rc = count_active(h, &na))) return rc;
            if (na == 0) break;
        }
        // the record step runs on the canonical copy of the state; the stepping kernels' own record (packed_valid != 0)
        // is not touched, so the trajectory is the one ldc_solve takes
        if ((rc = ensure_canonical(h))) return rc;
        const bool canon_is_master = h->packed_valid == 0;
        LDC_CUDA(cudaMemcpyAsync(h->fork, h->status, B * sizeof(CavityStatus), cudaMemcpyDeviceToDevice, h->stream));
        if (canon_is_master) {
            LDC_CUDA(cudaMemcpyAsync(fu, h->u, B * nf * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
            LDC_CUDA(cudaMemcpyAsync(fv, h->v, B * nf * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
            LDC_CUDA(cudaMemcpyAsync(fp, h->p, B * ni * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
        }
        if ((rc = hist_iterations(h, K, K + 1, max_iter, warmup, fz, ne, k))) return rc;
        LDC_CUDA(cudaMemcpyAsync(h->status, h->fork, B * sizeof(CavityStatus), cudaMemcpyDeviceToDevice, h->stream));
        if (canon_is_master) {
            LDC_CUDA(cudaMemcpyAsync(h->u, fu, B * nf * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
            LDC_CUDA(cudaMemcpyAsync(h->v, fv, B * nf * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
            LDC_CUDA(cudaMemcpyAsync(h->p, fp, B * ni * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
        } else {
            h->canon_valid = false;
        }
    }
    if (cur < max_iter) {
        int na = 1;
        if (started && (rc = count_active(h, &na))) return rc;
        if (na > 0 && (rc = run_steps(h, h->tol, cur, max_iter, max_iter, warmup, fz, ne))) return rc;
    }
    LDC_CUDA(cudaEventRecord(h->ev3, h->stream));
    LDC_CUDA(cudaEventSynchronize(h->ev3));
    float ms = 0;
    LDC_CUDA(cudaEventElapsedTime(&ms, h->ev2, h->ev3));
    h->last_ms = ms;   // the whole call: stepping kernels and record steps
    LDC_CUDA(cudaGetLastError());
    for (size_t b = 0; b < B; ++b)   // the device rows are h->hist_cap apart (a buffer of an earlier, longer call is reused)
        if ((rc = copy_out(h, hist + b * (size_t)n_rec * HIST_W, h->hist + b * h->hist_cap * HIST_W,
                           (size_t)n_rec * HIST_W * sizeof(double))))
            return rc;
    return export_status(h, iters, converged, diverged, rel_res, nullptr);
}

int ldc_diagnostics(ldc_t* h, const double* W, double* out) {
    int rc = check_ready(h, "ldc_diagnostics");
    if (rc) return rc;
    if (!W || !out) return fail(LDC_ERR_INVALID, "ldc_diagnostics: null argument");
    LDC_CUDA(cudaSetDevice(h->device));
    if ((rc = ensure_canonical(h))) return rc;
    const size_t nf = (size_t)h->n1x * h->n1y;
    double* dW = h->Rv;          // scratch: one full grid
    double* dOut = h->Rv + nf;   // scratch: [B][3]   (B * nf >= nf + 3 B needs B >= 2 or nf > 3: checked below)
    double* tmpW = nullptr;
    if ((size_t)h->B * nf < nf + 3 * (size_t)h->B) {
        LDC_CUDA(cudaMalloc((void**)&tmpW, (nf + 3 * h->B) * sizeof(double)));
        dW = tmpW;
        dOut = tmpW + nf;
    }
    if ((rc = copy_in(h, dW, W, nf))) return rc;
    k_vorticity<<<grid2d(h->n1x, h->n1y, h->B), kBlock2d, 0, h->stream>>>(h->u, h->v, h->Dx, h->DyT, h->pfull, h->n1x, h->n1y);
    k_diagnostics<<<h->B, RED_THREADS, 0, h->stream>>>(h->u, h->v, h->pfull, h->Dx, h->DyT, dW, dOut, h->n1x, h->n1y);
    h->launches += 2;
    LDC_CUDA(cudaGetLastError());
    rc = copy_out(h, out, dOut, (size_t)h->B * 3 * sizeof(double));
    if (tmpW) {
        cudaStreamSynchronize(h->stream);
        cudaFree(tmpW);
    }
    return rc;
}

int ldc_vortex_metrics(ldc_t* h, const double* Vai, const double* VbiT, const double* Va, const double* VbT,
                       const double* inv_lam, const double* x, const double* y, double* out, double* psi) {
    int rc = check_ready(h, "ldc_vortex_metrics");
    if (rc) return rc;
    if (!Vai || !VbiT || !Va || !VbT || !inv_lam || !x || !y || !out) return fail(LDC_ERR_INVALID, "ldc_vortex_metrics: null argument");
    LDC_CUDA(cudaSetDevice(h->device));
    if ((rc = ensure_canonical(h))) return rc;
    const int n1x = h->n1x, n1y = h->n1y, mx = h->mx, my = h->my, B = h->B;
    const size_t nf = (size_t)n1x * n1y, mm = (size_t)mx * my;
    // shared inputs: Vai Va (mx x mx) | VbiT VbT (my x my) | inv_lam A BT | x y | out
    double* M = nullptr;
    const size_t nM = 2 * (size_t)mx * mx + 2 * (size_t)my * my + mm + (size_t)mx * mx + (size_t)my * my + n1x + n1y + (size_t)B * VM_OUT;
    LDC_CUDA(cudaMalloc((void**)&M, nM * sizeof(double)));
    double *dVai = M, *dVa = dVai + (size_t)mx * mx, *dVbiT = dVa + (size_t)mx * mx, *dVbT = dVbiT + (size_t)my * my,
           *dInv = dVbT + (size_t)my * my, *dA = dInv + mm, *dBT = dA + (size_t)mx * mx, *dx = dBT + (size_t)my * my, *dy = dx + n1x,
           *dOut = dy + n1y;
    rc = copy_in(h, dVai, Vai, (size_t)mx * mx);
    if (!rc) rc = copy_in(h, dVa, Va, (size_t)mx * mx);
    if (!rc) rc = copy_in(h, dVbiT, VbiT, (size_t)my * my);
    if (!rc) rc = copy_in(h, dVbT, VbT, (size_t)my * my);
    if (!rc) rc = copy_in(h, dInv, inv_lam, mm);
    if (!rc) rc = copy_in(h, dx, x, n1x);
    if (!rc) rc = copy_in(h, dy, y, n1y);
    if (rc) { cudaFree(M); return rc; }
    // A = Dxx[1:-1,1:-1] (left operand), BT = Dyy[1:-1,1:-1]^T = DyyT[1:-1,1:-1] (right operand of psi B^T)
    LDC_CUDA(cudaMemcpy2DAsync(dA, mx * sizeof(double), h->Dxx + n1x + 1, n1x * sizeof(double), mx * sizeof(double), mx,
                               cudaMemcpyDeviceToDevice, h->stream));
    LDC_CUDA(cudaMemcpy2DAsync(dBT, my * sizeof(double), h->DyyT + n1y + 1, n1y * sizeof(double), my * sizeof(double), my,
                               cudaMemcpyDeviceToDevice, h->stream));
    // scratch of the tiled path, free outside a step: omega -> pfull, F -> Ru, T -> Rv, G / corrections -> u0, psi_int -> v0,
    // A psi -> Rp, psi B^T -> ptmp, R -> p0; psi on all nodes goes to Ru once F is no longer needed
    double *om = h->pfull, *F = h->Ru, *T = h->Rv, *G = h->u0, *P0 = h->v0, *AP = h->Rp, *PB = h->ptmp, *R = h->p0;
    const int TB = 256;
    const dim3 gm((unsigned)((mm + TB - 1) / TB), B), gf((unsigned)((nf + TB - 1) / TB), B);
    auto solve = [&](const double* rhs, double* dst) {   // dst = Va ((Vai rhs VbiT) * inv) VbT
        k_left_mul<<<grid2d(mx, my, B), kBlock2d, 0, h->stream>>>(dVai, rhs, T, mx, mx, my, nullptr);
        k_right_mul<<<grid2d(mx, my, B), kBlock2d, 0, h->stream>>>(T, dVbiT, dst, mx, my, my, nullptr);
        k_vm_scale<<<gm, TB, 0, h->stream>>>(dst, dInv, (int)mm);
        k_left_mul<<<grid2d(mx, my, B), kBlock2d, 0, h->stream>>>(dVa, dst, T, mx, mx, my, nullptr);
        k_right_mul<<<grid2d(mx, my, B), kBlock2d, 0, h->stream>>>(T, dVbT, dst, mx, my, my, nullptr);
        h->launches += 5;
    };
    k_vorticity<<<grid2d(n1x, n1y, B), kBlock2d, 0, h->stream>>>(h->u, h->v, h->Dx, h->DyT, om, n1x, n1y);
    k_vm_interior_neg<<<gm, TB, 0, h->stream>>>(om, F, n1x, n1y);
    solve(F, P0);
    // one step of iterative refinement against the (non-normal) interior operator: R = F - (A psi + psi B^T)
    k_left_mul<<<grid2d(mx, my, B), kBlock2d, 0, h->stream>>>(dA, P0, AP, mx, mx, my, nullptr);
    k_right_mul<<<grid2d(mx, my, B), kBlock2d, 0, h->stream>>>(P0, dBT, PB, mx, my, my, nullptr);
    k_vm_residual<<<gm, TB, 0, h->stream>>>(F, AP, PB, R, (int)mm);
    solve(R, G);
    double* psi_full = h->Ru;   // F is dead
    k_vm_embed<<<gf, TB, 0, h->stream>>>(P0, G, psi_full, n1x, n1y);
    k_vm_scan<<<B, VM_THREADS, 0, h->stream>>>(psi_full, om, dx, dy, n1x, n1y, dOut);
    h->launches += 7;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) rc = copy_out(h, out, dOut, (size_t)B * VM_OUT * sizeof(double));
    if (e == cudaSuccess && !rc && psi) rc = copy_out(h, psi, psi_full, (size_t)B * nf * sizeof(double));
    cudaStreamSynchronize(h->stream);
    cudaFree(M);
    if (e != cudaSuccess) return fail(LDC_ERR_CUDA, std::string("ldc_vortex_metrics: ") + cudaGetErrorString(e));
    return rc;
}

int ldc_validation_errors(ldc_t* h, int64_t nq, const int32_t* cell, const double* tx, const double* ty, const int32_t* mask,
                          const double* ref_u, const double* ref_v, double* out) {
    int rc = check_ready(h, "ldc_validation_errors");
    if (rc) return rc;
    if (nq < 0 || !out || (nq > 0 && (!cell || !tx || !ty || !mask || !ref_u || !ref_v)))
        return fail(LDC_ERR_INVALID, "ldc_validation_errors: bad argument");
    LDC_CUDA(cudaSetDevice(h->device));
    if ((rc = ensure_canonical(h))) return rc;
    const size_t n = (size_t)std::max<int64_t>(nq, 1), B = h->B;
    // one scratch block: tx ty ref_u ref_v [n doubles each] | partial [B][CHUNKS][5] | out [B][3] | cell mask [n ints each]
    double* d = nullptr;
    const size_t nd = 4 * n + B * VAL_CHUNKS * 5 + B * 3;
    LDC_CUDA(cudaMalloc((void**)&d, nd * sizeof(double) + 2 * n * sizeof(int)));
    double *dtx = d, *dty = d + n, *dru = d + 2 * n, *drv = d + 3 * n, *dpart = d + 4 * n, *dout = dpart + B * VAL_CHUNKS * 5;
    int *dcell = reinterpret_cast<int*>(d + nd), *dmask = dcell + n;
    // host pointers are range-checked here (a bad cell index would read outside the state arrays)
    if (nq > 0 && !is_device_ptr(cell)) {
        const int lim = (h->n1x - 1) * h->n1y;
        for (int64_t q = 0; q < nq; ++q)
            if (mask[q] && (cell[q] < 0 || cell[q] >= lim || cell[q] % h->n1y == h->n1y - 1)) {
                cudaFree(d);
                return fail(LDC_ERR_INVALID, "ldc_validation_errors: cell index outside the grid");
            }
    }
    rc = LDC_OK;
    if (nq > 0) {
        if (!rc) rc = copy_in(h, dtx, tx, nq);
        if (!rc) rc = copy_in(h, dty, ty, nq);
        if (!rc) rc = copy_in(h, dru, ref_u, nq);
        if (!rc) rc = copy_in(h, drv, ref_v, nq);
        const cudaMemcpyKind kind = is_device_ptr(cell) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
        if (!rc && (cudaMemcpyAsync(dcell, cell, nq * sizeof(int), kind, h->stream) != cudaSuccess ||
                    cudaMemcpyAsync(dmask, mask, nq * sizeof(int), kind, h->stream) != cudaSuccess ||
                    cudaStreamSynchronize(h->stream) != cudaSuccess))
            rc = fail(LDC_ERR_CUDA, "ldc_validation_errors: copy of the point plan failed");
    }
    if (!rc) {
        k_val_partial<<<dim3(VAL_CHUNKS, h->B), RED_THREADS, 0, h->stream>>>(h->u, h->v, h->n1x * h->n1y, h->n1y, nq, dcell, dtx,
                                                                           dty, dmask, dru, drv, dpart);
        k_val_final<<<(h->B + 127) / 128, 128, 0, h->stream>>>(dpart, h->B, dout);
        h->launches += 2;
        if (cudaGetLastError() != cudaSuccess) rc = fail(LDC_ERR_CUDA, "ldc_validation_errors: launch failed");
    }
    if (!rc) rc = copy_out(h, out, dout, B * 3 * sizeof(double));
    cudaStreamSynchronize(h->stream);
    cudaFree(d);
    return rc;
}

static int transfer_fields(ldc_t* src, ldc_t* dst, const double* Pxu, const double* Pyu, const double* Pxp, const double* Pyp,
                           const char* who) {
    if (!src || !dst || !Pxu || !Pyu || !Pxp || !Pyp) return fail(LDC_ERR_INVALID, std::string(who) + ": null argument");
    if (src->B != dst->B || src->device != dst->device)
        return fail(LDC_ERR_INVALID, std::string(who) + ": handles must share batch size and device");
    if (!dst->cases_set) return fail(LDC_ERR_STATE, std::string(who) + ": ldc_set_cases has not been called on the target");
    int rc;
    if ((rc = ensure_canonical(src))) return rc;
    // order the target stream after the source stream
    if (src->stream != dst->stream) {
        LDC_CUDA(cudaEventRecord(src->ev1, src->stream));
        LDC_CUDA(cudaStreamWaitEvent(dst->stream, src->ev1, 0));
    }
    ldc_t* h = dst;
    // transfer matrices live in the target's operator scratch (ptmp is big enough only for p; use a temp allocation)
    const size_t sz_xu = (size_t)dst->n1x * src->n1x, sz_yu = (size_t)dst->n1y * src->n1y;
    const size_t sz_xp = (size_t)dst->mx * src->mx, sz_yp = (size_t)dst->my * src->my;
    double* M = nullptr;
    LDC_CUDA(cudaMalloc((void**)&M, (sz_xu + 2 * sz_yu + sz_xp + 2 * sz_yp) * sizeof(double)));
    double *dPxu = M, *dPyu = dPxu + sz_xu, *dPyuT = dPyu + sz_yu, *dPxp = dPyuT + sz_yu, *dPyp = dPxp + sz_xp,
           *dPypT = dPyp + sz_yp;
    rc = copy_in(h, dPxu, Pxu, sz_xu);
    if (!rc) rc = copy_in(h, dPyu, Pyu, sz_yu);
    if (!rc) rc = copy_in(h, dPxp, Pxp, sz_xp);
    if (!rc) rc = copy_in(h, dPyp, Pyp, sz_yp);
    if (rc) { cudaFree(M); return rc; }
    const int T = 256;
    k_transpose<<<(int)((sz_yu + T - 1) / T), T, 0, h->stream>>>(dPyu, dPyuT, dst->n1y, src->n1y);
    k_transpose<<<(int)((sz_yp + T - 1) / T), T, 0, h->stream>>>(dPyp, dPypT, dst->my, src->my);
    // U_dst = Px U_src Py^T: columns first (left multiply), then rows (operators/transfer_operators.py:93-130)
    double* tmp = dst->Ru;  // dst.n1x x src.n1y per cavity; Ru is at least as large when dst is the finer grid
    double* tmp_alloc = nullptr;
    if ((size_t)dst->n1x * src->n1y > (size_t)dst->n1x * dst->n1y) {
        LDC_CUDA(cudaMalloc((void**)&tmp_alloc, (size_t)dst->B * dst->n1x * src->n1y * sizeof(double)));
        tmp = tmp_alloc;
    }
    const double* sf[2] = {src->u, src->v};
    double* df[2] = {dst->u, dst->v};
    for (int f = 0; f < 2; ++f) {
        k_left_mul<<<grid2d(dst->n1x, src->n1y, h->B), kBlock2d, 0, h->stream>>>(dPxu, sf[f], tmp, dst->n1x, src->n1x, src->n1y,
                                                                                 nullptr);
        k_right_mul<<<grid2d(dst->n1x, dst->n1y, h->B), kBlock2d, 0, h->stream>>>(tmp, dPyuT, df[f], dst->n1x, src->n1y,
                                                                                  dst->n1y, nullptr);
    }
    k_left_mul<<<grid2d(dst->mx, src->my, h->B), kBlock2d, 0, h->stream>>>(dPxp, src->p, tmp, dst->mx, src->mx, src->my, nullptr);
    k_right_mul<<<grid2d(dst->mx, dst->my, h->B), kBlock2d, 0, h->stream>>>(tmp, dPypT, dst->p, dst->mx, src->my, dst->my, nullptr);
    h->launches += 8;
    cudaError_t e = cudaGetLastError();
    cudaStreamSynchronize(h->stream);
    cudaFree(M);
    if (tmp_alloc) cudaFree(tmp_alloc);
    if (e != cudaSuccess) return fail(LDC_ERR_CUDA, std::string(who) + ": " + cudaGetErrorString(e));
    dst->canon_valid = true;
    dst->packed_valid = 0;
    return LDC_OK;
}

int ldc_prolongate(ldc_t* coarse, ldc_t* fine, const double* Pxu, const double* Pyu, const double* Pxp, const double* Pyp,
                   int bc_quirk) {
    if (fine) LDC_CUDA(cudaSetDevice(fine->device));
    int rc = transfer_fields(coarse, fine, Pxu, Pyu, Pxp, Pyp, "ldc_prolongate");
    if (rc) return rc;
    const int nfull = fine->n1x * fine->n1y;
    k_transfer_bc<<<dim3((nfull + 255) / 256, fine->B), 256, 0, fine->stream>>>(fine->u, fine->v, fine->ulid, fine->vlid,
                                                                                 fine->lidvel, fine->n1x, fine->n1y,
                                                                                 bc_quirk ? 1 : 0);
    fine->launches += 1;
    LDC_CUDA(cudaGetLastError());
    return LDC_OK;
}

int ldc_restrict(ldc_t* fine, ldc_t* coarse, const double* Rxu, const double* Ryu, const double* Rxp, const double* Ryp) {
    if (coarse) LDC_CUDA(cudaSetDevice(coarse->device));
    int rc = transfer_fields(fine, coarse, Rxu, Ryu, Rxp, Ryp, "ldc_restrict");
    if (rc) return rc;
    const int nfull = coarse->n1x * coarse->n1y;
    k_transfer_bc<<<dim3((nfull + 255) / 256, coarse->B), 256, 0, coarse->stream>>>(coarse->u, coarse->v, coarse->ulid,
                                                                                     coarse->vlid, coarse->lidvel, coarse->n1x,
                                                                                     coarse->n1y, 0);
    coarse->launches += 1;
    LDC_CUDA(cudaGetLastError());
    return LDC_OK;
}

int64_t ldc_launch_count(const ldc_t* h) { return h ? h->launches : 0; }
double ldc_last_kernel_ms(const ldc_t* h) { return h ? h->last_ms : 0.0; }

int ldc_fp64_peak(int device, int which, double* tflops) { return measure_fp64_peak(device, which, tflops); }

}  // extern "C"

// ---- allocation-free diagnostics ------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that may not allocate device
// memory (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same two kernels as
// ldc_diagnostics (k_vorticity, k_diagnostics); the vorticity field they hand over lives in the file-scope array below
// instead of in a handle's buffers.  Every GPU that loads the library holds the array (4 MiB) until the process exits.
constexpr size_t NOALLOC_MAX_DOUBLES = (size_t)LDC_NOALLOC_MAX_DOUBLES;
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_MAX_DOUBLES];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address is looked up once per device, with
// that device current, and it is that address the kernels are given.
constexpr int NOALLOC_MAX_DEVICES = 64;
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};
static std::mutex g_noalloc_mutex;

static int noalloc_scratch(int device, double** out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(LDC_ERR_INVALID, "ldc_diagnostics_noalloc: device index out of range");
    LDC_CUDA(cudaSetDevice(device));
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        double* p = nullptr;
        LDC_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) return fail(LDC_ERR_CUDA, "ldc_diagnostics_noalloc: cudaGetSymbolAddress returned a null address");
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return LDC_OK;
}

extern "C" int ldc_diagnostics_noalloc(int device, void* cuda_stream, int nx, int ny, int batch, const double* d_u,
                                       const double* d_v, const double* d_Dx, const double* d_DyT, const double* d_W,
                                       double* d_out) {
    if (nx < 2 || ny < 2 || batch < 1) return fail(LDC_ERR_INVALID, "ldc_diagnostics_noalloc: nx, ny >= 2 and batch >= 1 required");
    if (!d_u || !d_v || !d_Dx || !d_DyT || !d_W || !d_out) return fail(LDC_ERR_INVALID, "ldc_diagnostics_noalloc: null argument");
    const size_t n1x = (size_t)nx + 1, n1y = (size_t)ny + 1, need = (size_t)batch * n1x * n1y;
    if (batch > 65535 || need > NOALLOC_MAX_DOUBLES)
        return fail(LDC_ERR_INVALID, "ldc_diagnostics_noalloc: batch * (nx+1) * (ny+1) = " + std::to_string(need) +
                                         " doubles exceeds the static scratch of " + std::to_string(NOALLOC_MAX_DOUBLES) +
                                         " (LDC_NOALLOC_MAX_DOUBLES); use ldc_diagnostics, which allocates");
    double* om = nullptr;
    int rc = noalloc_scratch(device, &om);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    // The array keeps the previous call's vorticity.  k_vorticity writes every element that k_diagnostics reads, so nothing
    // here depends on a zero start; the part in use is cleared all the same, so that a stale value cannot be read by accident.
    LDC_CUDA(cudaMemsetAsync(om, 0, need * sizeof(double), st));
    k_vorticity<<<grid2d((int)n1x, (int)n1y, batch), kBlock2d, 0, st>>>(d_u, d_v, d_Dx, d_DyT, om, (int)n1x, (int)n1y);
    k_diagnostics<<<batch, RED_THREADS, 0, st>>>(d_u, d_v, om, d_Dx, d_DyT, d_W, d_out, (int)n1x, (int)n1y);
    LDC_CUDA(cudaGetLastError());
    return LDC_OK;
}
