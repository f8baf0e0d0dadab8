// Micro-benchmark: what an exchange of stage outputs between the CTAs of a thread-block cluster costs on B200 --
//   (a) barrier.cluster arrive.release + wait.acquire alone,
//   (b) st.shared::cluster stores (v2.f64 and scalar f64) of V bytes per CTA to 1 / 3 peers, followed by the barrier,
//   (c) cp.async.bulk shared::cta -> shared::cluster (peer) with mbarrier complete_tx on the peer, 1 / 4 issuing warps,
// for cluster sizes 4, 8, 16, with the shared-memory footprint of the planned kernel, plus how many such clusters the
// chip holds at once.  Design input for a cluster-resident even-odd kernel (state + operators in the cluster's shared memory).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dsmem_bw scripts/dsmem_bw.cu && ./dsmem_bw
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <vector>

__device__ __forceinline__ unsigned sm(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned cluster_rank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ unsigned mapa(unsigned addr, unsigned rank) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}

// mode 4: st.async stores, time until the receiver's mbarrier completes (the trailing cluster barrier not counted);
// mode 0: barrier only; 1: v2.f64 remote stores; 2: scalar f64 remote stores (pairs 32 B apart, the column-side pattern);
// 3: bulk copies smem -> peer smem.  `bytes` leave every CTA per iteration, spread over `npeers` peers.
__global__ void __launch_bounds__(384, 1) k_x(int mode, int cs, unsigned bytes, int npeers, int issuers, int iters,
                                               long long* __restrict__ out) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem);   // [2]
    double* src = reinterpret_cast<double*>(smem + 1024);
    double* dst = src + 8192;   // 64 KB further on
    const unsigned me = cluster_rank();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sm(bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int e = tid; e < 8192; e += blockDim.x) src[e] = e;
    __syncthreads();
    cluster_sync_all();
    const unsigned per_peer = bytes / npeers;
    long long t0 = clock64();
    long long t_data = 0;
    for (int it = 0; it < iters; ++it) {
        const long long t_it = clock64();
        if (mode == 1 && tid < 256) {
            for (int p = 0; p < npeers; ++p) {
                const unsigned r = mapa(sm(dst), (me + 1 + p) % cs);
                for (unsigned e = tid; e < per_peer / 16; e += 256)
                    asm volatile("st.shared::cluster.v2.f64 [%0], {%1, %2};" ::"r"(r + e * 16), "d"(1.0 + it), "d"(2.0) : "memory");
            }
        } else if (mode == 2 && tid < 256) {
            for (int p = 0; p < npeers; ++p) {
                const unsigned r = mapa(sm(dst), (me + 1 + p) % cs);
                for (unsigned e = tid; e < per_peer / 16; e += 256) {
                    // lane (g, t): two doubles 32 B apart, lanes of a quad 8 B... the transposed-store pattern of the epilogue
                    const unsigned o = (e >> 5) * 512 + (e & 3) * 64 + ((e >> 2) & 7) * 8;
                    asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(r + o), "d"(1.0 + it) : "memory");
                    asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(r + o + 32), "d"(2.0) : "memory");
                }
            }
        } else if (mode == 4) {
            // st.async (STAS) with byte-counted completion on the receiver's mbarrier: no release / acquire round trip
            if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sm(bar)), "r"(bytes) : "memory");
            if (tid < 256) {
                for (int p = 0; p < npeers; ++p) {
                    const unsigned r = mapa(sm(dst), (me + 1 + p) % cs), rb = mapa(sm(bar), (me + 1 + p) % cs);
                    for (unsigned e = tid; e < per_peer / 16; e += 256)
                        asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];"
                                     ::"r"(r + e * 16), "d"(1.0 + it), "d"(2.0), "r"(rb) : "memory");
                }
            }
            mbar_wait(sm(bar), it & 1);
            t_data += clock64() - t_it;
        } else if (mode == 3) {
            // the receiver expects `bytes` in total on its own barrier; `issuers` warps issue the copies, round robin
            if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sm(bar)), "r"(bytes) : "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            const unsigned piece = 6144;   // 12 blocks x 512 B: one k4-step of the twelve state blocks
            const unsigned npieces = per_peer / piece;
            if (lane == 0 && warp < issuers) {
                int n = 0;
                for (int p = 0; p < npeers; ++p) {
                    const unsigned peer = (me + 1 + p) % cs;
                    const unsigned rdst = mapa(sm(dst), peer), rbar = mapa(sm(bar), peer);
                    for (unsigned q = 0; q < npieces; ++q, ++n)
                        if (n % issuers == warp)
                            asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                         ::"r"(rdst + q * piece), "r"(sm(src) + q * piece), "r"(piece), "r"(rbar) : "memory");
                }
            }
            mbar_wait(sm(bar), it & 1);
        }
        cluster_sync_all();
    }
    long long t1 = clock64();
    if (tid == 0) out[blockIdx.x] = mode == 4 ? t_data / iters : (t1 - t0) / iters;
    if (dst[5] == -1.0) out[0] = 0;   // keep the stores alive
}

int main() {
    int dev = 0;
    cudaSetDevice(dev);
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, dev);
    printf("%s, %d SMs\n", prop.name, prop.multiProcessorCount);
    long long* out;
    cudaMalloc(&out, 4096 * sizeof(long long));
    const size_t smem_bytes = 200 * 1024;
    cudaFuncSetAttribute(k_x, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    cudaFuncSetAttribute(k_x, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cs : {4, 8, 16}) {
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cs;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        cfg.blockDim = dim3(384);
        cfg.dynamicSmemBytes = smem_bytes;
        cfg.gridDim = dim3(cs);
        int nclusters = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&nclusters, k_x, &cfg);
        printf("cluster size %2d, %zu KB smem, 384 threads: max active clusters %d (%s)\n", cs, smem_bytes / 1024, nclusters,
               cudaGetErrorString(e));
        if (nclusters <= 0) continue;
        for (int nclu : {1, nclusters}) {
            cfg.gridDim = dim3(cs * nclu);
            struct Case { int mode; unsigned bytes; int npeers, issuers; const char* what; };
            const Case cases[] = {
                {0, 0, 1, 1, "barrier only"},
                {1, 24576, 1, 1, "st v2.f64 24 KB -> 1 peer"},
                {1, 49152, 1, 1, "st v2.f64 48 KB -> 1 peer"},
                {1, 49152, 3, 1, "st v2.f64 48 KB -> 3 peers"},
                {2, 24576, 1, 1, "st f64 pairs 24 KB -> 1 peer"},
                {2, 49152, 3, 1, "st f64 pairs 48 KB -> 3 peers"},
                {4, 24576, 1, 1, "st.async v2 24 KB -> 1 peer (to data)"},
                {4, 49152, 1, 1, "st.async v2 48 KB -> 1 peer (to data)"},
                {4, 49152, 3, 1, "st.async v2 48 KB -> 3 peers (to data)"},
                {3, 24576, 1, 1, "bulk 24 KB -> 1 peer, 1 issuer"},
                {3, 49152, 1, 1, "bulk 48 KB -> 1 peer, 1 issuer"},
                {3, 49152, 1, 4, "bulk 48 KB -> 1 peer, 4 issuers"},
                {3, 36864, 3, 1, "bulk 36 KB -> 3 peers, 1 issuer"},
                {3, 36864, 3, 3, "bulk 36 KB -> 3 peers, 3 issuers"},
                {3, 36864, 3, 6, "bulk 36 KB -> 3 peers, 6 issuers"},
            };
            for (const Case& c : cases) {
                if (c.npeers >= cs) continue;
                const int iters = 200;
                int mode = c.mode, csz = cs, np = c.npeers, is = c.issuers;
                unsigned by = c.bytes;
                int it = iters;
                void* args[] = {&mode, &csz, &by, &np, &is, &it, &out};
                cudaError_t le = cudaLaunchKernelExC(&cfg, (const void*)k_x, args);
                cudaError_t se = cudaDeviceSynchronize();
                if (le != cudaSuccess || se != cudaSuccess) {
                    printf("  %-36s launch %s / %s\n", c.what, cudaGetErrorString(le), cudaGetErrorString(se));
                    return 1;
                }
                std::vector<long long> h(cfg.gridDim.x);
                cudaMemcpy(h.data(), out, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
                const long long mx = *std::max_element(h.begin(), h.end()), mn = *std::min_element(h.begin(), h.end());
                printf("  clusters %3d  %-36s %6lld .. %6lld cycles / iteration", nclu, c.what, mn, mx);
                if (c.bytes) printf("   (%.1f B/clk/CTA)", (double)c.bytes / (double)mx);
                printf("\n");
            }
        }
    }
    return 0;
}
