// Micro-benchmark: per-SM and chip-wide L2 -> shared-memory throughput of TMA bulk copies (cp.async.bulk, UBLKCP) as a
// function of copy size, copies in flight and CTAs running; and the issue cost of one copy.  Round-2 design input for
// k_stream_tma (how much operand traffic one SM can pull while it runs DMMAs).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_bw scripts/tma_bw.cu && ./tma_bw
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ unsigned sm(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// every warp's lane 0 runs its own ring of `depth` slots (issuers = blockDim.x / 32 independent copy streams per SM)
__global__ void __launch_bounds__(128, 1) k_bw(const double* __restrict__ src, size_t src_doubles, unsigned bytes, int depth,
                                               int iters, long long* __restrict__ out) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(smem) + warp * 16;
    unsigned char* ring = smem + 512 + (size_t)warp * depth * bytes;
    if ((threadIdx.x & 31) == 0) {
        for (int s = 0; s < depth; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sm(bars + s)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if ((threadIdx.x & 31) != 0) return;
    const size_t chunk = bytes / 8, nchunks = src_doubles / chunk;
    size_t pos = ((size_t)(blockIdx.x * nwarps + warp) * 7919) % nchunks;
    long long t_issue = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const int slot = it % depth;
        const unsigned bar = sm(bars + slot);
        if (it >= depth) {
            const unsigned parity = ((it / depth) - 1) & 1;
            unsigned ok;
            do {
                asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                             : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
            } while (!ok);
        }
        const long long ti = clock64();
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(sm(ring + (size_t)slot * bytes)), "l"(src + pos * chunk), "r"(bytes), "r"(bar) : "memory");
        t_issue += clock64() - ti;
        pos += 13;
        if (pos >= nchunks) pos -= nchunks;
    }
    for (int it = iters; it < iters + depth; ++it) {   // drain
        const int slot = it % depth;
        const unsigned bar = sm(bars + slot), parity = ((it / depth) - 1) & 1;
        unsigned ok;
        do {
            asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                         : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        } while (!ok);
    }
    const long long t1 = clock64();
    if (warp == 0) {
        out[2 * blockIdx.x] = t1 - t0;
        out[2 * blockIdx.x + 1] = t_issue;
    }
}

int main() {
    const size_t src_bytes = 48u << 20;   // L2-resident (126 MB L2)
    double* src;
    long long* out;
    cudaMalloc(&src, src_bytes);
    cudaMemset(src, 0, src_bytes);
    cudaMalloc(&out, 2 * 148 * sizeof(long long));
    cudaFuncSetAttribute(k_bw, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    printf("grid issuers  bytes  depth  B/clk/SM  GB/s/SM(at max clk)  chip TB/s  issue clk/copy\n");
    const int grids[] = {1, 148};
    const unsigned sizes[] = {2048, 8192, 16384, 32768};
    for (int g : grids)
      for (int nw : {1, 2, 4})
        for (unsigned bytes : sizes)
            for (int depth : {2, 4}) {
                if ((size_t)nw * depth * bytes > 192 * 1024) continue;
                const int iters = (int)((8u << 20) / bytes);   // 8 MB per issuing warp
                for (int rep = 0; rep < 2; ++rep)
                    k_bw<<<g, 32 * nw, 512 + (size_t)nw * depth * bytes>>>(src, src_bytes / 8, bytes, depth, iters, out);
                if (cudaDeviceSynchronize() != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                long long h[2 * 148];
                cudaMemcpy(h, out, 2 * g * sizeof(long long), cudaMemcpyDeviceToHost);
                double worst = 0, issue = 0;
                for (int b = 0; b < g; ++b) { worst = h[2 * b] > worst ? (double)h[2 * b] : worst; issue += (double)h[2 * b + 1] / iters / g; }
                const double bpc = (double)nw * iters * bytes / worst;
                printf("%4d %4d %9u %5d %9.2f %12.1f %14.2f %12.0f\n", g, nw, bytes, depth, bpc, bpc * clk_khz * 1e-6, bpc * g * clk_khz * 1e-9,
                       issue);
            }
    return 0;
}
