// Microbenchmark: what a persistent one-CTA-per-SM kernel can pull through HBM with bulk copies when every byte is
// read once and as many bytes are written (the data-gradient kernel's traffic: load a 64 KB slab image, store a 64 KB
// slab image), by buffers per issuing thread (1 = the buffer is refilled only after its store has been read out, like
// the in-place kernel; 2 = the next load is in flight during the store) and by issuing threads per CTA.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hbm_stream scripts/hbm_stream.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_store(void* dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_prefetch(const void* src, uint32_t bytes) { asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory"); }

// chunk = 64 KB moved as 4 x 16 KB requests.  Thread w (one lane of warp w < n_thr) owns chunks w, w + n_thr*grid ... of its CTA.
__global__ void __launch_bounds__(128, 1) stream_kernel(const uint8_t* src, uint8_t* dst, int n_chunks, int n_thr, int depth, int prefetch, int chunk_kb, int scatter) {
    extern __shared__ __align__(1024) uint8_t sm[];
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) != 0 || w >= n_thr) return;
    const uint32_t CH = chunk_kb * 1024u, RQ = CH / 4;
    const uint32_t base = ((smem_u32(sm) + 1023u) & ~1023u) + w * depth * CH;
    const uint32_t bars = ((smem_u32(sm) + 1023u) & ~1023u) + n_thr * depth * CH + w * 8 * depth;
    for (int i = 0; i < depth; ++i) mbar_init(bars + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const int stride = gridDim.x * n_thr, first = blockIdx.x * n_thr + w;
    int n_my = 0; for (int c = first; c < n_chunks; c += stride) ++n_my;
    auto issue = [&](int i) {
        int c = first + i * stride; const int b = i % depth;
        if (scatter) c = (int)(((long long)c * 7919) % n_chunks);
        mbar_expect(bars + 8 * b, CH);
        for (int r = 0; r < 4; ++r) tma_load(base + b * CH + r * RQ, src + (size_t)c * CH + r * RQ, RQ, bars + 8 * b);
        if (prefetch && i + depth < n_my) { int cn = first + (i + depth) * stride; if (scatter) cn = (int)(((long long)cn * 7919) % n_chunks); tma_prefetch(src + (size_t)cn * CH, CH); }
    };
    for (int i = 0; i < depth - 1 && i < n_my; ++i) issue(i);
    for (int i = 0; i < n_my; ++i) {
        const int b = i % depth;
        if (depth == 1) issue(i);
        else if (i + depth - 1 < n_my) {
            // the buffer the next load lands in was stored at iteration i-1: its store must have been read out
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            issue(i + depth - 1);
        }
        mbar_wait(bars + 8 * b, (i / depth) & 1);
        int c = first + i * stride;
        if (scatter) c = (int)(((long long)c * 7919) % n_chunks);
        for (int r = 0; r < 4; ++r) tma_store(dst + (size_t)c * CH + r * RQ, base + b * CH + r * RQ, RQ);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (depth == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main() {
    const size_t bytes = 1359ull << 20;                 // 1.36 GB read + 1.36 GB written, like the data-gradient kernel
    uint8_t *src, *dst, *flush;
    CK(cudaMalloc(&src, bytes)); CK(cudaMalloc(&dst, bytes)); CK(cudaMalloc(&flush, 256u << 20));
    CK(cudaMemset(src, 1, bytes));
    CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    { // cudaMemcpy D2D for reference
        float best = 1e9;
        for (int it = 0; it < 5; ++it) { CK(cudaMemset(flush, it, 256u << 20)); cudaEventRecord(e0); CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice)); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
        printf("cudaMemcpy D2D                                  : %.3f ms  %.0f GB/s\n", best, 2.0 * bytes / best / 1e6);
    }
    const int cfg[][4] = {{2, 1, 0, 64}, {2, 1, 1, 64}, {1, 3, 0, 64}, {4, 2, 0, 16}, {2, 3, 0, 32}};
    for (int scatter = 0; scatter < 2; ++scatter)
    for (auto& c : cfg) {
        const int n_thr = c[0], depth = c[1], pf = c[2], kb = c[3];
        const int n_chunks = (int)(bytes / (kb * 1024));
        const size_t smem = (size_t)n_thr * depth * kb * 1024 + 2048;
        if (smem > 227 * 1024) continue;
        float best = 1e9;
        for (int it = 0; it < 5; ++it) {
            CK(cudaMemset(flush, it, 256u << 20));
            cudaEventRecord(e0);
            stream_kernel<<<148, 128, smem>>>(src, dst, n_chunks, n_thr, depth, pf, kb, scatter);
            cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("%s threads/CTA %d  buffers/thread %d  L2 prefetch %d  chunk %2d KB : %.3f ms  %.0f GB/s\n", scatter ? "scattered " : "sequential", n_thr, depth, pf, kb, best, 2.0 * bytes / best / 1e6);
    }
    return 0;
}
