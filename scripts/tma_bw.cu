// Microbenchmark: per-SM throughput of TMA loads out of L2 into a shared-memory ring, 1-D bulk copies
// (cp.async.bulk) against 2-D tensor-map copies (cp.async.bulk.tensor), by request size and ring depth.
// One CTA per SM, one warp; every CTA streams the same 1.25 MB image (the forward kernel's access
// pattern).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_bw scripts/tma_bw.cu
#include <cuda.h>
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

// MODE 0: one lane of each of `blockDim/32` warps issues; MODE 1: the first `n_lanes` lanes of warp 0 issue
template <int TENSOR>
__global__ void __launch_bounds__(256, 1) lanes_kernel(const uint8_t* src, int img_bytes, int req_bytes, int depth, int n_req, int n_lanes, int nowait,
                                                       long long* cycles) {
    extern __shared__ __align__(1024) uint8_t sm[];
    const int lane = threadIdx.x;
    const uint32_t base = ((smem_u32(sm) + 1023u) & ~1023u) + lane * depth * req_bytes;
    const uint32_t bars = ((smem_u32(sm) + 1023u) & ~1023u) + n_lanes * depth * req_bytes + lane * 8 * depth;
    if (lane < n_lanes) {
        for (int i = 0; i < depth; ++i) mbar_init(bars + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (lane < n_lanes) {
        const int per_img = img_bytes / req_bytes;
        long long t0 = clock64();
        for (int i = 0; i < n_req + depth; ++i) {
            const int slot = i % depth;
            if (i >= depth && !(nowait && i < n_req)) mbar_wait(bars + 8 * slot, ((i / depth) - 1) & 1);
            if (i < n_req) {
                if (!nowait || i < depth) mbar_expect(bars + 8 * slot, nowait ? req_bytes * (n_req / depth) : req_bytes);
                const int r = (i * n_lanes + lane) % per_img;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(base + slot * req_bytes), "l"(src + (size_t)r * req_bytes), "r"(req_bytes), "r"(bars + 8 * slot) : "memory");
            }
        }
        if (lane == 0) cycles[blockIdx.x] = clock64() - t0;
    }
}

template <int TENSOR>
__global__ void __launch_bounds__(256, 1) stream_kernel(const uint8_t* src, const __grid_constant__ CUtensorMap tmap, int img_bytes, int req_bytes,
                                                        int depth, int n_req, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t sm[];
    const int warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const uint32_t base = ((smem_u32(sm) + 1023u) & ~1023u) + warp * depth * req_bytes;   // each warp streams through its own ring
    const uint32_t bars = ((smem_u32(sm) + 1023u) & ~1023u) + n_warps * depth * req_bytes + warp * 8 * depth;
    if ((threadIdx.x & 31) == 0) {
        for (int i = 0; i < depth; ++i) mbar_init(bars + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        const int per_img = img_bytes / req_bytes;
        long long t0 = clock64();
        for (int i = 0; i < n_req + depth; ++i) {
            const int slot = i % depth;
            const bool pf = blockIdx.x == 0 && warp == 0 && i >= 64 && i < 72;
            long long c0 = clock64();
            if (i >= depth) mbar_wait(bars + 8 * slot, ((i / depth) - 1) & 1);
            long long c1 = clock64();
            if (i < n_req) {
                mbar_expect(bars + 8 * slot, req_bytes);
                long long c2 = clock64();
                if (pf) { cycles[160 + (i - 64) * 4] = c0 - t0; cycles[161 + (i - 64) * 4] = c1 - c0; cycles[162 + (i - 64) * 4] = c2 - c1; }
                const int r = (i * n_warps + warp) % per_img;
                if (TENSOR) {
                    const int row = r * (req_bytes / 1024);
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                                 ::"r"(base + slot * req_bytes), "l"(&tmap), "r"(0), "r"(row), "r"(bars + 8 * slot) : "memory");
                } else {
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(base + slot * req_bytes), "l"(src + (size_t)r * req_bytes), "r"(req_bytes), "r"(bars + 8 * slot) : "memory");
                }
                if (pf) cycles[163 + (i - 64) * 4] = clock64() - c2;
            }
        }
        if (warp == 0) cycles[blockIdx.x] = clock64() - t0;
    }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const int img = 1280 * 1024;
    uint8_t* d; CK(cudaMalloc(&d, img)); CK(cudaMemset(d, 1, img));
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    long long* cyc; CK(cudaMalloc(&cyc, 256 * sizeof(long long)));
    EncodeFn encode = nullptr; cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qr));
    CK(cudaFuncSetAttribute(stream_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(stream_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    const int reqs[] = {4096, 8192, 16384, 32768};
    for (int tensor = 0; tensor < 1; ++tensor)
      for (int warps = 1; warps <= 1; warps *= 2)
        for (int ri = 1; ri < 3; ++ri)
            for (int depth = 4; depth <= 4; depth *= 2) {
                const int req = reqs[ri];
                if (warps * depth * req > 128 * 1024) continue;
                CUtensorMap tm; memset(&tm, 0, sizeof(tm));
                cuuint64_t dims[2] = {256, (cuuint64_t)img / 1024}; cuuint64_t strides[1] = {1024};
                cuuint32_t box[2] = {256, (cuuint32_t)req / 1024}; cuuint32_t es[2] = {1, 1};
                CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
                const int n_req = (64 << 20) / req / 8 / warps;         // 8 MB per SM
                const size_t smem = warps * depth * req + 8 * depth * warps + 2048;
                for (int rep = 0; rep < 2; ++rep) {
                    if (tensor) stream_kernel<1><<<sms, 32 * warps, smem>>>(d, tm, img, req, depth, n_req, cyc);
                    else stream_kernel<0><<<sms, 32 * warps, smem>>>(d, tm, img, req, depth, n_req, cyc);
                    CK(cudaDeviceSynchronize());
                }
                long long h[256]; CK(cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
                double mx = 0; for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
                printf("%s warps %d req %5d B depth %d: %.1f B/cycle/SM, %.0f cycles/request/warp\n", tensor ? "tensor2d" : "bulk1d  ", warps, req, depth,
                       (double)n_req * warps * req / mx, mx / n_req);
                { long long g[32]; CK(cudaMemcpy(g, cyc + 160, sizeof(g), cudaMemcpyDeviceToHost));
                  for (int i = 0; i < 8; ++i) printf("   iter %d @%lld: wait %lld expect %lld issue %lld\n", 64 + i, g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]); }
            }
    return 0;
}
