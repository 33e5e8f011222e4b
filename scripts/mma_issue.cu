// Microbenchmark: what the MMA-issuing warp can sustain.  One CTA per SM, tcgen05.mma M=128 N=256 K=16
// (128 tensor-pipe cycles each) issued by one elected lane of a converged warp, in groups of G MMAs with
// per-group overhead patterns:
//   0: nothing                      1: one (already complete) mbarrier try_wait per group
//   2: try_wait + tcgen05.commit    3: ld.shared poll + commit          4: commit only
// Prints tensor-pipe cycles per MMA (128 = pipe-bound).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I nerf_b200/csrc -o scripts/bin/mma_issue scripts/mma_issue.cu
#include "tc_common.cuh"
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(64, 1) issue_kernel(int pattern, int group, int n_groups, long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw), base = (raw + 1023u) & ~1023u;
    // A: 16 KB slab at +0, B: 2 x 16 KB at +16 KB, barriers at +48 KB
    const uint32_t bar_done = base + 49152, bar_ready = base + 49160, bar_sink = base + 49168, flag = base + 49184;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem_raw + (base - raw) + 49200);
    for (int i = threadIdx.x; i < 49152 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - raw))[i] = 0;
    if (threadIdx.x == 0) {
        tc::mbar_init(bar_done, 1); tc::mbar_init(bar_ready, 1); tc::mbar_init(bar_sink, 1);
        *reinterpret_cast<volatile uint32_t*>(smem_raw + (base - raw) + 49184) = 1;
        tc::fence_barrier_init();
    }
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tc::tmem_alloc(tc::smem_u32((const void*)tmem_slot), 512);
    tc::fence_proxy_async_smem();
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (threadIdx.x == 32) tc::mbar_arrive(bar_ready);              // phase 0 of bar_ready is complete from here on
    __syncthreads();
    if (warp == 0) {
        const uint32_t idesc = tc::umma_idesc_bf16(128, 256);
        long long t0 = clock64();
        for (int g = 0; g < n_groups; ++g) {
            if (pattern == 1 || pattern == 2) tc::mbar_wait(bar_ready, 0, 0);
            if (pattern == 3) {
                uint32_t v;
                do { asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(flag)); } while (v == 0);
            }
            if (tc::elect_one()) {
                for (int k = 0; k < group; ++k)
                    tc::umma_bf16(tmem + (g & 1) * 256, tc::umma_desc_sw128(base + (k & 3) * 32), tc::umma_desc_sw128(base + 16384 + (k & 3) * 32), idesc, 1);
                if (pattern >= 2) tc::umma_commit(bar_sink);
            }
            __syncwarp();
        }
        if (tc::elect_one()) tc::umma_commit(bar_done);
        __syncwarp();
        tc::mbar_wait(bar_done, 0, 0);
        if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) { tc::tcgen05_fence_after(); tc::tmem_dealloc(tmem, 512); }
}

int main() {
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    long long* cyc; CK(cudaMalloc(&cyc, 256 * sizeof(long long)));
    CK(cudaFuncSetAttribute(issue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    const char* names[] = {"plain", "try_wait", "try_wait+commit", "lds poll+commit", "commit"};
    for (int pattern = 0; pattern < 5; ++pattern)
        for (int group = 1; group <= 8; group *= 2) {
            const int n_groups = 4096 / group;
            for (int rep = 0; rep < 2; ++rep) { issue_kernel<<<sms, 64, 64 * 1024>>>(pattern, group, n_groups, cyc); CK(cudaDeviceSynchronize()); }
            long long h[256]; CK(cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
            double mx = 0; for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
            printf("%-16s group %d: %.1f cycles/MMA (%.0f per group)\n", names[pattern], group, mx / 4096.0, mx / n_groups);
        }
    return 0;
}
