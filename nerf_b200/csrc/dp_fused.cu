// Data-parallel exchange step of the training loop as ONE kernel over NVLink / NVSwitch peer memory:
//     cross-GPU barrier  ->  all-reduce of the flat gradient  ->  torch.optim.Adam on every rank
// (reference: the loop body train_conditional_nerf.py:130-135 run data-parallel, SURVEY 8e: "one all-reduce of the flat
// gradient buffer per step, then identical fused Adam on every rank").  The gradient lives in SYMMETRIC memory (same
// allocation on every rank, peer-mapped; multicast-bound when the fabric offers NVLS):
//   * NVLS path: every rank reads the SUM with multimem.ld_reduce.add.v4.f32 -- the switch adds the ranks' copies, each
//     GPU pulls 2.55 MB once, nothing is written back;
//   * peer path (no multicast object): one-shot all-reduce with plain P2P loads from every rank's buffer.
// Both orders of summation are the same on every rank, so the parameters stay bit-identical across ranks.
// Barrier: a monotone epoch per rank in a symmetric flag block.  Rank r's block 0 publishes `epoch` into slot r of EVERY
// rank's flags (st.release.sys after the kernel boundary that completed its gradient kernel); every block of every rank
// spins (bounded) on its LOCAL flags until all slots reach `epoch`.  The caller alternates between two gradient buffers by
// epoch parity, so a rank that runs ahead into the next step never overwrites a buffer a slower rank is still reading.
#include "common.cuh"
#include <math.h>

#define DP_MAX_WORLD 16

__device__ __forceinline__ void st_release_sys_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_add_f32x4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_f32x4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

struct DpArgs {
    const float* grads_mc;                 // multicast address of this epoch's gradient buffer (NVLS) or NULL
    const float* const* grads_peers;       // device array [world]: every rank's mapping of the gradient allocation
    uint64_t* const* flags_peers;          // device array [world]: every rank's mapping of the flag block ([world] uint64)
    int64_t grad_offset;                   // element offset of this epoch's buffer inside the allocation
    int64_t flag_offset;                   // element (uint64) offset of the flag block inside its allocation
    int rank, world;
    uint64_t epoch;
    float* p; float* m; float* v;
    int64_t n4;                            // float4 count (the buffers are padded to a multiple of 4)
    float lr, b1, b2, eps, bc1, bc2_sqrt, gscale;
};

__device__ __forceinline__ float adam1(float g, float& m, float& v, float p, const DpArgs& a) {
    g *= a.gscale;
    m = m + (g - m) * (1.0f - a.b1);                        // the arithmetic of adam_kernel (api.cu) / torch.optim.Adam
    v = a.b2 * v + (1.0f - a.b2) * g * g;
    return p - (a.lr / a.bc1) * (m / (sqrtf(v) / a.bc2_sqrt + a.eps));
}

__global__ void __launch_bounds__(256) dp_allreduce_adam_kernel(const DpArgs a) {
    // ---- barrier: publish, then wait for every rank (bounded: a lost peer traps instead of hanging the GPU)
    if (blockIdx.x == 0 && threadIdx.x < a.world) st_release_sys_u64(a.flags_peers[threadIdx.x] + a.flag_offset + a.rank, a.epoch);
    if (threadIdx.x < a.world) {
        const uint64_t* mine = a.flags_peers[a.rank] + a.flag_offset + threadIdx.x;
        unsigned long long t0 = 0;
        for (uint32_t it = 0; ld_acquire_sys_u64(mine) < a.epoch; ++it) {
            if ((it & 1023u) == 1023u) {
                unsigned long long now; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (!t0) t0 = now; else if (now - t0 > 20000000000ull) __trap();     // 20 s
            }
        }
    }
    __syncthreads();
    // ---- all-reduce + Adam, 16 bytes per thread and iteration
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 g;
        if (a.grads_mc) g = multimem_ld_reduce_add_f32x4(a.grads_mc + a.grad_offset + 4 * i);
        else {
            g = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int r = 0; r < a.world; ++r) {            // rank order: the same sum on every rank
                const float4 x = ld_relaxed_sys_f32x4(a.grads_peers[r] + a.grad_offset + 4 * i);
                g.x += x.x; g.y += x.y; g.z += x.z; g.w += x.w;
            }
        }
        float4 m = reinterpret_cast<float4*>(a.m)[i], v = reinterpret_cast<float4*>(a.v)[i], p = reinterpret_cast<float4*>(a.p)[i];
        p.x = adam1(g.x, m.x, v.x, p.x, a); p.y = adam1(g.y, m.y, v.y, p.y, a);
        p.z = adam1(g.z, m.z, v.z, p.z, a); p.w = adam1(g.w, m.w, v.w, p.w, a);
        reinterpret_cast<float4*>(a.m)[i] = m; reinterpret_cast<float4*>(a.v)[i] = v; reinterpret_cast<float4*>(a.p)[i] = p;
    }
}

extern "C" int nerf_dp_allreduce_adam(const float* grads_mc, const float* const* grads_peers, uint64_t* const* flags_peers,
                                      int64_t grad_offset, int64_t flag_offset, int rank, int world, uint64_t epoch,
                                      float* params, float* exp_avg, float* exp_avg_sq, int64_t n, int64_t step,
                                      float lr, float beta1, float beta2, float eps, float grad_scale, void* stream) {
    if (!grads_peers || !flags_peers || !params || !exp_avg || !exp_avg_sq || n < 0 || (n & 3) || (grad_offset & 3) || flag_offset < 0 || step < 1 ||
        world < 1 || world > DP_MAX_WORLD || rank < 0 || rank >= world || epoch == 0) return NERF_ERR_BAD_ARG;
    DpArgs a{};
    a.grads_mc = grads_mc; a.grads_peers = grads_peers; a.flags_peers = flags_peers; a.grad_offset = grad_offset; a.flag_offset = flag_offset;
    a.rank = rank; a.world = world; a.epoch = epoch; a.p = params; a.m = exp_avg; a.v = exp_avg_sq; a.n4 = n / 4;
    a.lr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps; a.gscale = grad_scale;
    a.bc1 = (float)(1.0 - pow((double)beta1, (double)step)); a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    int64_t blocks = (a.n4 + 255) / 256;
    const int64_t cap = (int64_t)nerf_num_sms() * 4; if (blocks > cap) blocks = cap; if (blocks < 1) blocks = 1;
    dp_allreduce_adam_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
