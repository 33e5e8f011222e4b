// Shared helpers for the nerf_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/nerf_b200.h"

#define NERF_EPS_T 1e-10f      // nerf/model.py:469  (alpha + 1e-10 inside the cumprod)
#define NERF_LAST_DIST 1e10f   // nerf/model.py:464  (last distance)
#define NERF_LN_EPS 1e-5f      // torch.nn.LayerNorm default

int nerf_set_error(cudaError_t e);   // records + returns (int)e
int nerf_set_error_text(int status, const char* text);   // records the text, returns status
#define NERF_CUDA_CHECK_LAUNCH() do { cudaError_t e__ = cudaGetLastError(); \
    if (e__ != cudaSuccess) return nerf_set_error(e__); } while (0)
#define NERF_CUDA(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) return nerf_set_error(e__); } while (0)

// Per-DEVICE caches (a process may touch more than one GPU): SM count and the opt-in shared-memory attribute.
#define NERF_MAX_DEVICES 64
static inline int nerf_current_device() { int dev = 0; cudaGetDevice(&dev); return dev; }
static inline int nerf_num_sms() {
    static int n[NERF_MAX_DEVICES] = {};
    const int dev = nerf_current_device();
    int v = (dev >= 0 && dev < NERF_MAX_DEVICES) ? n[dev] : 0;
    if (!v) {
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev); if (v <= 0) v = 148;
        if (dev >= 0 && dev < NERF_MAX_DEVICES) n[dev] = v;
    }
    return v;
}
// cudaFuncAttributeMaxDynamicSharedMemorySize once per (kernel, device)
template <auto Kernel>
static inline cudaError_t nerf_smem_opt_in(int bytes) {
    static bool done[NERF_MAX_DEVICES] = {};
    const int dev = nerf_current_device();
    if (dev >= 0 && dev < NERF_MAX_DEVICES && done[dev]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && dev >= 0 && dev < NERF_MAX_DEVICES) done[dev] = true;
    return e;
}
// tuning knobs from the environment, read once per process (not per launch)
#include <stdlib.h>
static inline uint32_t nerf_env_u32(const char* name, uint32_t dflt) { const char* e = getenv(name); return e ? (uint32_t)strtoul(e, nullptr, 0) : dflt; }
static inline bool nerf_env_set(const char* name) { return getenv(name) != nullptr; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- Philox4x32-7 (counter based; keyed by (seed), counter = (ray, sample)) ----
// Seven rounds: the Crush-resistant variant of Salmon et al. (Random123, SC'11; ten rounds are their safety-margin
// default).  Every round is two 32 x 32 -> 64 multiplies on the fma pipe, which is what bound the in-kernel-jitter sampler
// at 0.50 of the HBM peak with ten; the streams are checked as distributions in tests/test_gpu_train_parity.py.
#define NERF_PHILOX_ROUNDS 7
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
    uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
    uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}
__device__ __forceinline__ void philox4x32(uint64_t seed, uint64_t ctr_lo, uint32_t ctr_hi, uint32_t (&out)[4]) {
    uint32_t c[4] = {(uint32_t)ctr_lo, (uint32_t)(ctr_lo >> 32), ctr_hi, 0u};
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int i = 0; i < NERF_PHILOX_ROUNDS; ++i) { philox_round(c, k0, k1); k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
    out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
}
// uniform in [0,1) with 24 random bits (same grid as torch.rand for fp32)
__device__ __forceinline__ float u01_from_bits(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// torch.linspace(start, end, steps) element i in fp32 (ATen RangeFactories: symmetric halves)
__device__ __forceinline__ float linspace_at(float start, float end, int steps, int i) {
    if (steps == 1) return start;
    float step = (end - start) / (float)(steps - 1);
    return (i < steps / 2) ? __fadd_rn(start, __fmul_rn(step, (float)i))
                           : __fsub_rn(end, __fmul_rn(step, (float)(steps - 1 - i)));
}

// ---------------------------------------------------------------- camera ray --
// x = (col - 0.5 (W-1)) / f ; y = (row - 0.5 (H-1)) / f ; ray = (x, -y, -1)
// (nerf/model.py:276-278), each op rounded separately like eager torch.
__device__ __forceinline__ void camera_ray(int row, int col, int H, int W, float focal, float (&r)[3]) {
    float cx = 0.5f * (float)(W - 1), cy = 0.5f * (float)(H - 1);   // python double -> fp32 scalar
    r[0] = __fdiv_rn(__fsub_rn((float)col, cx), focal);
    r[1] = -__fdiv_rn(__fsub_rn((float)row, cy), focal);
    r[2] = -1.0f;
}
// d_i = sum_j R[i][j] ray[j]  ((camera_r * rays[..., None, :]).sum(-1), model.py:367)
__device__ __forceinline__ void rotate_ray(const float* __restrict__ R, const float (&r)[3], float (&d)[3]) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float p0 = __fmul_rn(R[i * 3 + 0], r[0]), p1 = __fmul_rn(R[i * 3 + 1], r[1]), p2 = __fmul_rn(R[i * 3 + 2], r[2]);
        d[i] = __fadd_rn(__fadd_rn(p0, p1), p2);
    }
}

