// Library bookkeeping, the flat parameter layout, fused Adam and the MSE loss.
#include "common.cuh"
#include <string.h>
#include <stdio.h>
#include <math.h>

static thread_local char g_err[256] = "no error";

int nerf_set_error(cudaError_t e) {
    const char* n = cudaGetErrorName(e); const char* s = cudaGetErrorString(e);
    snprintf(g_err, sizeof(g_err), "%s: %s", n ? n : "?", s ? s : "?");
    return (int)e;
}
// argument errors that deserve more than a status code: the text is what nerf_last_error() returns next
int nerf_set_error_text(int status, const char* text) { snprintf(g_err, sizeof(g_err), "%s", text); return status; }
extern "C" int nerf_abi_version(void) { return 1; }
extern "C" const char* nerf_last_error(void) { return g_err; }

// state_dict order of examples/nerf.pth: density.{w,b}, color.{w,b},
// block_0.{0,2,3,5,6,8,9,11}.{w,b}, block_1.*, block_2.{0,2,3,5}.{w,b}
extern "C" int64_t nerf_g1_param_layout(int Dx, int Dd, int64_t* off) {
    if (Dx < 3 || Dd < 3) return NERF_ERR_BAD_ARG;
    const int64_t H = NERF_HIDDEN, EX = 2 * NERF_PE_X_OCTAVES * Dx, ED = 2 * NERF_PE_D_OCTAVES * Dd;
    int64_t p = 0; int n = 0;
    auto put = [&](int64_t sz) { if (off) off[n] = p; ++n; p += sz; };
    put(H); put(1); put(3 * H); put(3);
    const int64_t in0[3] = {EX, H + EX, H + ED};
    const int units[3] = {4, 4, 2};
    for (int b = 0; b < 3; ++b)
        for (int u = 0; u < units[b]; ++u) {
            put(H * (u == 0 ? in0[b] : H)); put(H);   // Linear weight [H,in], bias
            put(H); put(H);                           // LayerNorm gamma, beta
        }
    return p;
}

// ------------------------------------------------------------------ Adam ------
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t n, float lr, float b1, float b2, float eps,
                            float bc1, float bc2_sqrt, float gscale) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // torch.optim.Adam (single tensor path): m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2
    // denom = sqrt(v)/sqrt(bc2) + eps ; p -= (lr/bc1) m / denom
    float gi = g[i] * gscale;
    float mi = m[i] + (gi - m[i]) * (1.0f - b1);          // lerp form used by torch
    float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - (lr / bc1) * (mi / denom);
}

extern "C" int nerf_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq,
                              int64_t n, int64_t step, float lr, float beta1, float beta2, float eps,
                              float grad_scale, void* stream) {
    if (!params || !grads || !exp_avg || !exp_avg_sq || n < 0 || step < 1) return NERF_ERR_BAD_ARG;
    if (n == 0) return NERF_OK;
    double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, (float)bc1, (float)sqrt(bc2), grad_scale);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// ------------------------------------------------------------------ loss ------
__global__ void mse_kernel(const float* __restrict__ x, const float* __restrict__ y, int64_t n,
                           float* __restrict__ loss, float* __restrict__ g) {
    __shared__ float part[8];
    float acc = 0.f;
    const float inv = 1.0f / (float)n;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float d = x[i] - y[i];
        acc += d * d;
        if (g) g[i] = 2.0f * d * inv;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : 0.f;
        v = warp_sum(v);
        if (threadIdx.x == 0) atomicAdd(loss, v * inv);
    }
}

extern "C" int nerf_mse_loss(const float* pixels, const float* target, int64_t n, float* loss_out,
                             float* g_pixels, void* stream) {
    if (!pixels || !target || !loss_out || n <= 0) return NERF_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    NERF_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float), st));
    int64_t blocks = (n + 255) / 256; if (blocks > 1024) blocks = 1024;
    mse_kernel<<<(unsigned)blocks, 256, 0, st>>>(pixels, target, n, loss_out, g_pixels);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
