// G1 positional encoder (standalone form, used by the fp32 path, tests and the
// bandwidth sweep) and the hierarchical inverse-CDF sampler.
// Layout: nerf/model.py:235-240 (per coordinate [L sin | L cos], flattened);
// frequencies pi*2^j.
#include "common.cuh"

#define PI_F32 3.14159274101257324f   // float(np.pi)

template <int L>
__device__ __forceinline__ void pe_write(float v, float* __restrict__ dst) {
    // dst is 16-byte aligned (2L floats, L in {10, 6} -> 80 B / 48 B groups)
    float tmp[2 * L];
#pragma unroll
    for (int j = 0; j < L; ++j) {
        float f = PI_F32 * (float)(1 << j);         // exact: power-of-two scaling
        float s, c;
        sincosf(__fmul_rn(v, f), &s, &c);
        tmp[j] = s; tmp[L + j] = c;
    }
#pragma unroll
    for (int q = 0; q < (2 * L) / 4; ++q)
        reinterpret_cast<float4*>(dst)[q] = make_float4(tmp[4 * q], tmp[4 * q + 1], tmp[4 * q + 2], tmp[4 * q + 3]);
}

__global__ void encode_x_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                const float* __restrict__ t, const float* __restrict__ states_x, int Sx,
                                int64_t B, int S, float norm, float* __restrict__ ex,
                                float* __restrict__ points_out) {
    const int Dx = 3 + Sx;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;    // (point, coordinate)
    int64_t total = B * (int64_t)S * Dx;
    if (i >= total) return;
    int64_t pt = i / Dx; int c = (int)(i - pt * Dx);
    int64_t ray = pt / S;
    float v;
    if (c < 3) {
        float p = __fadd_rn(rays_o[ray * 3 + c], __fmul_rn(rays_d[ray * 3 + c], t[pt]));
        if (points_out) points_out[pt * 3 + c] = p;
        v = __fdiv_rn(p, norm);
    } else {
        v = states_x[ray * Sx + (c - 3)];
    }
    if (ex) pe_write<NERF_PE_X_OCTAVES>(v, ex + (size_t)pt * (2 * NERF_PE_X_OCTAVES * Dx) + c * 2 * NERF_PE_X_OCTAVES);
}

__global__ void encode_d_kernel(const float* __restrict__ rays_d, const float* __restrict__ states_d, int Sd,
                                int64_t B, int normalize_direction, float* __restrict__ ed) {
    const int Dd = 3 + Sd;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;    // (ray, coordinate)
    if (i >= B * Dd) return;
    int64_t ray = i / Dd; int c = (int)(i - ray * Dd);
    float v;
    if (c < 3) {
        float dx = rays_d[ray * 3], dy = rays_d[ray * 3 + 1], dz = rays_d[ray * 3 + 2];
        v = rays_d[ray * 3 + c];
        if (normalize_direction) {
            float n = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
            v = __fdiv_rn(v, n);
        }
    } else {
        v = states_d[ray * Sd + (c - 3)];
    }
    pe_write<NERF_PE_D_OCTAVES>(v, ed + (size_t)ray * (2 * NERF_PE_D_OCTAVES * Dd) + c * 2 * NERF_PE_D_OCTAVES);
}

extern "C" int nerf_encode_g1(const float* rays_o, const float* rays_d, const float* t,
                              const float* states_x, int Sx, const float* states_d, int Sd,
                              int64_t B, int S, float normalize_position, int normalize_direction,
                              float* ex, float* ed, float* points_out, void* stream) {
    if (B == 0 && S > 0) return NERF_OK;
    if (!rays_o || !rays_d || B < 0 || S <= 0 || Sx < 0 || Sd < 0) return NERF_ERR_BAD_ARG;
    if ((Sx > 0 && !states_x) || (Sd > 0 && !states_d)) return NERF_ERR_BAD_ARG;
    if ((ex || points_out) && !t) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    if (ex || points_out) {
        int64_t total = B * (int64_t)S * (3 + Sx);
        encode_x_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
            rays_o, rays_d, t, states_x, Sx, B, S, normalize_position, ex, points_out);
        NERF_CUDA_CHECK_LAUNCH();
    }
    if (ed) {
        int64_t total = B * (int64_t)(3 + Sd);
        encode_d_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
            rays_d, states_d, Sd, B, normalize_direction, ed);
        NERF_CUDA_CHECK_LAUNCH();
    }
    return NERF_OK;
}

// ------------------------------------------------ hierarchical sampler -------
// One warp per ray.  cdf over the S-2 interior weights (+1e-5) on the S-1 midpoint
// bins (oracle/g1.py:sample_pdf); each lane inverts n_fine/32 uniforms by binary
// search in shared memory, then the S coarse and n_fine fine positions are merged
// by rank (both lists are ascending) so no sort is needed.
#define PDF_MAX_S 256
#define PDF_MAX_F 256
__global__ void __launch_bounds__(128) sample_pdf_merge_kernel(
        const float* __restrict__ t, const float* __restrict__ weights, int S, int n_fine,
        const float* __restrict__ u_fine, int64_t B, float* __restrict__ t_out) {
    __shared__ float s_mid[4][PDF_MAX_S];
    __shared__ float s_cdf[4][PDF_MAX_S];
    __shared__ float s_t[4][PDF_MAX_S];
    __shared__ float s_f[4][PDF_MAX_F];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int64_t ray = (int64_t)blockIdx.x * 4 + w;
    if (ray >= B) return;
    const float* tr = t + ray * S;
    const float* wr = weights + ray * S;
    float* mid = s_mid[w]; float* cdf = s_cdf[w]; float* tc = s_t[w]; float* tf = s_f[w];
    for (int k = lane; k < S; k += 32) tc[k] = tr[k];
    __syncwarp();
    for (int k = lane; k < S - 1; k += 32) mid[k] = 0.5f * (tc[k + 1] + tc[k]);
    // inclusive scan of (w[1..S-2] + 1e-5) in segments of 32
    float carry = 0.f;
    const int nb = S - 2;
    for (int base = 0; base < nb; base += 32) {
        int k = base + lane;
        float v = (k < nb) ? wr[k + 1] + 1e-5f : 0.f;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { float n = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += n; }
        v += carry;
        if (k < nb) cdf[k + 1] = v;
        carry = __shfl_sync(0xffffffffu, v, 31);
    }
    if (lane == 0) cdf[0] = 0.f;
    __syncwarp();
    const float total = carry;
    for (int k = lane; k < nb; k += 32) cdf[k + 1] = cdf[k + 1] / total;   // pdf = w / sum, cumsum (same order of ops up to rounding)
    __syncwarp();
    const int nc = S - 1;                     // cdf / mid length
    for (int j = lane; j < n_fine; j += 32) {
        float u = u_fine ? u_fine[ray * n_fine + j]
                         : (n_fine == 1 ? 0.f : linspace_at(0.f, 1.f, n_fine, j));
        // searchsorted(cdf, u, right=True): first index with cdf[idx] > u
        int lo = 0, hi = nc;
        while (lo < hi) { int m = (lo + hi) >> 1; if (cdf[m] > u) hi = m; else lo = m + 1; }
        int below = max(lo - 1, 0), above = min(lo, nc - 1);
        float cb = cdf[below], ca = cdf[above];
        float denom = ca - cb; if (denom < 1e-5f) denom = 1.f;
        tf[j] = mid[below] + (u - cb) / denom * (mid[above] - mid[below]);
    }
    __syncwarp();
    // merge by rank: coarse k goes to k + #{fine < t_k}; fine j to j + #{coarse <= f_j}
    float* out = t_out + ray * (S + n_fine);
    // (the fine list is ascending only when u is; ranks are counted, so any u works)
    for (int k = lane; k < S; k += 32) {
        float v = tc[k]; int r = 0;
        for (int m = 0; m < n_fine; ++m) r += (tf[m] < v);
        out[k + r] = v;
    }
    for (int j = lane; j < n_fine; j += 32) {
        float v = tf[j]; int lo = 0, hi = S;
        while (lo < hi) { int m = (lo + hi) >> 1; if (tc[m] <= v) lo = m + 1; else hi = m; }
        int r = lo;
        for (int m = 0; m < n_fine; ++m) { float f = tf[m]; r += (f < v) || (f == v && m < j); }
        out[r] = v;
    }
}

extern "C" int nerf_sample_pdf_merge(const float* t, const float* weights, int S, int n_fine,
                                     const float* u_fine, int64_t B, float* t_out, void* stream) {
    if (B == 0 && S >= 3 && n_fine >= 1) return NERF_OK;
    if (!t || !weights || !t_out || S < 3 || S > PDF_MAX_S || n_fine < 1 || n_fine > PDF_MAX_F || B < 0) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    sample_pdf_merge_kernel<<<(unsigned)((B + 3) / 4), 128, 0, (cudaStream_t)stream>>>(t, weights, S, n_fine, u_fine, B, t_out);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
