// Ray generation, world transform, stratified samplers, G1 positional encoder and
// the batched PixelRayDataset gatherer.  HBM-bound, one thread per output element
// group, coalesced stores.  Reference: nerf/model.py:242-278, 336-367, 412-435,
// 216-240 (PE layout); nerf/dataset.py:283-316.
#include "common.cuh"

__global__ void generate_rays_kernel(const float* __restrict__ cam_o, const float* __restrict__ cam_r,
                                     int H, int W, float focal, int64_t ray_begin, int64_t n,
                                     float* __restrict__ rays_o, float* __restrict__ rays_d) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t id = ray_begin + i;
    int col = (int)(id % W); int64_t q = id / W;
    int row = (int)(q % H);  int cam = (int)(q / H);
    float r[3], d[3];
    camera_ray(row, col, H, W, focal, r);
    rotate_ray(cam_r + (size_t)cam * 9, r, d);
    if (rays_d) { rays_d[i * 3 + 0] = d[0]; rays_d[i * 3 + 1] = d[1]; rays_d[i * 3 + 2] = d[2]; }
    if (rays_o) { rays_o[i * 3 + 0] = cam_o[cam * 3 + 0]; rays_o[i * 3 + 1] = cam_o[cam * 3 + 1]; rays_o[i * 3 + 2] = cam_o[cam * 3 + 2]; }
}

extern "C" int nerf_generate_rays(const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                                  float focal, int64_t ray_begin, int64_t ray_end,
                                  float* rays_o, float* rays_d, void* stream) {
    if (!cam_o || !cam_r || n_cams <= 0 || H <= 0 || W <= 0 || !(focal != 0.0f)) return NERF_ERR_BAD_ARG;
    if (ray_begin < 0 || ray_end < ray_begin || ray_end > (int64_t)n_cams * H * W) return NERF_ERR_BAD_ARG;
    int64_t n = ray_end - ray_begin;
    if (n == 0) return NERF_OK;
    generate_rays_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        cam_o, cam_r, H, W, focal, ray_begin, n, rays_o, rays_d);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// ------------------------------------------------------- dataset gatherer ----
__global__ void gather_pixel_rays_kernel(const float* __restrict__ images, const float* __restrict__ poses,
                                         int pose_rows, const float* __restrict__ states_x, int Sx,
                                         int N, int H, int W, int C, float focal,
                                         const int64_t* __restrict__ idx, int64_t B,
                                         float* __restrict__ rays_o, float* __restrict__ rays_d,
                                         float* __restrict__ pixels, float* __restrict__ states_out,
                                         int64_t* __restrict__ bhw) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    int64_t id = idx[i];                       // dataset.py:283-291
    int wi = (int)(id % W); id /= W;
    int hi = (int)(id % H); id /= H;
    int bi = (int)(id % N);
    const float* pose = poses + (size_t)bi * pose_rows * 4;
    float R[9], r[3], d[3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) R[a * 3 + b] = pose[a * 4 + b];
    camera_ray(hi, wi, H, W, focal, r);
    rotate_ray(R, r, d);
#pragma unroll
    for (int a = 0; a < 3; ++a) { rays_o[i * 3 + a] = pose[a * 4 + 3]; rays_d[i * 3 + a] = d[a]; }
    const float* px = images + (((size_t)bi * H + hi) * W + wi) * C;
    for (int c = 0; c < C; ++c) pixels[i * C + c] = px[c];
    if (states_out) for (int s = 0; s < Sx; ++s) states_out[i * Sx + s] = states_x[(size_t)bi * Sx + s];
    if (bhw) { bhw[i * 3 + 0] = bi; bhw[i * 3 + 1] = hi; bhw[i * 3 + 2] = wi; }
}

extern "C" int nerf_gather_pixel_rays(const float* images, const float* poses, int pose_rows,
                                      const float* states_x, int Sx, int N, int H, int W, int C,
                                      float focal, const int64_t* idx, int64_t B,
                                      float* rays_o, float* rays_d, float* pixels, float* states_out,
                                      int64_t* bhw, void* stream) {
    if (!images || !poses || !idx || !rays_o || !rays_d || !pixels) return NERF_ERR_BAD_ARG;
    if ((pose_rows != 3 && pose_rows != 4) || N <= 0 || H <= 0 || W <= 0 || C <= 0 || B < 0) return NERF_ERR_BAD_ARG;
    if (states_out && (!states_x || Sx <= 0)) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    gather_pixel_rays_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        images, poses, pose_rows, states_x, Sx, N, H, W, C, focal, idx, B, rays_o, rays_d, pixels, states_out, bhw);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// ------------------------------------------------------------- samplers ------
// mode 0: ladder = linspace(near, far, S);  mode 1: ladder = 2^linspace(lo, 0, S), result * scale
__device__ __forceinline__ float ladder_at(const float* __restrict__ t_base, int mode, float a, float b, int S, int k) {
    if (t_base) return t_base[k];
    float v = linspace_at(a, b, S, k);
    return mode == 0 ? v : exp2f(v);
}

__global__ void sample_stratified_kernel(const float* __restrict__ t_base, int mode, float a, float b,
                                         float scale, int S, int64_t B, const float* __restrict__ u,
                                         int jitter, uint64_t seed, int64_t ray_id_base,
                                         float* __restrict__ t_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 4 samples
    int64_t total = B * (int64_t)S;
    int64_t e0 = i * 4;
    if (e0 >= total) return;
    float out[4];
    if ((S & 3) == 0) {
        // fast path: the 4 samples belong to one ray and share one Philox block (counter k>>2), one
        // ray/sample split and a 6-entry window of the ladder -- same values as the generic path
        const int64_t ray = (total < (1ll << 31)) ? (int64_t)((uint32_t)e0 / (uint32_t)S) : e0 / S;   // 32-bit divide when it fits
        const int k0 = (int)(e0 - ray * S);
        float lad[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) { const int k = k0 - 1 + j; lad[j] = (k >= 0 && k < S) ? ladder_at(t_base, mode, a, b, S, k) : 0.f; }
        uint32_t r4[4] = {0u, 0u, 0u, 0u};
        float4 uv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (jitter) {
            if (u) uv = *reinterpret_cast<const float4*>(u + e0);
            else { philox4x32(seed, (uint64_t)(ray_id_base + ray), (uint32_t)(k0 >> 2), r4);
                   uv = make_float4(u01_from_bits(r4[0]), u01_from_bits(r4[1]), u01_from_bits(r4[2]), u01_from_bits(r4[3])); }
        }
        const float uu[4] = {uv.x, uv.y, uv.z, uv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = k0 + j; const float tk = lad[j + 1];
            float v = tk;
            if (jitter) {
                float lo = tk, hi = tk;
                if (k > 0)     lo = __fmul_rn(0.5f, __fadd_rn(tk, lad[j]));
                if (k < S - 1) hi = __fmul_rn(0.5f, __fadd_rn(lad[j + 2], tk));
                v = __fadd_rn(lo, __fmul_rn(__fsub_rn(hi, lo), uu[j]));
            }
            out[j] = (mode == 0) ? v : __fmul_rn(v, scale);
        }
    } else
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        int64_t e = e0 + j;
        if (e >= total) { out[j] = 0.f; continue; }
        int64_t ray = e / S; int k = (int)(e - ray * S);
        float tk = ladder_at(t_base, mode, a, b, S, k);
        float v = tk;
        if (jitter) {
            // nerf/model.py:425-433: bins bounded by the midpoints of the ladder
            float lo = tk, hi = tk;
            if (k > 0)     lo = __fmul_rn(0.5f, __fadd_rn(tk, ladder_at(t_base, mode, a, b, S, k - 1)));
            if (k < S - 1) hi = __fmul_rn(0.5f, __fadd_rn(ladder_at(t_base, mode, a, b, S, k + 1), tk));
            float uu;
            if (u) uu = u[e];
            else { uint32_t r4[4]; philox4x32(seed, (uint64_t)(ray_id_base + ray), (uint32_t)(k >> 2), r4); uu = u01_from_bits(r4[k & 3]); }
            v = __fadd_rn(lo, __fmul_rn(__fsub_rn(hi, lo), uu));
        }
        out[j] = (mode == 0) ? v : __fmul_rn(v, scale);
    }
    if (e0 + 3 < total && ((((uintptr_t)(t_out + e0)) & 15) == 0)) {
        *reinterpret_cast<float4*>(t_out + e0) = make_float4(out[0], out[1], out[2], out[3]);
    } else {
        for (int j = 0; j < 4 && e0 + j < total; ++j) t_out[e0 + j] = out[j];
    }
}

// Streaming form for S % 4 == 0 (every configuration of the reference): the stratified bins are the same for
// every ray, so each persistent block builds (lower edge, width) of the S bins once in shared memory with exactly
// the per-sample arithmetic of the generic kernel (0.5*(t_k + t_k-1), hi - lo, separately rounded) and then
// streams: per thread and iteration one Philox block (or one 16-byte load of u), two 16-byte shared loads,
// 4 x (mul, add) and one 16-byte streaming store.  No divisions in the loop when S is a power of two.
// Values are bit-identical to sample_stratified_kernel.
template <bool JITTER, bool HAS_U>
__global__ void __launch_bounds__(256) sample_stream_kernel(const float* __restrict__ t_base, int mode, float a, float b,
                                                            float scale, int S, int log2S, int64_t total,
                                                            const float* __restrict__ u, uint64_t seed, int64_t ray_id_base,
                                                            float* __restrict__ t_out) {
    extern __shared__ __align__(16) float smp_sm[];
    float* s_lo = smp_sm; float* s_w = smp_sm + S;
    for (int k = threadIdx.x; k < S; k += blockDim.x) {
        const float tk = ladder_at(t_base, mode, a, b, S, k);
        float lo = tk, hi = tk;
        if (JITTER) {
            if (k > 0)     lo = __fmul_rn(0.5f, __fadd_rn(tk, ladder_at(t_base, mode, a, b, S, k - 1)));
            if (k < S - 1) hi = __fmul_rn(0.5f, __fadd_rn(ladder_at(t_base, mode, a, b, S, k + 1), tk));
        }
        s_lo[k] = lo; s_w[k] = __fsub_rn(hi, lo);
    }
    __syncthreads();
    const bool small = total < (1ll << 31);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t e0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; e0 < total; e0 += stride) {
        int64_t ray; int k0;
        if (log2S >= 0) { ray = e0 >> log2S; k0 = (int)e0 & (S - 1); }
        else { ray = small ? (int64_t)((uint32_t)e0 / (uint32_t)S) : e0 / S; k0 = (int)(e0 - ray * S); }
        float4 v = *reinterpret_cast<const float4*>(s_lo + k0);
        if (JITTER) {
            const float4 w = *reinterpret_cast<const float4*>(s_w + k0);
            float4 uv;
            if (HAS_U) uv = __ldcs(reinterpret_cast<const float4*>(u + e0));
            else {
                uint32_t r4[4];
                philox4x32(seed, (uint64_t)(ray_id_base + ray), (uint32_t)(k0 >> 2), r4);
                uv = make_float4(u01_from_bits(r4[0]), u01_from_bits(r4[1]), u01_from_bits(r4[2]), u01_from_bits(r4[3]));
            }
            v.x = __fadd_rn(v.x, __fmul_rn(w.x, uv.x)); v.y = __fadd_rn(v.y, __fmul_rn(w.y, uv.y));
            v.z = __fadd_rn(v.z, __fmul_rn(w.z, uv.z)); v.w = __fadd_rn(v.w, __fmul_rn(w.w, uv.w));
        }
        if (mode != 0) { v.x = __fmul_rn(v.x, scale); v.y = __fmul_rn(v.y, scale); v.z = __fmul_rn(v.z, scale); v.w = __fmul_rn(v.w, scale); }
        __stcs(reinterpret_cast<float4*>(t_out + e0), v);
    }
}

static int launch_sampler(const float* t_base, int mode, float a, float b, float scale, int S, int64_t B,
                          const float* u, int jitter, uint64_t seed, int64_t ray_id_base, float* t_out, void* stream) {
    if (B == 0 && S > 0) return NERF_OK;                       // zero rays: nothing to do (empty tensors carry null pointers)
    if (S <= 0 || B < 0 || !t_out) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    int64_t groups = (B * (int64_t)S + 3) / 4;
    if ((S & 3) == 0 && S <= 4096 && (((uintptr_t)t_out) & 15) == 0 && (!u || (((uintptr_t)u) & 15) == 0)) {
        int log2S = -1;
        if ((S & (S - 1)) == 0) { log2S = 0; while ((1 << log2S) < S) ++log2S; }
        const int64_t need = (groups + 255) / 256, cap = (int64_t)nerf_num_sms() * 8;
        const unsigned grid = (unsigned)(need < cap ? need : cap);
        const size_t smem = (size_t)2 * S * sizeof(float);
        const int64_t total = B * (int64_t)S;
        cudaStream_t st = (cudaStream_t)stream;
        if (!jitter) sample_stream_kernel<false, false><<<grid, 256, smem, st>>>(t_base, mode, a, b, scale, S, log2S, total, u, seed, ray_id_base, t_out);
        else if (u)  sample_stream_kernel<true, true><<<grid, 256, smem, st>>>(t_base, mode, a, b, scale, S, log2S, total, u, seed, ray_id_base, t_out);
        else         sample_stream_kernel<true, false><<<grid, 256, smem, st>>>(t_base, mode, a, b, scale, S, log2S, total, u, seed, ray_id_base, t_out);
        NERF_CUDA_CHECK_LAUNCH();
        return NERF_OK;
    }
    sample_stratified_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        t_base, mode, a, b, scale, S, B, u, jitter, seed, ray_id_base, t_out);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int nerf_sample_linear(float near_plane, float far_plane, const float* t_base, int S,
                                  int64_t B, const float* u, int jitter, uint64_t seed,
                                  int64_t ray_id_base, float* t_out, void* stream) {
    return launch_sampler(t_base, 0, near_plane, far_plane, 1.0f, S, B, u, jitter, seed, ray_id_base, t_out, stream);
}

extern "C" int nerf_sample_geometric(float box_diag, const float* t_base, int S, int64_t B,
                                     const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                                     float* t_out, void* stream) {
    return launch_sampler(t_base, 1, -9.43633744014f, 0.0f, box_diag, S, B, u, jitter, seed, ray_id_base, t_out, stream);
}

// ------------------------------------------------------- density noise -------
// sigma += N(0,1) * std  (nerf/model.py:652-654), counter-based so that a ray's
// noise does not depend on how rays are sharded: Philox keyed by (seed ^ tag,
// global ray id, sample/4), Box-Muller on pairs.
__global__ void density_noise_kernel(float* __restrict__ sigma, int S, int64_t B, float std,
                                     uint64_t seed, int64_t ray_id_base) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;     // one thread per 4 samples of a ray
    const int groups = (S + 3) / 4;
    if (i >= B * groups) return;
    int64_t ray = i / groups; int g = (int)(i - ray * groups);
    uint32_t r4[4];
    philox4x32(seed ^ 0x6E6F697365ull, (uint64_t)(ray_id_base + ray), (uint32_t)g, r4);
    float z[4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float u1 = ((float)(r4[2 * h] >> 8) + 1.0f) * (1.0f / 16777216.0f);     // (0,1]
        float u2 = u01_from_bits(r4[2 * h + 1]);
        float rad = sqrtf(-2.0f * logf(u1)), sn, cs;
        sincospif(2.0f * u2, &sn, &cs);
        z[2 * h] = rad * cs; z[2 * h + 1] = rad * sn;
    }
    for (int j = 0; j < 4; ++j) { int k = g * 4 + j; if (k < S) sigma[ray * S + k] += z[j] * std; }
}

extern "C" int nerf_add_density_noise(float* sigma, int S, int64_t B, float std, uint64_t seed,
                                      int64_t ray_id_base, void* stream) {
    if (B == 0 && S > 0) return NERF_OK;
    if (!sigma || S <= 0 || B < 0) return NERF_ERR_BAD_ARG;
    if (B == 0 || std == 0.0f) return NERF_OK;
    int64_t n = B * ((S + 3) / 4);
    density_noise_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(sigma, S, B, std, seed, ray_id_base);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
