// Volume-rendering composite: forward warp product-scan and backward reverse
// (suffix) scan.  One warp per ray, 32 samples per segment, any S.
// Reference arithmetic: nerf/model.py:462-469 (weights), :660 (colour).
//   delta_k = |p_{k+1} - p_k| (last = 1e10);  a_k = exp(-relu(sigma_k) delta_k)
//   T_k = prod_{j<k} (a_j + 1e-10);  w_k = (1 - a_k) T_k;  pixel = sum_k w_k sigmoid(rgb_k)
// HBM-bound: 20 S + 20 (+24) bytes per ray forward, 36 S + 40 (+24) backward.
#include "common.cuh"
#include "composite_tile.cuh"

struct RayGeom {            // how sample positions are obtained
    const float* points;    // [B,S,3] or null
    const float* t;         // [B,S]   (needed if points == null, or for depth)
    float o[3], d[3];
};

__device__ __forceinline__ void point_at(const RayGeom& g, int64_t base, int k, float (&p)[3]) {
    if (g.points) {
        const float* q = g.points + (base + k) * 3;
        p[0] = q[0]; p[1] = q[1]; p[2] = q[2];
    } else {
        float tk = g.t[base + k];
#pragma unroll
        for (int c = 0; c < 3; ++c) p[c] = __fadd_rn(g.o[c], __fmul_rn(g.d[c], tk));
    }
}
__device__ __forceinline__ float delta_at(const RayGeom& g, int64_t base, int k, int S) {
    if (k >= S - 1) return NERF_LAST_DIST;
    float p0[3], p1[3];
    point_at(g, base, k, p0); point_at(g, base, k + 1, p1);
    float dx = __fsub_rn(p1[0], p0[0]), dy = __fsub_rn(p1[1], p0[1]), dz = __fsub_rn(p1[2], p0[2]);
    return sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// exclusive product scan over the warp; returns T for this lane, updates carry
__device__ __forceinline__ float warp_excl_prod(float v, float& carry, int lane) {
    float incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { float n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl *= n; }
    float prev = __shfl_up_sync(0xffffffffu, incl, 1);
    float T = carry * (lane == 0 ? 1.0f : prev);
    carry = carry * __shfl_sync(0xffffffffu, incl, 31);
    return T;
}

__device__ __forceinline__ void load_geom(RayGeom& g, const float* points, const float* rays_o,
                                          const float* rays_d, const float* t, int64_t ray) {
    g.points = points; g.t = t;
    if (!points) {
#pragma unroll
        for (int c = 0; c < 3; ++c) { g.o[c] = rays_o[ray * 3 + c]; g.d[c] = rays_d[ray * 3 + c]; }
    }
}

__global__ void __launch_bounds__(256) composite_fwd_kernel(
        const float* __restrict__ sigma, const float* __restrict__ rgb, const float* __restrict__ points,
        const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ t,
        int64_t B, int S, float* __restrict__ pixels, float* __restrict__ depth, float* __restrict__ acc,
        float* __restrict__ weights) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < B; ray += warps) {
        RayGeom g; load_geom(g, points, rays_o, rays_d, t, ray);
        const int64_t base = ray * S;
        float carry = 1.0f, pr = 0.f, pg = 0.f, pb = 0.f, dep = 0.f, ac = 0.f;
        for (int s0 = 0; s0 < S; s0 += 32) {
            const int k = s0 + lane; const bool valid = k < S;
            float a = 1.0f, one_minus_eps_a = 1.0f;
            if (valid) {
                float sg = fmaxf(sigma[base + k], 0.0f);
                a = expf(__fmul_rn(-sg, delta_at(g, base, k, S)));
                one_minus_eps_a = __fadd_rn(a, NERF_EPS_T);
            }
            float T = warp_excl_prod(one_minus_eps_a, carry, lane);
            if (valid) {
                float w = __fmul_rn(__fsub_rn(1.0f, a), T);
                if (weights) weights[base + k] = w;
                const float* c = rgb + (base + k) * 3;
                pr += w * sigmoidf_(c[0]); pg += w * sigmoidf_(c[1]); pb += w * sigmoidf_(c[2]);
                ac += w;
                if (depth) dep += w * t[base + k];
            }
        }
        pr = warp_sum(pr); pg = warp_sum(pg); pb = warp_sum(pb); ac = warp_sum(ac); dep = warp_sum(dep);
        if (lane == 0) {
            if (pixels) { pixels[ray * 3] = pr; pixels[ray * 3 + 1] = pg; pixels[ray * 3 + 2] = pb; }
            if (depth) depth[ray] = dep;
            if (acc) acc[ray] = ac;
        }
    }
}

template <int NSEG>
__global__ void __launch_bounds__(256) composite_bwd_kernel(
        const float* __restrict__ sigma, const float* __restrict__ rgb, const float* __restrict__ points,
        const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ t,
        int64_t B, int S, const float* __restrict__ g_pixels, const float* __restrict__ g_depth,
        const float* __restrict__ g_acc, const float* __restrict__ g_weights,
        float* __restrict__ d_sigma, float* __restrict__ d_rgb) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < B; ray += warps) {
        RayGeom g; load_geom(g, points, rays_o, rays_d, t, ray);
        const int64_t base = ray * S;
        const float gr = g_pixels ? g_pixels[ray * 3] : 0.f, gg = g_pixels ? g_pixels[ray * 3 + 1] : 0.f,
                    gb = g_pixels ? g_pixels[ray * 3 + 2] : 0.f;
        const float gd = g_depth ? g_depth[ray] : 0.f, ga = g_acc ? g_acc[ray] : 0.f;
        float coef[NSEG], Ts[NSEG], ws[NSEG], inv[NSEG];   // per segment, this lane's sample
        float carry = 1.0f;
#pragma unroll
        for (int seg = 0; seg < NSEG; ++seg) {
            const int k = seg * 32 + lane; const bool valid = k < S;
            float a = 1.0f, ae = 1.0f, dl = 0.f, sg = 0.f;
            if (valid) {
                sg = sigma[base + k];
                dl = delta_at(g, base, k, S);
                a = expf(__fmul_rn(-fmaxf(sg, 0.f), dl));
                ae = __fadd_rn(a, NERF_EPS_T);
            }
            float T = (seg * 32 < S) ? warp_excl_prod(ae, carry, lane) : 1.0f;
            float s = 0.f, w = 0.f;
            if (valid) {
                w = __fmul_rn(__fsub_rn(1.0f, a), T);
                const float* c = rgb + (base + k) * 3;
                float s0 = sigmoidf_(c[0]), s1 = sigmoidf_(c[1]), s2 = sigmoidf_(c[2]);
                s = gr * s0 + gg * s1 + gb * s2 + ga;
                if (g_depth) s += gd * t[base + k];
                if (g_weights) s += g_weights[base + k];
                float* o = d_rgb + (base + k) * 3;
                o[0] = w * gr * s0 * (1.0f - s0); o[1] = w * gg * s1 * (1.0f - s1); o[2] = w * gb * s2 * (1.0f - s2);
            }
            coef[seg] = (valid && sg > 0.f) ? dl * a : 0.f;    // delta a relu'(sigma)
            Ts[seg] = T * s; ws[seg] = w * s; inv[seg] = 1.0f / ae;
        }
        float tail = 0.f;                                     // sum over later segments
#pragma unroll
        for (int seg = NSEG - 1; seg >= 0; --seg) {
            if (seg * 32 >= S) continue;
            float incl = ws[seg];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { float n = __shfl_down_sync(0xffffffffu, incl, o); if (lane + o < 32) incl += n; }
            float suffix = tail + (incl - ws[seg]);           // sum_{j>k} w_j s_j
            const int k = seg * 32 + lane;
            if (k < S) d_sigma[base + k] = coef[seg] * (Ts[seg] - suffix * inv[seg]);
            tail += __shfl_sync(0xffffffffu, incl, 0);
        }
    }
}

// The loss side of a training step in one pass per ray: sigma += N(0,1) std (the draw of density_noise_kernel),
// composite forward, mean((pixel - target)^2) with its gradient, composite backward.  Same arithmetic, in the same
// order, as density_noise_kernel -> composite_fwd_kernel -> mse_kernel -> composite_bwd_kernel<NSEG>; the ray's
// samples stay in registers between the sweeps (S <= 32 NSEG).
template <int NSEG>
__global__ void __launch_bounds__(256) composite_mse_train_kernel(
        const float* __restrict__ sigma, const float* __restrict__ rgb,
        const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ t,
        const float* __restrict__ target, int64_t B, int S, float noise_std, uint64_t seed, int64_t ray_id_base,
        float* __restrict__ pixels, float* __restrict__ loss, float* __restrict__ d_sigma, float* __restrict__ d_rgb) {
    __shared__ float part[8];
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float inv_n = 1.0f / (float)(B * 3);
    float sq = 0.f;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < B; ray += warps) {
        RayGeom g; load_geom(g, nullptr, rays_o, rays_d, t, ray);
        const int64_t base = ray * S;
        float coef[NSEG], Ts[NSEG], ws[NSEG], inv[NSEG], c0[NSEG], c1[NSEG], c2[NSEG];
        float carry = 1.0f, pr = 0.f, pg = 0.f, pb = 0.f;
#pragma unroll
        for (int seg = 0; seg < NSEG; ++seg) {
            const int k = seg * 32 + lane; const bool valid = k < S;
            float a = 1.0f, ae = 1.0f, dl = 0.f, sg = 0.f;
            if (valid) {
                sg = sigma[base + k];
                if (noise_std != 0.0f) {
                    uint32_t r4[4];
                    philox4x32(seed ^ 0x6E6F697365ull, (uint64_t)(ray_id_base + ray), (uint32_t)(k >> 2), r4);
                    const int h = (k >> 1) & 1;
                    float u1 = ((float)(r4[2 * h] >> 8) + 1.0f) * (1.0f / 16777216.0f);
                    float u2 = u01_from_bits(r4[2 * h + 1]);
                    float rad = sqrtf(-2.0f * logf(u1)), sn, cs;
                    sincospif(2.0f * u2, &sn, &cs);
                    sg += ((k & 1) ? rad * sn : rad * cs) * noise_std;
                }
                dl = delta_at(g, base, k, S);
                a = expf(__fmul_rn(-fmaxf(sg, 0.f), dl));
                ae = __fadd_rn(a, NERF_EPS_T);
            }
            float T = (seg * 32 < S) ? warp_excl_prod(ae, carry, lane) : 1.0f;
            float w = 0.f; c0[seg] = c1[seg] = c2[seg] = 0.f;
            if (valid) {
                w = __fmul_rn(__fsub_rn(1.0f, a), T);
                const float* c = rgb + (base + k) * 3;
                c0[seg] = sigmoidf_(c[0]); c1[seg] = sigmoidf_(c[1]); c2[seg] = sigmoidf_(c[2]);
                pr += w * c0[seg]; pg += w * c1[seg]; pb += w * c2[seg];
            }
            coef[seg] = (valid && sg > 0.f) ? dl * a : 0.f;
            Ts[seg] = T; ws[seg] = w; inv[seg] = 1.0f / ae;
        }
        pr = warp_sum(pr); pg = warp_sum(pg); pb = warp_sum(pb);
        const float er = pr - target[ray * 3], eg = pg - target[ray * 3 + 1], eb = pb - target[ray * 3 + 2];
        const float gr = 2.0f * er * inv_n, gg = 2.0f * eg * inv_n, gb = 2.0f * eb * inv_n;
        if (lane == 0) {
            sq += er * er; sq += eg * eg; sq += eb * eb;
            if (pixels) { pixels[ray * 3] = pr; pixels[ray * 3 + 1] = pg; pixels[ray * 3 + 2] = pb; }
        }
#pragma unroll
        for (int seg = 0; seg < NSEG; ++seg) {
            const int k = seg * 32 + lane;
            const float s0 = c0[seg], s1 = c1[seg], s2 = c2[seg], w = ws[seg];
            const float s = gr * s0 + gg * s1 + gb * s2 + 0.f;
            if (k < S) {
                float* o = d_rgb + (base + k) * 3;
                o[0] = w * gr * s0 * (1.0f - s0); o[1] = w * gg * s1 * (1.0f - s1); o[2] = w * gb * s2 * (1.0f - s2);
            }
            Ts[seg] *= s; ws[seg] = w * s;
        }
        float tail = 0.f;
#pragma unroll
        for (int seg = NSEG - 1; seg >= 0; --seg) {
            if (seg * 32 >= S) continue;
            float incl = ws[seg];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { float n = __shfl_down_sync(0xffffffffu, incl, o); if (lane + o < 32) incl += n; }
            float suffix = tail + (incl - ws[seg]);
            const int k = seg * 32 + lane;
            if (k < S) d_sigma[base + k] = coef[seg] * (Ts[seg] - suffix * inv[seg]);
            tail += __shfl_sync(0xffffffffu, incl, 0);
        }
    }
    if (lane == 0) part[threadIdx.x >> 5] = sq;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < 8 ? part[threadIdx.x] : 0.f;
        v = warp_sum(v);
        if (threadIdx.x == 0) atomicAdd(loss, v * inv_n);
    }
}

static unsigned grid_for_rays(int64_t B) {
    int64_t blocks = (B + 7) / 8;                              // 8 warps (rays) per 256-thread block
    int64_t cap = (int64_t)nerf_num_sms() * 8 * 4;             // grid-stride beyond 4 waves
    return (unsigned)(blocks < cap ? blocks : cap);
}

// A thread per ray needs many rays to cover the machine: below these counts the warp-per-ray kernels are faster
// (measured on B200 at S = 32..256: 4096 x 64 takes 12 + 11 us against 18 + 25 us; equal at 2^14 / 2^15 rays).
#define CT_MIN_RAYS_FWD 16384
#define CT_MIN_RAYS_BWD 32768

extern "C" int nerf_composite_fwd(const float* sigma, const float* rgb, const float* points,
                                  const float* rays_o, const float* rays_d, const float* t,
                                  int64_t B, int S, float* pixels, float* depth, float* acc,
                                  float* weights, void* stream) {
    if (B == 0 && S > 0) return NERF_OK;
    if (!sigma || B < 0 || S <= 0) return NERF_ERR_BAD_ARG;
    if (!points && (!rays_o || !rays_d || !t)) return NERF_ERR_BAD_ARG;
    if (depth && !t) return NERF_ERR_BAD_ARG;
    if (pixels && !rgb) return NERF_ERR_BAD_ARG;
    if (!rgb) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    if (!points && B >= CT_MIN_RAYS_FWD && S % CT_CH == 0 && ct_aligned16(sigma) && ct_aligned16(rgb) && ct_aligned16(t) && (!weights || ct_aligned16(weights))) {
        composite_fwd_tile_kernel<<<(unsigned)((B + CT_RAYS - 1) / CT_RAYS), CT_RAYS, 0, (cudaStream_t)stream>>>(
            sigma, rgb, rays_o, rays_d, t, B, S, pixels, depth, acc, weights);
        NERF_CUDA_CHECK_LAUNCH();
        return NERF_OK;
    }
    composite_fwd_kernel<<<grid_for_rays(B), 256, 0, (cudaStream_t)stream>>>(
        sigma, rgb, points, rays_o, rays_d, t, B, S, pixels, depth, acc, weights);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int nerf_composite_bwd(const float* sigma, const float* rgb, const float* points,
                                  const float* rays_o, const float* rays_d, const float* t,
                                  int64_t B, int S, const float* g_pixels, const float* g_depth,
                                  const float* g_acc, const float* g_weights,
                                  const float* saved_pixels, const float* saved_depth, const float* saved_acc,
                                  float* d_sigma, float* d_rgb, void* stream) {
    if (B == 0 && S > 0) return NERF_OK;
    if (!sigma || !rgb || !d_sigma || !d_rgb || B < 0 || S <= 0) return NERF_ERR_BAD_ARG;
    if (!points && (!rays_o || !rays_d || !t)) return NERF_ERR_BAD_ARG;
    if (g_depth && !t) return NERF_ERR_BAD_ARG;
    if (S > 1024) return NERF_ERR_UNSUPPORTED;
    if (B == 0) return NERF_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (!points && !g_weights && B >= CT_MIN_RAYS_BWD && S % CT_CH == 0 && ct_aligned16(sigma) && ct_aligned16(rgb) && ct_aligned16(t) &&
        ct_aligned16(d_sigma) && ct_aligned16(d_rgb)) {
        composite_bwd_tile_kernel<<<(unsigned)((B + CT_RAYS - 1) / CT_RAYS), CT_RAYS, 0, st>>>(
            sigma, rgb, rays_o, rays_d, t, B, S, g_pixels, g_depth, g_acc, saved_pixels, saved_depth, saved_acc, d_sigma, d_rgb);
        NERF_CUDA_CHECK_LAUNCH();
        return NERF_OK;
    }
    unsigned grid = grid_for_rays(B);
#define LAUNCH_BWD(N) composite_bwd_kernel<N><<<grid, 256, 0, st>>>(sigma, rgb, points, rays_o, rays_d, t, B, S, \
        g_pixels, g_depth, g_acc, g_weights, d_sigma, d_rgb)
    if (S <= 32) LAUNCH_BWD(1); else if (S <= 64) LAUNCH_BWD(2); else if (S <= 128) LAUNCH_BWD(4);
    else if (S <= 256) LAUNCH_BWD(8); else LAUNCH_BWD(32);
#undef LAUNCH_BWD
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int nerf_composite_mse_train(const float* sigma, const float* rgb, const float* rays_o, const float* rays_d,
                                        const float* t, const float* target, int64_t B, int S,
                                        float noise_std, uint64_t seed, int64_t ray_id_base,
                                        float* pixels, float* loss_out, float* d_sigma, float* d_rgb, void* stream) {
    if (!loss_out || B < 0 || S <= 0) return NERF_ERR_BAD_ARG;
    if (S > 256) return NERF_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    NERF_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float), st));
    if (B == 0) return NERF_OK;
    if (!sigma || !rgb || !rays_o || !rays_d || !t || !target || !d_sigma || !d_rgb) return NERF_ERR_BAD_ARG;
    unsigned grid = grid_for_rays(B);
#define LAUNCH_CMT(N) composite_mse_train_kernel<N><<<grid, 256, 0, st>>>(sigma, rgb, rays_o, rays_d, t, target, B, S, \
        noise_std, seed, ray_id_base, pixels, loss_out, d_sigma, d_rgb)
    if (S <= 32) LAUNCH_CMT(1); else if (S <= 64) LAUNCH_CMT(2); else if (S <= 128) LAUNCH_CMT(4); else LAUNCH_CMT(8);
#undef LAUNCH_CMT
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
