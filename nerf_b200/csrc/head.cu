// HEAD generation of the reference model (the mip-NeRF style network whose source IS in the reference
// tree): integrated positional encoding of conical frustums, the Linear->LayerNorm->ReLU trunk with a
// 1+3+50 output head, and the colour + segmentation composite.  SURVEY 8(a) rows a5-H, a6-H, a7, a11.
// fp32 throughout: this is the <=1e-3 parity path, pinned against outputs of the unmodified reference
// (tests/golden/head_ref.npz).
#include "mlp_fp32_gemm.cuh"

// ------------------------------------------------------------ a5-H: encoder ----
// nerf/model.py:112-136 (cast_rays), :74-81 (conical_frustum_to_gaussian, stable form), :35-45
// (lift_gaussian, diagonal), :139-163 (integrated_pos_enc) and :24-30 (expected_sin):
//   per interval [t0,t1]: mu=(t0+t1)/2, hw=(t1-t0)/2,
//   t_mean = mu + 2 mu hw^2 / (3 mu^2 + hw^2)
//   t_var  = hw^2/3 - 4/15 hw^4 (12 mu^2 - hw^2) / (3 mu^2 + hw^2)^2
//   r_var  = r^2 (mu^2/4 + 5/12 hw^2 - 4/15 hw^4 / (3 mu^2 + hw^2))
//   mean = o + d t_mean;  cov = t_var d^2 + r_var (1 - d^2 / max(|d|^2, 1e-10))
//   feat[deg*3+c] = exp(-cov_c 4^deg / 2) sin(mean_c 2^deg),  feat[3 n_deg + deg*3+c] = same with sin(. + pi/2)
// One thread per (interval, degree): consecutive threads write consecutive triples (coalesced).
__global__ void head_encode_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                   const float* __restrict__ t, int64_t B, int S, float base_radius,
                                   int min_deg, int n_deg, float* __restrict__ mean_out,
                                   float* __restrict__ cov_out, float* __restrict__ feat_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n_int = B * (int64_t)(S - 1);
    if (i >= n_int * n_deg) return;
    const int64_t iv = i / n_deg; const int deg = (int)(i - iv * n_deg);
    const int64_t ray = iv / (S - 1); const int k = (int)(iv - ray * (S - 1));
    const float t0 = t[ray * S + k], t1 = t[ray * S + k + 1];
    const float mu = __fmul_rn(__fadd_rn(t0, t1), 0.5f), hw = __fmul_rn(__fsub_rn(t1, t0), 0.5f);
    const float mu2 = __fmul_rn(mu, mu), hw2 = __fmul_rn(hw, hw), hw4 = __fmul_rn(hw2, hw2);
    const float den = __fadd_rn(__fmul_rn(3.0f, mu2), hw2);
    const float t_mean = __fadd_rn(mu, __fdiv_rn(__fmul_rn(__fmul_rn(2.0f, mu), hw2), den));
    const float t_var = __fsub_rn(__fdiv_rn(hw2, 3.0f),
                                  __fmul_rn((float)(4.0 / 15.0), __fdiv_rn(__fmul_rn(hw4, __fsub_rn(__fmul_rn(12.0f, mu2), hw2)), __fmul_rn(den, den))));
    const float r2 = __fmul_rn(base_radius, base_radius);
    const float r_var = __fmul_rn(r2, __fsub_rn(__fadd_rn(__fdiv_rn(mu2, 4.0f), __fmul_rn((float)(5.0 / 12.0), hw2)),
                                                __fdiv_rn(__fmul_rn((float)(4.0 / 15.0), hw4), den)));
    float d[3], o[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { d[c] = rays_d[ray * 3 + c]; o[c] = rays_o[ray * 3 + c]; }
    const float mag = fmaxf(__fadd_rn(__fadd_rn(__fmul_rn(d[0], d[0]), __fmul_rn(d[1], d[1])), __fmul_rn(d[2], d[2])), 1e-10f);
    const float scale = exp2f((float)(min_deg + deg)), scale2 = __fmul_rn(scale, scale);
    float* f = feat_out + iv * (6 * n_deg);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float dsq = __fmul_rn(d[c], d[c]);
        const float mean = __fadd_rn(__fmul_rn(d[c], t_mean), o[c]);
        const float cov = __fadd_rn(__fmul_rn(t_var, dsq), __fmul_rn(r_var, __fsub_rn(1.0f, __fdiv_rn(dsq, mag))));
        if (deg == 0) { if (mean_out) mean_out[iv * 3 + c] = mean; if (cov_out) cov_out[iv * 3 + c] = cov; }
        const float y = __fmul_rn(mean, scale), yv = __fmul_rn(cov, scale2);
        const float amp = expf(__fmul_rn(-0.5f, yv));
        f[deg * 3 + c] = __fmul_rn(amp, sinf(y));
        f[3 * n_deg + deg * 3 + c] = __fmul_rn(amp, sinf(__fadd_rn(y, (float)(0.5 * 3.141592653589793))));
    }
}

extern "C" int nerf_head_encode(const float* rays_o, const float* rays_d, const float* t, int64_t B, int S,
                                float base_radius, int min_deg, int n_deg, float* mean_out, float* cov_out,
                                float* feat_out, void* stream) {
    if (B == 0 && S >= 2 && n_deg > 0) return NERF_OK;
    if (!rays_o || !rays_d || !t || !feat_out || B < 0 || S < 2 || n_deg <= 0 || n_deg > 32) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    const int64_t n = B * (int64_t)(S - 1) * n_deg;
    head_encode_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        rays_o, rays_d, t, B, S, base_radius, min_deg, n_deg, mean_out, cov_out, feat_out);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// -------------------------------------------------------------- a6-H: trunk ----
// prediction_heads (nerf/model.py:525-542): 5 x (Linear -> LayerNorm(1e-5, affine) -> ReLU), then
// Linear 256 -> n_out.  params_host: 22 DEVICE pointers in state_dict order
// (W0,b0,g0,beta0, W1,b1,g1,beta1, ..., W4,b4,g4,beta4, W5,b5).  Workspace: two [P,256] buffers.
extern "C" int64_t nerf_head_mlp_workspace_bytes(int64_t P) {
    if (P < 0) return NERF_ERR_BAD_ARG;
    return 2 * ((P * HID + 63) & ~(int64_t)63) * (int64_t)sizeof(float);
}

extern "C" int nerf_head_mlp_fwd(const float* const* params_host, const float* feat, int64_t P, int n_in, int n_out,
                                 float* out, void* workspace, int64_t workspace_bytes, void* stream) {
    if (!params_host || !feat || !out || P < 0 || n_in <= 0 || (n_in & 3) || n_out <= 0 || n_out > HID) return NERF_ERR_BAD_ARG;
    if (P == 0) return NERF_OK;
    if (!workspace || workspace_bytes < nerf_head_mlp_workspace_bytes(P)) return NERF_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    float* buf[2] = {(float*)workspace, (float*)workspace + ((P * HID + 63) & ~(int64_t)63)};
    const unsigned tiles = (unsigned)((P + BM - 1) / BM);
    for (int l = 0; l < 5; ++l) {
        ASource s{};
        s.row_div2 = 1;
        if (l == 0) { s.seg2 = feat; s.K2 = n_in; }
        else { s.n_prev = buf[(l - 1) & 1]; s.gamma = params_host[4 * (l - 1) + 2]; s.beta = params_host[4 * (l - 1) + 3]; s.K1 = HID; s.relu = 1; }
        mlp_tile_kernel<MODE_FWD_LN><<<tiles, 256, 0, st>>>(s, params_host[4 * l], l == 0 ? n_in : HID, params_host[4 * l + 1],
                                                           buf[l & 1], nullptr, nullptr, P, HID);
        NERF_CUDA_CHECK_LAUNCH();
    }
    ASource s{};
    s.row_div2 = 1; s.n_prev = buf[0]; s.gamma = params_host[18]; s.beta = params_host[19]; s.K1 = HID; s.relu = 1;
    mlp_tile_kernel<MODE_LINEAR><<<tiles, 256, 0, st>>>(s, params_host[20], HID, params_host[21], out, nullptr, nullptr, P, n_out);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// ------------------------------------------- a9-a11: colour + segmentation composite ----
// nerf/model.py:656-663 on the trunk's interleaved output rows [density | colour(3) | segmentation(n_seg)]:
//   w = alpha_compositing_coefficients(points = Gaussian means, density)        (:437-469)
//   image = sum_k w_k sigmoid(colour_k)                                        (:660)
//   seg_c = logsumexp_k( log(w_k + 1e-10) + log_softmax(seg_k)_c )             (:661-663)
// One warp per ray.  Per 32-sample segment: lane = sample for the transmittance product scan, then
// lane = class for the segmentation: the warp reads one sample's class row coalesced, reduces its
// log-softmax normaliser, and every class lane folds the sample into its own running logsumexp.
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(256) head_composite_kernel(
        const float* __restrict__ out, int n_out, int n_seg, const float* __restrict__ points,
        const float* __restrict__ noise, float noise_std, int64_t B, int S,
        float* __restrict__ pixels, float* __restrict__ seg_out, float* __restrict__ weights, int seg_is_log_softmax) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < B; ray += warps) {
        const int64_t base = ray * S;
        float carry = 1.0f, pr = 0.f, pg = 0.f, pb = 0.f;
        float M0 = -INFINITY, S0 = 0.f, M1 = -INFINITY, S1 = 0.f;       // running logsumexp of classes lane, lane+32
        for (int s0 = 0; s0 < S; s0 += 32) {
            const int k = s0 + lane; const bool valid = k < S;
            float a = 1.0f, ae = 1.0f, w = 0.f;
            if (valid) {
                float sg = out[(base + k) * n_out];
                if (noise) sg = __fadd_rn(sg, __fmul_rn(noise[base + k], noise_std));
                float dl = NERF_LAST_DIST;
                if (k < S - 1) {
                    const float* p0 = points + (base + k) * 3;
                    const float dx = __fsub_rn(p0[3], p0[0]), dy = __fsub_rn(p0[4], p0[1]), dz = __fsub_rn(p0[5], p0[2]);
                    dl = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
                }
                a = expf(__fmul_rn(-fmaxf(sg, 0.f), dl));
                ae = __fadd_rn(a, NERF_EPS_T);
            }
            // exclusive product scan of (a + eps)
            float incl = ae;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { float n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl *= n; }
            const float prev = __shfl_up_sync(0xffffffffu, incl, 1);
            const float T = carry * (lane == 0 ? 1.0f : prev);
            carry = carry * __shfl_sync(0xffffffffu, incl, 31);
            if (valid) {
                w = __fmul_rn(__fsub_rn(1.0f, a), T);
                if (weights) weights[base + k] = w;
                const float* c = out + (base + k) * n_out + 1;
                pr += w / (1.0f + expf(-c[0])); pg += w / (1.0f + expf(-c[1])); pb += w / (1.0f + expf(-c[2]));
            }
            if (seg_out) {
                const int nv = (S - s0 < 32) ? S - s0 : 32;
                for (int kk = 0; kk < nv; ++kk) {
                    const float lw = logf(__fadd_rn(__shfl_sync(0xffffffffu, w, kk), 1e-10f));
                    const float* row = out + (base + s0 + kk) * n_out + 4;
                    const float x0 = lane < n_seg ? row[lane] : -INFINITY;
                    const float x1 = lane + 32 < n_seg ? row[lane + 32] : -INFINITY;
                    float lse = 0.f;
                    if (!seg_is_log_softmax) {
                        const float m = warp_max(fmaxf(x0, x1));
                        lse = m + logf(warp_sum(expf(x0 - m) + expf(x1 - m)));
                    }
                    const float v0 = lw + (x0 - lse), v1 = lw + (x1 - lse);
                    if (lane < n_seg) { const float mn = fmaxf(M0, v0); S0 = S0 * expf(M0 - mn) + expf(v0 - mn); M0 = mn; }
                    if (lane + 32 < n_seg) { const float mn = fmaxf(M1, v1); S1 = S1 * expf(M1 - mn) + expf(v1 - mn); M1 = mn; }
                }
            }
        }
        pr = warp_sum(pr); pg = warp_sum(pg); pb = warp_sum(pb);
        if (lane == 0 && pixels) { pixels[ray * 3] = pr; pixels[ray * 3 + 1] = pg; pixels[ray * 3 + 2] = pb; }
        if (seg_out) {
            if (lane < n_seg) seg_out[ray * n_seg + lane] = M0 + logf(S0);
            if (lane + 32 < n_seg) seg_out[ray * n_seg + lane + 32] = M1 + logf(S1);
        }
    }
}

extern "C" int nerf_head_composite(const float* out, int n_out, int n_seg, const float* points, const float* noise,
                                   float noise_std, int64_t B, int S, float* pixels, float* seg_out, float* weights,
                                   int seg_is_log_softmax, void* stream) {
    if (B == 0 && S > 0) return NERF_OK;
    if (!out || !points || B < 0 || S <= 0 || n_seg < 0 || n_seg > 64 || n_out != 4 + n_seg) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    int64_t blocks = (B + 7) / 8, cap = (int64_t)nerf_num_sms() * 32;
    head_composite_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(
        out, n_out, n_seg, points, noise, noise_std, B, S, pixels, n_seg ? seg_out : nullptr, weights, seg_is_log_softmax);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
