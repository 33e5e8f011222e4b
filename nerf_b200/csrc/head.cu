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
                const float lw_lane = logf(__fadd_rn(w, 1e-10f));       // once per lane (= sample) here, not once per sample by every lane
                for (int kk = 0; kk < nv; ++kk) {
                    const float lw = __shfl_sync(0xffffffffu, lw_lane, kk);
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

// ============================================================ HEAD backward (fp32) ====
// The reference ships no training script for this generation, but its model is an nn.Module whose
// render_rays is differentiable (autograd through nerf/model.py:596-668); these kernels are that
// backward: composite (colour + segmentation logsumexp) -> output Linear -> 5 x (ReLU, LayerNorm, Linear).
#include "mlp_fp32_rows.cuh"

// ---- autograd of nerf/model.py:652-663 w.r.t. the trunk's output rows -----------------------------
//   pixel_c = sum_k w_k sigmoid(c_kc),   seg_c = logsumexp_k v_kc,  v_kc = log(w_k + 1e-10) + log_softmax(x_k)_c
//   d c_kc = dP_c w_k s (1 - s)
//   d v_kc = dS_c exp(v_kc - seg_c);  d x_kc = d v_kc - softmax(x_k)_c sum_c' d v_kc'
//   d w_k  = sum_c dP_c s_kc + (sum_c d v_kc) / (w_k + 1e-10)
//   w_k = (1 - a_k) T_k, T_k = prod_{j<k} (a_j + 1e-10):
//   d a_k  = -T_k d w_k + (sum_{j>k} d w_j w_j) / (a_k + 1e-10);   d sigma_k = [sigma_k > 0] (-delta_k a_k) d a_k
// One warp per ray, per-ray scratch (a, T, d w, -delta a mask) in shared memory: pass A lane = sample
// (forward product scan, colour terms), pass B lane = class (one sample row at a time, coalesced),
// pass C lane = sample in reverse (suffix-sum scan).  d_out rows are [P, ldo] with the columns past
// n_out zeroed: the row layout the trunk's weight-gradient / data-gradient tiles read with 16-byte loads.
__global__ void __launch_bounds__(256) head_composite_bwd_kernel(
        const float* __restrict__ out, int n_out, int n_seg, const float* __restrict__ points,
        const float* __restrict__ noise, float noise_std, int64_t B, int S,
        const float* __restrict__ seg_fwd, const float* __restrict__ d_pixels, const float* __restrict__ d_seg,
        float* __restrict__ d_out, int ldo) {
    extern __shared__ float head_bwd_sm[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    float* sa = head_bwd_sm + (size_t)wib * 4 * S; float* sT = sa + S; float* sdw = sT + S; float* sfac = sdw + S;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < B; ray += warps) {
        const int64_t base = ray * S;
        float dP[3] = {0.f, 0.f, 0.f};
        if (d_pixels) { dP[0] = d_pixels[ray * 3]; dP[1] = d_pixels[ray * 3 + 1]; dP[2] = d_pixels[ray * 3 + 2]; }
        // ---- pass A: forward transmittance scan, colour gradients, d w (colour part) ----
        float carry = 1.0f;
        for (int s0 = 0; s0 < S; s0 += 32) {
            const int k = s0 + lane; const bool valid = k < S;
            float a = 1.0f, ae = 1.0f, fac = 0.f;
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
                fac = sg > 0.f ? -dl * a : 0.f;
            }
            float incl = ae;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { float n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl *= n; }
            const float prev = __shfl_up_sync(0xffffffffu, incl, 1);
            const float T = carry * (lane == 0 ? 1.0f : prev);
            carry = carry * __shfl_sync(0xffffffffu, incl, 31);
            if (valid) {
                const float w = __fmul_rn(__fsub_rn(1.0f, a), T);
                const float* c = out + (base + k) * n_out + 1;
                float* g = d_out + (base + k) * (int64_t)ldo;
                float dwc = 0.f;
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    const float s = 1.0f / (1.0f + expf(-c[ch]));
                    dwc = fmaf(dP[ch], s, dwc);
                    g[1 + ch] = dP[ch] * w * s * (1.0f - s);
                }
                sa[k] = a; sT[k] = T; sdw[k] = dwc; sfac[k] = fac;
            }
        }
        __syncwarp();
        // ---- pass B: segmentation logsumexp, lane = class ----
        const bool seg_on = n_seg > 0 && d_seg != nullptr;
        float sc0 = 0.f, sc1 = 0.f, dS0 = 0.f, dS1 = 0.f;
        if (seg_on) {
            if (lane < n_seg) { sc0 = seg_fwd[ray * n_seg + lane]; dS0 = d_seg[ray * n_seg + lane]; }
            if (lane + 32 < n_seg) { sc1 = seg_fwd[ray * n_seg + lane + 32]; dS1 = d_seg[ray * n_seg + lane + 32]; }
        }
        for (int k = 0; k < S; ++k) {
            float* g = d_out + (base + k) * (int64_t)ldo + 4;
            float dx0 = 0.f, dx1 = 0.f;
            if (seg_on) {
                const float w = __fmul_rn(__fsub_rn(1.0f, sa[k]), sT[k]);
                const float we = __fadd_rn(w, 1e-10f);
                const float lw = logf(we);
                const float* row = out + (base + k) * n_out + 4;
                const float x0 = lane < n_seg ? row[lane] : -INFINITY;
                const float x1 = lane + 32 < n_seg ? row[lane + 32] : -INFINITY;
                const float m = warp_max(fmaxf(x0, x1));
                const float e0 = expf(x0 - m), e1 = expf(x1 - m);
                const float se = warp_sum(e0 + e1);
                const float lse = m + logf(se);
                const float dv0 = lane < n_seg ? dS0 * expf(lw + (x0 - lse) - sc0) : 0.f;
                const float dv1 = lane + 32 < n_seg ? dS1 * expf(lw + (x1 - lse) - sc1) : 0.f;
                const float sum = warp_sum(dv0 + dv1);
                dx0 = dv0 - (e0 / se) * sum; dx1 = dv1 - (e1 / se) * sum;
                if (lane == 0) sdw[k] += sum / we;
            }
            if (4 + lane < ldo) g[lane] = lane < n_seg ? dx0 : 0.f;
            if (4 + lane + 32 < ldo) g[lane + 32] = lane + 32 < n_seg ? dx1 : 0.f;
        }
        __syncwarp();
        // ---- pass C: reverse suffix sums of d w_j w_j, density gradient ----
        float R = 0.f;
        for (int s0 = ((S - 1) / 32) * 32; s0 >= 0; s0 -= 32) {
            const int k = s0 + lane; const bool valid = k < S;
            float a = 1.f, T = 0.f, dw = 0.f, fac = 0.f;
            if (valid) { a = sa[k]; T = sT[k]; dw = sdw[k]; fac = sfac[k]; }
            const float x = dw * __fmul_rn(__fsub_rn(1.0f, a), T);
            float incl = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { float n = __shfl_down_sync(0xffffffffu, incl, o); if (lane + o < 32) incl += n; }
            float excl = __shfl_down_sync(0xffffffffu, incl, 1); if (lane == 31) excl = 0.f;
            if (valid) {
                const float da = -T * dw + (R + excl) / __fadd_rn(a, NERF_EPS_T);
                d_out[(base + k) * (int64_t)ldo] = fac == 0.f ? 0.f : da * fac;
            }
            R += __shfl_sync(0xffffffffu, incl, 0);
        }
        __syncwarp();
    }
}

extern "C" int nerf_head_composite_bwd(const float* out, int n_out, int n_seg, const float* points, const float* noise,
                                       float noise_std, int64_t B, int S, const float* seg_fwd, const float* d_pixels,
                                       const float* d_seg, float* d_out, int ldo, void* stream) {
    if (B == 0 && S > 0) return NERF_OK;
    if (!out || !points || !d_out || B < 0 || S <= 0 || n_seg < 0 || n_seg > 64 || n_out != 4 + n_seg || ldo < n_out || (ldo & 3))
        return NERF_ERR_BAD_ARG;
    if (d_seg && !seg_fwd) return NERF_ERR_BAD_ARG;
    if (B == 0) return NERF_OK;
    if (S > 3072) return NERF_ERR_UNSUPPORTED;                   // per-ray scratch must fit 48 KB
    int wpb = (int)(49152 / (16 * (int64_t)S)); if (wpb > 8) wpb = 8; if (wpb < 1) wpb = 1;
    int64_t blocks = (B + wpb - 1) / wpb, cap = (int64_t)nerf_num_sms() * 32;
    head_composite_bwd_kernel<<<(unsigned)(blocks < cap ? blocks : cap), wpb * 32, (size_t)wpb * 16 * S, (cudaStream_t)stream>>>(
        out, n_out, n_seg, points, noise, noise_std, B, S, seg_fwd, d_pixels, d_seg, d_out, ldo);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// ---- one trunk unit's ReLU + LayerNorm backward (unit = Linear -> LayerNorm -> ReLU, model.py:525-540) ----
//   y = relu(gamma n + beta):  da = [gamma n + beta > 0] dy;  d gamma += da n;  d beta += da
//   dz = rstd (dn - mean(dn) - n mean(dn n)),  dn = da gamma;  d bias += dz
// One warp per point row, block reduction of the column sums through shared atomics.
__global__ void __launch_bounds__(256) head_row_bwd_kernel(
        const float* __restrict__ G, const float* __restrict__ n, const float* __restrict__ rstd,
        const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ dz_out,
        float* __restrict__ d_gamma, float* __restrict__ d_beta, float* __restrict__ d_bias, int64_t P) {
    __shared__ float red[3 * HID];
    const int lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 3 * HID; i += blockDim.x) red[i] = 0.f;
    __syncthreads();
    float g[8], bt[8], s_dg[8], s_db[8], s_dbias[8];
    load_row8(gamma, lane, g); load_row8(beta, lane, bt);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s_dg[j] = 0.f; s_db[j] = 0.f; s_dbias[j] = 0.f; }
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t m = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; m < P; m += warps) {
        float nn[8], dy[8], dn[8], c1 = 0.f, c2 = 0.f;
        load_row8(n + m * HID, lane, nn); load_row8(G + m * HID, lane, dy);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float da = fmaf(g[j], nn[j], bt[j]) > 0.f ? dy[j] : 0.f;
            s_dg[j] = fmaf(da, nn[j], s_dg[j]); s_db[j] += da;
            dn[j] = da * g[j]; c1 += dn[j]; c2 = fmaf(dn[j], nn[j], c2);
        }
        c1 = warp_sum(c1) * (1.0f / HID); c2 = warp_sum(c2) * (1.0f / HID);
        const float rs = rstd[m];
        float dz[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { dz[j] = rs * (dn[j] - c1 - nn[j] * c2); s_dbias[j] += dz[j]; }
        store_row8(dz_out + m * HID, lane, dz);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int col = col_of(lane, j);
        atomicAdd(&red[col], s_dg[j]); atomicAdd(&red[HID + col], s_db[j]); atomicAdd(&red[2 * HID + col], s_dbias[j]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < HID; i += blockDim.x) {
        atomicAdd(d_gamma + i, red[i]); atomicAdd(d_beta + i, red[HID + i]); atomicAdd(d_bias + i, red[2 * HID + i]);
    }
}

// column sums of d_out [P, ldo] -> bias gradient of the output Linear (first n_out columns)
__global__ void __launch_bounds__(256) head_colsum_kernel(const float* __restrict__ d_out, int ldo, int n_out, int64_t P,
                                                          float* __restrict__ d_bias) {
    __shared__ float red[64];
    if (threadIdx.x < 64) red[threadIdx.x] = 0.f;
    __syncthreads();
    const int col = threadIdx.x & 63, rg = threadIdx.x >> 6;
    float s = 0.f;
    if (col < n_out)
        for (int64_t m = (int64_t)blockIdx.x * 4 + rg; m < P; m += (int64_t)gridDim.x * 4) s += d_out[m * ldo + col];
    atomicAdd(&red[col], s);
    __syncthreads();
    if (threadIdx.x < n_out) atomicAdd(d_bias + threadIdx.x, red[threadIdx.x]);
}

// Training workspace (floats): n[5][P*256] | mu[5][P] | rstd[5][P] | G[P*256] | dz[P*256]
struct HeadWs { float* n[5]; float* mu[5]; float* rstd[5]; float* G; float* dz; };
static int64_t head_carve(int64_t P, void* base, HeadWs* ws) {
    const int64_t act = (P * HID + 63) & ~(int64_t)63, vec = (P + 63) & ~(int64_t)63;
    float* b = (float*)base; int64_t o = 0;
    for (int l = 0; l < 5; ++l) { if (ws) ws->n[l] = b + o; o += act; }
    for (int l = 0; l < 5; ++l) { if (ws) ws->mu[l] = b + o; o += vec; }
    for (int l = 0; l < 5; ++l) { if (ws) ws->rstd[l] = b + o; o += vec; }
    if (ws) ws->G = b + o; o += act;
    if (ws) ws->dz = b + o; o += act;
    return o * (int64_t)sizeof(float);
}

extern "C" int64_t nerf_head_mlp_train_workspace_bytes(int64_t P) {
    if (P < 0) return NERF_ERR_BAD_ARG;
    return head_carve(P, nullptr, nullptr);
}

extern "C" int nerf_head_mlp_fwd_train(const float* const* params_host, const float* feat, int64_t P, int n_in, int n_out,
                                       float* out, void* workspace, int64_t workspace_bytes, void* stream) {
    if (!params_host || !feat || !out || P < 0 || n_in <= 0 || (n_in & 3) || n_out <= 0 || n_out > 64) return NERF_ERR_BAD_ARG;
    if (P == 0) return NERF_OK;
    HeadWs ws; if (!workspace || head_carve(P, workspace, &ws) > workspace_bytes) return NERF_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned tiles = (unsigned)((P + BM - 1) / BM);
    for (int l = 0; l < 5; ++l) {
        ASource s{};
        s.row_div2 = 1;
        if (l == 0) { s.seg2 = feat; s.K2 = n_in; }
        else { s.n_prev = ws.n[l - 1]; s.gamma = params_host[4 * (l - 1) + 2]; s.beta = params_host[4 * (l - 1) + 3]; s.K1 = HID; s.relu = 1; }
        mlp_tile_kernel<MODE_FWD_LN><<<tiles, 256, 0, st>>>(s, params_host[4 * l], l == 0 ? n_in : HID, params_host[4 * l + 1],
                                                           ws.n[l], ws.mu[l], ws.rstd[l], P, HID);
        NERF_CUDA_CHECK_LAUNCH();
    }
    ASource s{};
    s.row_div2 = 1; s.n_prev = ws.n[4]; s.gamma = params_host[18]; s.beta = params_host[19]; s.K1 = HID; s.relu = 1;
    mlp_tile_kernel<MODE_LINEAR><<<tiles, 256, 0, st>>>(s, params_host[20], HID, params_host[21], out, nullptr, nullptr, P, n_out);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int nerf_head_mlp_bwd(const float* const* params_host, const float* feat, int64_t P, int n_in, int n_out,
                                 const float* d_out, int ldo, float* const* grads_host, void* workspace,
                                 int64_t workspace_bytes, void* stream) {
    if (!params_host || !grads_host || !feat || !d_out || P < 0 || n_in <= 0 || (n_in & 3) || n_out <= 0 || n_out > 64 || ldo != 64)
        return NERF_ERR_BAD_ARG;
    if (P == 0) return NERF_OK;
    HeadWs ws; if (!workspace || head_carve(P, workspace, &ws) > workspace_bytes) return NERF_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned tiles = (unsigned)((P + BM - 1) / BM);
    const int64_t rows_per_split = 4096, splits = (P + rows_per_split - 1) / rows_per_split;
    int64_t rb = (P + 7) / 8, cap = (int64_t)nerf_num_sms() * 4;
    const unsigned rgrid = (unsigned)(rb < cap ? rb : cap);
    // output Linear: d b5, d W5 = d_out^T relu(gamma4 n4 + beta4), G = d_out W5
    {
        int64_t cb = (P + 3) / 4; if (cb > cap) cb = cap;
        head_colsum_kernel<<<(unsigned)cb, 256, 0, st>>>(d_out, ldo, n_out, P, grads_host[21]);
        NERF_CUDA_CHECK_LAUNCH();
        ASource src{}; src.row_div2 = 1; src.n_prev = ws.n[4]; src.gamma = params_host[18]; src.beta = params_host[19]; src.K1 = HID; src.relu = 1;
        dim3 wg(HID / 64, 1, (unsigned)splits);
        mlp_wgrad_kernel<<<wg, 256, 0, st>>>(d_out, src, HID, grads_host[20], P, rows_per_split, ldo, n_out);
        NERF_CUDA_CHECK_LAUNCH();
        ASource d{}; d.n_prev = d_out; d.K1 = ldo;
        mlp_tile_kernel<MODE_DGRAD><<<tiles, 256, 0, st>>>(d, params_host[20], HID, nullptr, ws.G, nullptr, nullptr, P, n_out);
        NERF_CUDA_CHECK_LAUNCH();
    }
    for (int l = 4; l >= 0; --l) {
        head_row_bwd_kernel<<<rgrid, 256, 0, st>>>(ws.G, ws.n[l], ws.rstd[l], params_host[4 * l + 2], params_host[4 * l + 3], ws.dz,
                                                  grads_host[4 * l + 2], grads_host[4 * l + 3], grads_host[4 * l + 1], P);
        NERF_CUDA_CHECK_LAUNCH();
        ASource src{}; src.row_div2 = 1;
        if (l == 0) { src.seg2 = feat; src.K2 = n_in; }
        else { src.n_prev = ws.n[l - 1]; src.gamma = params_host[4 * (l - 1) + 2]; src.beta = params_host[4 * (l - 1) + 3]; src.K1 = HID; src.relu = 1; }
        const int Kt = l == 0 ? n_in : HID;
        dim3 wg((unsigned)((Kt + 63) / 64), HID / 64, (unsigned)splits);
        mlp_wgrad_kernel<<<wg, 256, 0, st>>>(ws.dz, src, Kt, grads_host[4 * l], P, rows_per_split);
        NERF_CUDA_CHECK_LAUNCH();
        if (l > 0) {
            ASource d{}; d.n_prev = ws.dz; d.K1 = HID;
            mlp_tile_kernel<MODE_DGRAD><<<tiles, 256, 0, st>>>(d, params_host[4 * l], HID, nullptr, ws.G, nullptr, nullptr, P);
            NERF_CUDA_CHECK_LAUNCH();
        }
    }
    return NERF_OK;
}
