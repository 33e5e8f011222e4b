// Weight packing for the tcgen05 path: fp32 flat parameters (state_dict order) ->
//   * the bf16 B operand as k-cores ([128 out-features x 8 in-features], no swizzle, 2 KB; see
//     tc_layout.cuh) per (unit, n-half) in the order the MMA warp consumes them, with the PREVIOUS
//     unit's LayerNorm gamma folded into the 256 hidden columns and bias' as three bf16 parts in
//     the columns that meet the slabs' ones;
//   * fp32 epilogue vectors: bias' = b + W[:, :256] beta_prev per layer, and the two heads
//     with the affine of the unit they read folded in.
// (LN(x) = gamma n + beta, so W (gamma n + beta) + b = (W diag gamma) n + (W beta + b).)
#include "tc_layout.cuh"

// bias' -> three bf16 parts whose fp32 sum reproduces it to ~2^-24
__device__ __forceinline__ void split_bias3(float b, float (&p)[3]) {
    p[0] = __bfloat162float(__float2bfloat16_rn(b));
    p[1] = __bfloat162float(__float2bfloat16_rn(b - p[0]));
    p[2] = __bfloat162float(__float2bfloat16_rn(b - p[0] - p[1]));
}

// one thread per 16-byte row of a k-core: (unit, n-half, k-core, n) -> 8 consecutive k of output feature n
__global__ void pack_stage_kernel(const float* __restrict__ params, PackLayout L, const float* __restrict__ vec,
                                  uint8_t* __restrict__ packed) {
    const int l = blockIdx.y;
    const int hk = L.half_kcores[l];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * hk * 128) return;
    const int r = i & 127, kc = (i >> 7) % hk, nh = (i >> 7) / hk;
    const int n = nh * 128 + r;
    const int Kt = L.Kt[l];
    const float* W = params + L.off_W[l];
    const float* gprev = (l > 0) ? params + L.off_gamma[l - 1] : nullptr;
    const bool blk = tc_unit_has_bias_block(l);
    const int hid_cores = (l == 0) ? 0 : 32;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    int c = kc;
    if (blk && c < TF_BIAS_KCORES) {                      // bias block against slab columns 48..63
        if (c == 1) { float p3[3]; split_bias3(vec[TC_VEC_BIAS + l * 256 + n], p3); v[4] = p3[0]; v[5] = p3[1]; v[6] = p3[2]; }
    } else {
        if (blk) c -= TF_BIAS_KCORES;
        if (c < hid_cores) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const int k = c * 8 + j; v[j] = W[(size_t)n * Kt + k] * gprev[k]; }
        } else {
            const int nenc_cores = hk - hid_cores - (blk ? TF_BIAS_KCORES : 0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int k2 = (c - hid_cores) * 8 + j;   // index into the appended encoding columns
                const int kcol = (l == 0 ? 0 : 256) + k2;
                const int n_enc = (l == 8) ? L.ED : 2 * NERF_PE_X_OCTAVES * 3;       // position columns only (states: per-ray bias)
                if (k2 < n_enc) v[j] = W[(size_t)n * Kt + kcol];
                else if (k2 >= nenc_cores * 8 - 64 + TC_ONES_COL && k2 < nenc_cores * 8 - 64 + TC_ONES_COL + 3) {   // the last slab's ones columns
                    float p3[3]; split_bias3(vec[TC_VEC_BIAS + l * 256 + n], p3); v[j] = p3[k2 - (nenc_cores * 8 - 64 + TC_ONES_COL)];
                }
            }
        }
    }
    uint8_t* dst = packed + L.unit_off[l] + ((size_t)(nh * hk + kc)) * TF_KCORE_BYTES + (r >> 3) * TF_B_SBO + (r & 7) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(tc::pack_bf16x2(v[0], v[1]), tc::pack_bf16x2(v[2], v[3]),
                                                tc::pack_bf16x2(v[4], v[5]), tc::pack_bf16x2(v[6], v[7]));
}

__global__ void pack_vectors_kernel(const float* __restrict__ params, PackLayout L, float* __restrict__ vec) {
    // one warp per output scalar: 10*256 biases, then 4*256 head weights (elementwise), then 4 head biases
    const int lane = threadIdx.x & 31;
    int w = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (w < 10 * 256) {
        int l = w >> 8, n = w & 255;
        float s = 0.f;
        if (l > 0) {
            const float* W = params + L.off_W[l] + (size_t)n * L.Kt[l];
            const float* beta = params + L.off_beta[l - 1];
            for (int k = lane; k < 256; k += 32) s = fmaf(W[k], beta[k], s);
            s = warp_sum(s);
        }
        if (lane == 0) vec[TC_VEC_BIAS + w] = params[L.off_b[l] + n] + s;
    } else if (w < 10 * 256 + 4) {
        int c = w - 10 * 256;                      // 0 = density, 1..3 = colour rows
        const float* wh = (c == 0) ? params + L.off_head_w[0] : params + L.off_head_w[1] + (size_t)(c - 1) * 256;
        const int unit = (c == 0) ? 7 : 9;
        const float* g = params + L.off_gamma[unit];
        const float* beta = params + L.off_beta[unit];
        float s = 0.f;
        for (int k = lane; k < 256; k += 32) {
            vec[TC_VEC_HEADW + c * 256 + k] = wh[k] * g[k];
            s = fmaf(wh[k], beta[k], s);
        }
        s = warp_sum(s);
        if (lane == 0) vec[TC_VEC_HEADB + c] = ((c == 0) ? params[L.off_head_b[0]] : params[L.off_head_b[1] + c - 1]) + s;
    }
}

// fp32 copy of the state columns of units 0 and 4, transposed so that the per-ray bias GEMM reads it coalesced:
// sw[u][k][n] = W_u[n, (u ? 256 : 0) + 60 + k]
__global__ void pack_state_weights_kernel(const float* __restrict__ params, PackLayout L, float* __restrict__ sw) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * 256 * L.EXs) return;
    const int n = i & 255, k = (i >> 8) % L.EXs, u = i / (L.EXs * 256);
    const int l = u ? 4 : 0;
    sw[i] = params[L.off_W[l] + (size_t)n * L.Kt[l] + (u ? 256 : 0) + 2 * NERF_PE_X_OCTAVES * 3 + k];
}

// ---- fp32-grade image (fused_tc_x3.cu): the same folded weights, split hi = bf16(w'), lo = bf16(w' - hi) ----
// folded weight of unit l, output feature n, K index kk of the unit's slab sequence [hidden 256 | encoding 64]
__device__ __forceinline__ float folded_weight(const float* __restrict__ params, const PackLayout& L, const float* __restrict__ vec,
                                               int l, int n, int kk) {
    const int Kt = L.Kt[l];
    const float* W = params + L.off_W[l];
    const int hid = (l == 0) ? 0 : 256;
    if (kk < hid) return W[(size_t)n * Kt + kk] * params[L.off_gamma[l - 1] + kk];
    const int k2 = kk - hid;
    const int n_enc = (l == 8) ? L.ED : 2 * NERF_PE_X_OCTAVES * 3;
    if (k2 < n_enc) return W[(size_t)n * Kt + hid + k2];
    if (k2 >= TC_ONES_COL && k2 < TC_ONES_COL + 3) { float p3[3]; split_bias3(vec[TC_VEC_BIAS + l * 256 + n], p3); return p3[k2 - TC_ONES_COL]; }
    return 0.f;
}
// one thread per 16-byte row of a k-core: (unit, n-half, k-core of the half's sequence, n)
__global__ void pack_stage_x3_kernel(const float* __restrict__ params, PackLayout L, X3Layout X, const float* __restrict__ vec,
                                     uint8_t* __restrict__ packed) {
    const int l = blockIdx.y;
    const bool blk = tc_unit_has_bias_block(l);
    const int cores = (int)(X.half_bytes[l] / TF_KCORE_BYTES);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * cores * 128) return;
    const int r = i & 127, kc = (i >> 7) % cores, nh = (i >> 7) / cores;
    const int n = nh * 128 + r;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    bool lo_part = false;
    if (blk && kc < TF_BIAS_KCORES) {
        if (kc == 1) { float p3[3]; split_bias3(vec[TC_VEC_BIAS + l * 256 + n], p3); v[4] = p3[0]; v[5] = p3[1]; v[6] = p3[2]; }
    } else {
        const int c = kc - (blk ? TF_BIAS_KCORES : 0);              // group g = c / 8: [4 k-cores hi | 4 k-cores lo]
        const int g = c >> 3, kq = c & 3;
        lo_part = (c & 4) != 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = folded_weight(params, L, vec, l, n, g * 32 + kq * 8 + j);
    }
    if (lo_part) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = v[j] - __bfloat162float(__float2bfloat16_rn(v[j]));
    }
    uint8_t* dst = packed + X.unit_off[l] + (size_t)nh * X.half_bytes[l] + (size_t)kc * TF_KCORE_BYTES + (r >> 3) * TF_B_SBO + (r & 7) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(tc::pack_bf16x2(v[0], v[1]), tc::pack_bf16x2(v[2], v[3]),
                                                tc::pack_bf16x2(v[4], v[5]), tc::pack_bf16x2(v[6], v[7]));
}

extern "C" int64_t nerf_g1_packed_x3_bytes(void) {
    X3Layout X; make_x3_layout(X);
    return tc_packed_x3_total(X);
}

extern "C" int nerf_g1_pack_weights_x3(const float* params, void* packed, void* stream) {
    if (!params || !packed) return NERF_ERR_BAD_ARG;
    PackLayout L; int rc = make_pack_layout(3, 3, L); if (rc) return rc;
    X3Layout X; make_x3_layout(X);
    cudaStream_t st = (cudaStream_t)stream;
    float* vec = reinterpret_cast<float*>((uint8_t*)packed + X.weight_bytes);
    const int warps = 10 * 256 + 4;
    pack_vectors_kernel<<<(warps * 32 + 255) / 256, 256, 0, st>>>(params, L, vec);
    NERF_CUDA_CHECK_LAUNCH();
    int64_t max_cores = 0;
    for (int l = 0; l < 10; ++l) max_cores = X.half_bytes[l] / TF_KCORE_BYTES > max_cores ? X.half_bytes[l] / TF_KCORE_BYTES : max_cores;
    pack_stage_x3_kernel<<<dim3((unsigned)((2 * max_cores * 128 + 255) / 256), 10), 256, 0, st>>>(params, L, X, vec, (uint8_t*)packed);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int64_t nerf_g1_packed_bytes(int Dx, int Dd) {
    PackLayout L;
    if (make_pack_layout(Dx, Dd, L) != NERF_OK) return NERF_ERR_BAD_ARG;
    return tc_packed_total(L);
}

extern "C" int nerf_g1_pack_weights(const float* params, int Dx, int Dd, void* packed, void* stream) {
    if (!params || !packed) return NERF_ERR_BAD_ARG;
    PackLayout L;
    int rc = make_pack_layout(Dx, Dd, L); if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    float* vec = reinterpret_cast<float*>((uint8_t*)packed + tc_packed_vec_offset(L));
    int warps = 10 * 256 + 4;
    pack_vectors_kernel<<<(warps * 32 + 255) / 256, 256, 0, st>>>(params, L, vec);
    NERF_CUDA_CHECK_LAUNCH();
    int max_hk = 0;
    for (int l = 0; l < 10; ++l) max_hk = L.half_kcores[l] > max_hk ? L.half_kcores[l] : max_hk;
    pack_stage_kernel<<<dim3((2 * max_hk * 128 + 255) / 256, 10), 256, 0, st>>>(params, L, vec, (uint8_t*)packed);
    NERF_CUDA_CHECK_LAUNCH();
    if (L.EXs) {
        pack_state_weights_kernel<<<(2 * 256 * L.EXs + 255) / 256, 256, 0, st>>>(
            params, L, reinterpret_cast<float*>((uint8_t*)packed + tc_packed_state_w_offset(L)));
        NERF_CUDA_CHECK_LAUNCH();
    }
    return NERF_OK;
}
