// Weight packing for the tcgen05 path: fp32 flat parameters (state_dict order) ->
//   * bf16 B-operand stages, 16 KB each = [128 out-features x 64 in-features], K-major,
//     128B-swizzled, in the exact order the MMA warp consumes them
//     (layer, k-slab, n-half), with the PREVIOUS unit's LayerNorm gamma folded into the
//     256 hidden columns;
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

__global__ void pack_stage_kernel(const float* __restrict__ params, PackLayout L, const float* __restrict__ vec,
                                  uint8_t* __restrict__ packed) {
    // one thread per (stage, row r in [0,128), 16-byte chunk c in [0,8))
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)L.n_stages * 128 * 8) return;
    int c = (int)(i & 7), r = (int)((i >> 3) & 127), stage = (int)(i >> 10);
    int l = 0;
    while (l < 9 && stage >= L.stage_begin[l + 1]) ++l;
    int rel = stage - L.stage_begin[l];
    int ks = rel >> 1, nh = rel & 1;
    int n = nh * 128 + r;
    const int Kt = L.Kt[l];
    const float* W = params + L.off_W[l];
    const float* gprev = (l > 0) ? params + L.off_gamma[l - 1] : nullptr;
    const int hid_slabs = (l == 0) ? 0 : 4;
    uint32_t out[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float v[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int kk = c * 8 + q * 2 + h;            // column inside the slab
            float w = 0.f;
            if (ks < hid_slabs) {
                int k = ks * 64 + kk;
                w = W[(size_t)n * Kt + k] * gprev[k];
            } else {
                int k2 = (ks - hid_slabs) * 64 + kk;      // index into the appended encoding columns
                int kcol = (l == 0 ? 0 : 256) + k2;
                if (kcol < Kt) w = W[(size_t)n * Kt + kcol];
                else if (ks == L.nk[l] - 1 && kk >= TC_ONES_COL && kk < TC_ONES_COL + 3) {   // the slab's ones columns
                    float p3[3]; split_bias3(vec[TC_VEC_BIAS + l * 256 + n], p3); w = p3[kk - TC_ONES_COL];
                }
            }
            v[h] = w;
        }
        out[q] = tc::pack_bf16x2(v[0], v[1]);
    }
    *reinterpret_cast<uint4*>(packed + (size_t)stage * TC_STAGE_BYTES + tc::sw128_offset(r, c)) =
        make_uint4(out[0], out[1], out[2], out[3]);
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

// one thread per (unit, output feature): the 16-byte k-core-1 row of the unit's bias block
__global__ void pack_bias_blocks_kernel(const float* __restrict__ vec, uint8_t* __restrict__ blocks) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 10 * 256) return;
    const int l = i >> 8, n = i & 255, nh = n >> 7, r = n & 127;
    float p3[3]; split_bias3(vec[TC_VEC_BIAS + i], p3);
    uint8_t* row = blocks + (size_t)(l * 2 + nh) * TC_BIASBLK_BYTES + (r >> 3) * TC_BIASBLK_SBO + (r & 7) * 16;
    *reinterpret_cast<uint4*>(row) = make_uint4(0u, 0u, 0u, 0u);                                        // k-core 0: slab columns 48..55
    *reinterpret_cast<uint4*>(row + TC_BIASBLK_LBO) = make_uint4(0u, 0u, tc::pack_bf16x2(p3[0], p3[1]), tc::pack_bf16x2(p3[2], 0.f));
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
    int64_t n = (int64_t)L.n_stages * 128 * 8;
    pack_stage_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(params, L, vec, (uint8_t*)packed);
    NERF_CUDA_CHECK_LAUNCH();
    pack_bias_blocks_kernel<<<10, 256, 0, st>>>(vec, (uint8_t*)packed + tc_packed_biasblk_offset(L));
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
