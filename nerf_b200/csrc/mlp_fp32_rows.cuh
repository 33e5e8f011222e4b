// Row-wise pieces of the fp32 MLP: the two heads (density 256->1, colour 256->3)
// and the LayerNorm/ReLU backward that turns dL/dy of a unit into dL/dz, with all
// per-column reductions (d gamma, d beta, d bias, head weight grads).
// One warp per point row; lane owns columns lane*4..+3 and 128+lane*4..+3.
#pragma once
#include "common.cuh"
#define HID NERF_HIDDEN

__device__ __forceinline__ void load_row8(const float* __restrict__ p, int lane, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p + lane * 4), b = *reinterpret_cast<const float4*>(p + 128 + lane * 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void store_row8(float* __restrict__ p, int lane, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p + lane * 4) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 128 + lane * 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ int col_of(int lane, int j) { return (j < 4) ? lane * 4 + j : 128 + lane * 4 + (j - 4); }

// out[m][c] = sum_k (gamma_k n[m][k] + beta_k) Wh[c][k] + bh[c]
template <int NH>
__global__ void __launch_bounds__(256) head_fwd_kernel(
        const float* __restrict__ n, const float* __restrict__ gamma, const float* __restrict__ beta,
        const float* __restrict__ Wh, const float* __restrict__ bh, float* __restrict__ out, int64_t P) {
    const int lane = threadIdx.x & 31;
    float g[8], b[8], w[NH][8];
    load_row8(gamma, lane, g); load_row8(beta, lane, b);
#pragma unroll
    for (int c = 0; c < NH; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) w[c][j] = Wh[c * HID + col_of(lane, j)];   // scalar: color.weight is not 16B aligned
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t m = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; m < P; m += warps) {
        float x[8]; load_row8(n + m * HID, lane, x);
        float s[NH];
#pragma unroll
        for (int c = 0; c < NH; ++c) s[c] = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float y = fmaf(g[j], x[j], b[j]);
#pragma unroll
            for (int c = 0; c < NH; ++c) s[c] = fmaf(y, w[c][j], s[c]);
        }
#pragma unroll
        for (int c = 0; c < NH; ++c) { s[c] = warp_sum(s[c]); }
        if (lane == 0) {
#pragma unroll
            for (int c = 0; c < NH; ++c) out[m * NH + c] = s[c] + bh[c];
        }
    }
}

struct RowBwdArgs {
    const float* G;        // [P,256] dL/dy from the next unit's dgrad, or null
    const float* dh;       // [P,NH] head gradient, or null
    const float* Wh;       // [NH,256]
    const float* n;        // [P,256] this unit's normalised output
    const float* mu; const float* rstd;   // [P]
    const float* gamma;    // this unit's LayerNorm gamma
    const float* beta;
    float* dz;             // [P,256] out
    float* d_gamma; float* d_beta; float* d_bias;   // [256] accumulated
    float* d_Wh; float* d_bh;                        // [NH,256], [NH] accumulated (if dh)
    int64_t P;
};

template <int NH>   // NH = 0: no head on this unit
__global__ void __launch_bounds__(256) row_bwd_kernel(RowBwdArgs a) {
    constexpr int NHS = NH > 0 ? NH : 1;
    __shared__ float red[(3 + NHS) * HID + 4];
    const int lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (3 + NHS) * HID + 4; i += blockDim.x) red[i] = 0.f;
    __syncthreads();
    float g[8], bt[8], wh[NHS][8];
    load_row8(a.gamma, lane, g); load_row8(a.beta, lane, bt);
#pragma unroll
    for (int c = 0; c < NHS; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) wh[c][j] = (NH > 0) ? a.Wh[c * HID + col_of(lane, j)] : 0.f;
    float s_dg[8], s_db[8], s_dbias[8], s_dwh[NHS][8], s_dbh[NHS];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s_dg[j] = 0.f; s_db[j] = 0.f; s_dbias[j] = 0.f; }
#pragma unroll
    for (int c = 0; c < NHS; ++c) { s_dbh[c] = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s_dwh[c][j] = 0.f; }

    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t m = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; m < a.P; m += warps) {
        float nn[8], dy[8];
        load_row8(a.n + m * HID, lane, nn);
        if (a.G) load_row8(a.G + m * HID, lane, dy);
        else {
#pragma unroll
            for (int j = 0; j < 8; ++j) dy[j] = 0.f;
        }
        if (NH > 0) {
            float dh[NHS];
#pragma unroll
            for (int c = 0; c < NHS; ++c) dh[c] = a.dh[m * NH + c];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float y = fmaf(g[j], nn[j], bt[j]);
#pragma unroll
                for (int c = 0; c < NHS; ++c) { dy[j] = fmaf(dh[c], wh[c][j], dy[j]); s_dwh[c][j] = fmaf(dh[c], y, s_dwh[c][j]); }
            }
            if (lane == 0) {
#pragma unroll
                for (int c = 0; c < NHS; ++c) s_dbh[c] += dh[c];
            }
        }
        const float mu = a.mu[m], rstd = a.rstd[m];
        const float n0 = __fmul_rn(__fsub_rn(0.f, mu), rstd);     // value a zeroed (ReLU-clamped) element normalises to
        float dn[8], c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            s_dg[j] = fmaf(dy[j], nn[j], s_dg[j]); s_db[j] += dy[j];
            dn[j] = dy[j] * g[j]; c1 += dn[j]; c2 = fmaf(dn[j], nn[j], c2);
        }
        c1 = warp_sum(c1) * (1.0f / HID); c2 = warp_sum(c2) * (1.0f / HID);
        float dz[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float da = rstd * (dn[j] - c1 - nn[j] * c2);
            dz[j] = (nn[j] > n0) ? da : 0.f;
            s_dbias[j] += dz[j];
        }
        store_row8(a.dz + m * HID, lane, dz);
    }
    // block reduction through shared atomics (conflict free across lanes), then one global atomic per column
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        int col = col_of(lane, j);
        atomicAdd(&red[col], s_dg[j]); atomicAdd(&red[HID + col], s_db[j]); atomicAdd(&red[2 * HID + col], s_dbias[j]);
        if (NH > 0) {
#pragma unroll
            for (int c = 0; c < NHS; ++c) atomicAdd(&red[(3 + c) * HID + col], s_dwh[c][j]);
        }
    }
    if (NH > 0 && lane == 0) {
#pragma unroll
        for (int c = 0; c < NHS; ++c) atomicAdd(&red[(3 + NHS) * HID + c], s_dbh[c]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < HID; i += blockDim.x) {
        atomicAdd(a.d_gamma + i, red[i]); atomicAdd(a.d_beta + i, red[HID + i]); atomicAdd(a.d_bias + i, red[2 * HID + i]);
        if (NH > 0) {
#pragma unroll
            for (int c = 0; c < NHS; ++c) atomicAdd(a.d_Wh + c * HID + i, red[(3 + c) * HID + i]);
        }
    }
    if (NH > 0 && threadIdx.x < NH) atomicAdd(a.d_bh + threadIdx.x, red[(3 + NHS) * HID + threadIdx.x]);
}
