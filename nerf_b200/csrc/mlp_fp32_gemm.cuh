// fp32 FFMA tiles for the G1 coordinate MLP ("fp32 mode": the <=1e-3 parity path).
// One kernel template does  C[64 x 256] = A[64 x K] * B  per CTA with
//   MODE_FWD  : A = [gamma*n_prev+beta | seg2], B = W^T (W is [256][Kt] row-major),
//               epilogue = +bias, ReLU, LayerNorm statistics, normalise  (the whole
//               Linear->ReLU->LayerNorm unit of SURVEY A.3 without its affine, which
//               the NEXT consumer applies when it loads the row)
//   MODE_DGRAD: A = dz [P, lda = src.K1], B = W[:n_out, :256] (natural rows; n_out = rows of W = width of dz,
//               256 for a trunk unit, 54 for HEAD's output Linear), epilogue = plain store.
//   MODE_FWD_LN: HEAD trunk unit Linear->LayerNorm->ReLU (nerf/model.py:525-542): epilogue = +bias,
//               LayerNorm statistics, normalise; the consumer applies the affine AND the ReLU on load
//               (ASource.relu).   MODE_LINEAR: plain Linear with n_out <= 256 outputs (HEAD's 256->54).
#pragma once
#include "common.cuh"

#define HID NERF_HIDDEN
#define BM 64
#define BK 16
#define A_STRIDE 68      // 64 + 4: keeps float4 alignment, <=2-way conflicts on the transposing store

struct ASource {
    const float* n_prev;     // [P,256] normalised activations of the previous unit (or null: K1 = 0)
    const float* gamma;      // previous LayerNorm affine (applied on load)
    const float* beta;
    const float* seg2;       // extra input columns (PE of position: one row per point; PE of dir: one row per ray)
    int K1, K2, row_div2;
    int relu;                // HEAD: the producing unit is Linear->LayerNorm->ReLU, so relu(gamma*n+beta)
};

__device__ __forceinline__ float4 load_a4(const ASource& s, int64_t row, int k, int64_t P) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row >= P) return v;
    if (k < s.K1) {
        float4 n = *reinterpret_cast<const float4*>(s.n_prev + row * HID + k);
        float4 g = *reinterpret_cast<const float4*>(s.gamma + k);
        float4 b = *reinterpret_cast<const float4*>(s.beta + k);
        v.x = fmaf(g.x, n.x, b.x); v.y = fmaf(g.y, n.y, b.y); v.z = fmaf(g.z, n.z, b.z); v.w = fmaf(g.w, n.w, b.w);
        if (s.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    } else if (k - s.K1 < s.K2) {
        v = *reinterpret_cast<const float4*>(s.seg2 + (row / s.row_div2) * s.K2 + (k - s.K1));
    }
    return v;
}

enum { MODE_FWD = 0, MODE_DGRAD = 1, MODE_FWD_LN = 2, MODE_LINEAR = 3 };

template <int MODE>
__global__ void __launch_bounds__(256, 2) mlp_tile_kernel(
        ASource src,                       // FWD: inputs; DGRAD: n_prev = dz, K1 = 256, gamma = null
        const float* __restrict__ W, int Kt,        // weight [256][Kt]
        const float* __restrict__ bias,             // FWD
        float* __restrict__ out,                    // FWD: n_out [P,256]; DGRAD: G [P,256]
        float* __restrict__ mu_out, float* __restrict__ rstd_out, int64_t P, int n_out = HID) {
    constexpr bool IS_FWD = MODE != MODE_DGRAD;               // A from ASource, B = W^T
    __shared__ __align__(16) float As[2][BK][A_STRIDE];
    __shared__ __align__(16) float Bs[2][BK][HID];
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int64_t row0 = (int64_t)blockIdx.x * BM;
    const int K = IS_FWD ? (src.K1 + src.K2) : n_out;
    const int nchunks = (K + BK - 1) / BK;

    // accumulators as packed pairs: the inner product issues fma.rn.f32x2 (FFMA2, two IEEE fp32 FMAs per
    // instruction -- same results as scalar FFMA, half the issue slots / register-file reads)
    float2 acc2[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc2[i][j] = make_float2(0.f, 0.f);

    float4 ra; float4 rb[4];
    auto gload = [&](int chunk) {
        const int k0 = chunk * BK;
        {   // A tile: 64 rows x 16 k, one float4 per thread
            const int r = tid >> 2, kq = (tid & 3) * 4;
            if (IS_FWD) ra = load_a4(src, row0 + r, k0 + kq, P);
            else ra = (row0 + r < P && k0 + kq < src.K1) ? *reinterpret_cast<const float4*>(src.n_prev + (row0 + r) * src.K1 + k0 + kq)
                                     : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (IS_FWD) {     // B[k][n] = W[n][k]: thread = output feature n, 16 consecutive k
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                int k = k0 + q * 4;
                rb[q] = (k < Kt && tid < n_out) ? *reinterpret_cast<const float4*>(W + (size_t)tid * Kt + k) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        } else {                    // B[kk][j] = W[n0+kk][j], j < 256: natural rows
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                int idx = tid + q * 256; int kk = idx >> 6, j4 = (idx & 63) * 4;
                rb[q] = (k0 + kk < n_out) ? *reinterpret_cast<const float4*>(W + (size_t)(k0 + kk) * Kt + j4) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    };
    auto sstore = [&](int buf) {
        {
            const int r = tid >> 2, kq = (tid & 3) * 4;
            As[buf][kq + 0][r] = ra.x; As[buf][kq + 1][r] = ra.y; As[buf][kq + 2][r] = ra.z; As[buf][kq + 3][r] = ra.w;
        }
        if (IS_FWD) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                Bs[buf][q * 4 + 0][tid] = rb[q].x; Bs[buf][q * 4 + 1][tid] = rb[q].y;
                Bs[buf][q * 4 + 2][tid] = rb[q].z; Bs[buf][q * 4 + 3][tid] = rb[q].w;
            }
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                int idx = tid + q * 256; int kk = idx >> 6, j4 = (idx & 63) * 4;
                *reinterpret_cast<float4*>(&Bs[buf][kk][j4]) = rb[q];
            }
        }
    };

    gload(0); sstore(0); __syncthreads();
    for (int c = 0; c < nchunks; ++c) {
        const int buf = c & 1;
        if (c + 1 < nchunks) gload(c + 1);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8]);
            float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8 + 4]);
            float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
            float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][k][128 + tx * 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float2 b[4] = {make_float2(b0.x, b0.y), make_float2(b0.z, b0.w), make_float2(b1.x, b1.y), make_float2(b1.z, b1.w)};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float2 aa = make_float2(a[i], a[i]);
#pragma unroll
                for (int j = 0; j < 4; ++j) acc2[i][j] = __ffma2_rn(aa, b[j], acc2[i][j]);
            }
        }
        if (c + 1 < nchunks) { sstore(buf ^ 1); }
        __syncthreads();
    }

    // ---- epilogue: thread owns rows ty*8..+7, columns tx*4..+3 and 128+tx*4..+3 ----
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc[i][2 * j] = acc2[i][j].x; acc[i][2 * j + 1] = acc2[i][j].y; }
    if (MODE == MODE_LINEAR) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int64_t row = row0 + ty * 8 + i;
            if (row >= P) continue;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int col = (j < 4 ? 0 : 128) + tx * 4 + (j & 3);
                if (col < n_out) out[row * n_out + col] = acc[i][j] + bias[col];
            }
        }
    } else if (MODE == MODE_FWD || MODE == MODE_FWD_LN) {
        float bv[8];
        { float4 b0 = *reinterpret_cast<const float4*>(bias + tx * 4), b1 = *reinterpret_cast<const float4*>(bias + 128 + tx * 4);
          bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w; bv[4] = b1.x; bv[5] = b1.y; bv[6] = b1.z; bv[7] = b1.w; }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) { float a = acc[i][j] + bv[j]; if (MODE == MODE_FWD) a = fmaxf(a, 0.f); acc[i][j] = a; s += a; }
            const float mean = warp_sum(s) * (1.0f / HID);
            float q = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) { float d = __fsub_rn(acc[i][j], mean); q = fmaf(d, d, q); }
            const float var = warp_sum(q) * (1.0f / HID);
            const float rstd = 1.0f / sqrtf(var + NERF_LN_EPS);
            const int64_t row = row0 + ty * 8 + i;
            if (row < P) {
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = __fmul_rn(__fsub_rn(acc[i][j], mean), rstd);
                *reinterpret_cast<float4*>(out + row * HID + tx * 4) = make_float4(o[0], o[1], o[2], o[3]);
                *reinterpret_cast<float4*>(out + row * HID + 128 + tx * 4) = make_float4(o[4], o[5], o[6], o[7]);
                if (tx == 0 && mu_out) { mu_out[row] = mean; rstd_out[row] = rstd; }
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int64_t row = row0 + ty * 8 + i;
            if (row < P) {
                *reinterpret_cast<float4*>(out + row * HID + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
                *reinterpret_cast<float4*>(out + row * HID + 128 + tx * 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
            }
        }
    }
}

// ---- weight gradient: dW[n][k] += sum_m dz[m][n] * x[m][k],  x = [gamma*n_prev+beta | seg2] ----
// 64(n) x 64(k) tile per CTA, rows [m_begin, m_end) reduced in chunks of 16, atomics at the end.
// dz is [P, ldz] and only its first n_rows columns are weight rows (HEAD's output Linear: ldz 64, 54 rows).
static __global__ void __launch_bounds__(256) mlp_wgrad_kernel(
        const float* __restrict__ dz, ASource src, int Kt, float* __restrict__ dW,
        int64_t P, int64_t rows_per_split, int ldz = HID, int n_rows = HID) {
    __shared__ __align__(16) float As[2][BK][64];
    __shared__ __align__(16) float Bs[2][BK][64];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int k0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
    const int64_t m_begin = (int64_t)blockIdx.z * rows_per_split;
    int64_t m_end = m_begin + rows_per_split; if (m_end > P) m_end = P;
    if (m_begin >= m_end) return;
    float2 acc2[4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) acc2[i][j] = make_float2(0.f, 0.f);
    const int lr = tid >> 4, lc = (tid & 15) * 4;     // loader: row within chunk, column group
    float4 ra, rb;
    auto gload = [&](int64_t m0) {
        int64_t m = m0 + lr;
        ra = (m < m_end) ? *reinterpret_cast<const float4*>(dz + m * ldz + n0 + lc) : make_float4(0.f, 0.f, 0.f, 0.f);
        rb = (m < m_end) ? load_a4(src, m, k0 + lc, P) : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    auto sstore = [&](int buf) {
        *reinterpret_cast<float4*>(&As[buf][lr][lc]) = ra;
        *reinterpret_cast<float4*>(&Bs[buf][lr][lc]) = rb;
    };
    gload(m_begin); sstore(0); __syncthreads();
    int buf = 0;
    for (int64_t m0 = m_begin; m0 < m_end; m0 += BK) {
        if (m0 + BK < m_end) gload(m0 + BK);
#pragma unroll
        for (int mm = 0; mm < BK; ++mm) {
            float4 a = *reinterpret_cast<const float4*>(&As[buf][mm][ty * 4]);
            float4 b = *reinterpret_cast<const float4*>(&Bs[buf][mm][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
            const float2 bv[2] = {make_float2(b.x, b.y), make_float2(b.z, b.w)};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 aa = make_float2(av[i], av[i]);
#pragma unroll
                for (int j = 0; j < 2; ++j) acc2[i][j] = __ffma2_rn(aa, bv[j], acc2[i][j]);
            }
        }
        if (m0 + BK < m_end) sstore(buf ^ 1);
        __syncthreads();
        buf ^= 1;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int n = n0 + ty * 4 + i, k = k0 + tx * 4 + j;
            const float v = (j & 1) ? acc2[i][j >> 1].y : acc2[i][j >> 1].x;
            if (k < Kt && n < n_rows) atomicAdd(dW + (size_t)n * Kt + k, v);
        }
}
