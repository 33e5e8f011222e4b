// HEAD generation (the mip-NeRF style network of nerf/model.py in the reference tree) on tcgen05:
// integrated positional encoding + the Linear->LayerNorm->ReLU trunk + the 1+3+50 output head in ONE
// persistent kernel, bf16 operands, fp32 accumulation in TMEM.  Same skeleton as the G1 forward
// (fused_tc.cu): CTA pairs (cta_group::2, M = 256), two 128-interval tile slots per CTA ping-ponging
// between the tensor pipe and the epilogue warps, weights streamed as unswizzled k-cores through a TMA
// ring, activations never leaving the SM.  Differences that shape the kernel:
//   * LayerNorm comes BEFORE the ReLU and keeps its affine (it cannot be folded into the next Linear
//     through the ReLU): pass 2 of the epilogue is relu(z * (rstd gamma_j) + (beta_j - mean rstd gamma_j));
//   * unit 0 reads the 96 IPE features (K padded to 128) straight out of the slot's activation buffer,
//     which is free at that point -- no separate encoding slab;
//   * every unit's bias rides in the accumulator through a constant [128 x 16] "ones" A block and a
//     4 KB bias block (bias as hi+mid+lo bf16 parts) in front of the unit's first weight stage;
//   * the output head is the sixth unit with N = 64 (k-cores of 32 rows per CTA, 512 B), its epilogue
//     writes rows of n_out floats.
// SURVEY 8(a) rows a5-H, a6-H, a7 (nerf/model.py:24-163, 525-594).
#include "tc_layout.cuh"
#include <stdlib.h>

#define HT_THREADS 320
#define HT_RING 3
#define HT_UNITS 6
#define HT_SM_ACT 0                                       // 2 x 64 KB
#define HT_SM_RING (2 * 65536)                            // HT_RING x 20 KB
#define HT_SM_ONES (HT_SM_RING + HT_RING * TF_SLOT_BYTES) // [128 x 16] bf16 no-swizzle K-major: (1,1,1,0,...) per row
#define HT_SM_VEC (HT_SM_ONES + 4096)                     // gamma[5][256], beta[5][256] fp32
#define HT_SM_BAR (HT_SM_VEC + 10 * 256 * 4)
#define HT_SM_TOTAL (HT_SM_BAR + 256)
#define HT_SMEM_BYTES (HT_SM_TOTAL + 1024)
#define HT_ONES_LBO 128
#define HT_ONES_SBO 256

#define HT_OUT_N 64                                        // MMA N of the output head (n_out <= 64)
struct HeadTcLayout { int nk[HT_UNITS]; int half_kcores[HT_UNITS]; int kcore_bytes[HT_UNITS]; int64_t unit_off[HT_UNITS]; int64_t weight_bytes, vec_off, total; };
static inline void make_head_tc_layout(HeadTcLayout& L) {
    int64_t o = 0;
    for (int l = 0; l < HT_UNITS; ++l) {
        L.nk[l] = l == 0 ? 2 : 4;
        L.half_kcores[l] = TF_BIAS_KCORES + 8 * L.nk[l];
        L.kcore_bytes[l] = (l < 5 ? 128 : HT_OUT_N / 2) * 16;       // rows per CTA half x 16 B
        L.unit_off[l] = o; o += 2ll * L.half_kcores[l] * L.kcore_bytes[l];
    }
    L.weight_bytes = o; L.vec_off = o; L.total = o + 10 * 256 * 4;
}

struct HeadTcArgs {
    const uint8_t* stages; const float* vec;
    int nk[HT_UNITS]; int half_kcores[HT_UNITS]; int kcore_bytes[HT_UNITS]; int64_t unit_off[HT_UNITS];
    const float* rays_o; const float* rays_d; const float* t;     // [B,3], [B,3], fenceposts [B,S]
    int64_t B; int S; int64_t P; int n_tiles;                     // P = B (S-1) intervals
    float base_radius; int min_deg;
    float* out; int n_out; float* mean_out;                       // [P,n_out], [P,3]
    int log_softmax_seg;                                          // segmentation columns leave as log_softmax
    uint32_t wait_hint;
};

// ------------------------------------------------------------------ weight image ----
__device__ __forceinline__ void ht_split3(float b, float (&p)[3]) {
    p[0] = __bfloat162float(__float2bfloat16_rn(b));
    p[1] = __bfloat162float(__float2bfloat16_rn(b - p[0]));
    p[2] = __bfloat162float(__float2bfloat16_rn(b - p[0] - p[1]));
}

struct HeadParams { const float* p[22]; };                        // state_dict order: (W,b,gamma,beta) x 5, W5, b5

// one thread per 16-byte row of a k-core: (unit, n-half, k-core, n)
__global__ void head_pack_kernel(HeadParams hp, HeadTcLayout L, int n_in, int n_out, uint8_t* __restrict__ packed) {
    const int l = blockIdx.y;
    const int hk = L.half_kcores[l], rows = L.kcore_bytes[l] / 16;     // output features per CTA half
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * hk * rows) return;
    const int r = i % rows, kc = (i / rows) % hk, nh = (i / rows) / hk;
    const int n = nh * rows + r;
    const float* W = hp.p[l < 5 ? 4 * l : 20]; const float* bias = hp.p[l < 5 ? 4 * l + 1 : 21];
    const int K = l == 0 ? n_in : 256, N = l < 5 ? 256 : n_out;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    if (n < N) {
        if (kc == 0) { float p3[3]; ht_split3(bias[n], p3); v[0] = p3[0]; v[1] = p3[1]; v[2] = p3[2]; }   // x ones block columns 0..2
        else if (kc >= TF_BIAS_KCORES) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const int k = (kc - TF_BIAS_KCORES) * 8 + j; if (k < K) v[j] = W[(size_t)n * K + k]; }
        }
    }
    uint8_t* dst = packed + L.unit_off[l] + ((size_t)(nh * hk + kc)) * L.kcore_bytes[l] + (r >> 3) * TF_B_SBO + (r & 7) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(tc::pack_bf16x2(v[0], v[1]), tc::pack_bf16x2(v[2], v[3]),
                                                tc::pack_bf16x2(v[4], v[5]), tc::pack_bf16x2(v[6], v[7]));
}
__global__ void head_pack_vec_kernel(HeadParams hp, float* __restrict__ vec) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 5 * 256) return;
    const int l = i >> 8, n = i & 255;
    vec[l * 512 + n] = hp.p[4 * l + 2][n];
    vec[l * 512 + 256 + n] = hp.p[4 * l + 3][n];
}

extern "C" int64_t nerf_head_packed_bytes(void) { HeadTcLayout L; make_head_tc_layout(L); return L.total; }

extern "C" int nerf_head_pack_weights(const float* const* params_host, int n_in, int n_out, void* packed, void* stream) {
    if (!params_host || !packed || n_in <= 0 || n_in > 128 || n_out <= 0 || n_out > HT_OUT_N) return NERF_ERR_BAD_ARG;
    HeadTcLayout L; make_head_tc_layout(L);
    HeadParams hp; for (int i = 0; i < 22; ++i) { if (!params_host[i]) return NERF_ERR_BAD_ARG; hp.p[i] = params_host[i]; }
    cudaStream_t st = (cudaStream_t)stream;
    int max_hk = 0; for (int l = 0; l < HT_UNITS; ++l) max_hk = L.half_kcores[l] > max_hk ? L.half_kcores[l] : max_hk;
    head_pack_kernel<<<dim3((2 * max_hk * 128 + 255) / 256, HT_UNITS), 256, 0, st>>>(hp, L, n_in, n_out, (uint8_t*)packed);
    NERF_CUDA_CHECK_LAUNCH();
    head_pack_vec_kernel<<<5, 256, 0, st>>>(hp, reinterpret_cast<float*>((uint8_t*)packed + L.vec_off));
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// ------------------------------------------------------------------ device pieces ----
__device__ __forceinline__ float2 hf2(uint32_t a, uint32_t b) { return make_float2(__uint_as_float(a), __uint_as_float(b)); }
__device__ __forceinline__ void ht_store_chunk(uint32_t slab_addr, uint32_t row, uint32_t c16, uint4 v) {
    const uint32_t a = slab_addr + tc::sw128_offset(row, c16);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// sin and cos of x (|x| < 2^17) with a three-term Cody-Waite reduction, then the MUFU approximations on
// the reduced argument (|r| <= pi): features are rounded to bf16, this is ~1e-6 absolute.
__device__ __forceinline__ void ht_sincos(float x, float& s, float& c) {
    const float k = rintf(x * 0.15915494309189535f);
    float r = fmaf(k, -6.28125f, x);                       // 2 pi = 6.28125 + 1.9353071795864769e-3 (hi part: 9 significant bits)
    r = fmaf(k, -1.93500518798828125e-3f, r);
    r = fmaf(k, -3.0199159819e-7f, r);
    s = __sinf(r); c = __cosf(r);
}

// integrated positional encoding of one interval -> the row's 128 K-columns (96 features, 32 zeros)
// of the slot's activation slabs 0 and 1, and the Gaussian mean for the compositing kernel.
__device__ __forceinline__ void ht_encode_row(const HeadTcArgs& a, int tile, uint32_t row, uint32_t act_addr, int64_t& p_out, bool& valid_out) {
    const int64_t p = (int64_t)tile * 128 + row;
    const bool valid = p < a.P;
    p_out = p; valid_out = valid;
    float mean[3] = {0.f, 0.f, 0.f}, cov[3] = {0.f, 0.f, 0.f};
    if (valid) {
        const int64_t ray = p / (a.S - 1); const int k = (int)(p - ray * (a.S - 1));
        const float t0 = a.t[ray * a.S + k], t1 = a.t[ray * a.S + k + 1];
        const float mu = 0.5f * (t0 + t1), hw = 0.5f * (t1 - t0);
        const float mu2 = mu * mu, hw2 = hw * hw, hw4 = hw2 * hw2;
        const float den = 3.0f * mu2 + hw2;
        const float t_mean = mu + (2.0f * mu * hw2) / den;
        const float t_var = hw2 / 3.0f - (float)(4.0 / 15.0) * ((hw4 * (12.0f * mu2 - hw2)) / (den * den));
        const float r_var = a.base_radius * a.base_radius * (mu2 / 4.0f + (float)(5.0 / 12.0) * hw2 - ((float)(4.0 / 15.0) * hw4) / den);
        float d[3], o[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) { d[c] = a.rays_d[ray * 3 + c]; o[c] = a.rays_o[ray * 3 + c]; }
        const float mag = fmaxf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2], 1e-10f);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float dsq = d[c] * d[c];
            mean[c] = fmaf(d[c], t_mean, o[c]);
            cov[c] = t_var * dsq + r_var * (1.0f - dsq / mag);
            a.mean_out[p * 3 + c] = mean[c];
        }
    }
    // feature index deg*3+c (sin block, 48 columns) then 48 + deg*3+c (shifted-sin = cos block); packed two per register
    uint32_t pk[48];
    float prev_s = 0.f, prev_c = 0.f;
    float scale = exp2f((float)a.min_deg);
#pragma unroll
    for (int deg = 0; deg < 16; ++deg) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const int idx = deg * 3 + c;
            float s, co; ht_sincos(mean[c] * scale, s, co);
            const float amp = exp2f(-0.72134752044f * cov[c] * scale * scale);      // exp(-cov 4^deg / 2)
            s *= amp; co *= amp;
            if (idx & 1) { pk[idx >> 1] = tc::pack_bf16x2(prev_s, s); pk[24 + (idx >> 1)] = tc::pack_bf16x2(prev_c, co); }
            else { prev_s = s; prev_c = co; }
        }
        scale = scale + scale;
    }
#pragma unroll
    for (int c16 = 0; c16 < 8; ++c16)                                  // columns 0..63
        ht_store_chunk(act_addr, row, c16, make_uint4(pk[c16 * 4], pk[c16 * 4 + 1], pk[c16 * 4 + 2], pk[c16 * 4 + 3]));
#pragma unroll
    for (int c16 = 0; c16 < 8; ++c16) {                                // columns 64..127 (96.. are zero)
        const uint4 v = c16 < 4 ? make_uint4(pk[32 + c16 * 4], pk[33 + c16 * 4], pk[34 + c16 * 4], pk[35 + c16 * 4]) : make_uint4(0u, 0u, 0u, 0u);
        ht_store_chunk(act_addr + TC_SLAB_BYTES, row, c16, v);
    }
}

// LayerNorm (with affine) -> ReLU -> bf16 row of the next unit's A operand.
__device__ __noinline__ void ht_epilogue_ln(uint32_t taddr, uint32_t act_addr, uint32_t row, uint32_t gamma_addr) {
    uint32_t r0[32], r1[32];
    float2 s2 = make_float2(0.f, 0.f), q2 = make_float2(0.f, 0.f), s3 = s2, q3 = q2;
    auto pass1 = [&](const uint32_t (&rc)[32]) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float2 z0 = hf2(rc[j], rc[j + 1]), z1 = hf2(rc[j + 2], rc[j + 3]);
            s2 = __fadd2_rn(s2, z0); s3 = __fadd2_rn(s3, z1);
            q2 = __ffma2_rn(z0, z0, q2); q3 = __ffma2_rn(z1, z1, q3);
        }
    };
    tc::tmem_ld32(taddr, r0); tc::tmem_ld_wait();
#pragma unroll 1
    for (int c2 = 0; c2 < 4; ++c2) {
        tc::tmem_ld32(taddr + (2 * c2 + 1) * 32, r1);
        pass1(r0);
        tc::tmem_ld_wait();
        if (c2 < 3) tc::tmem_ld32(taddr + (2 * c2 + 2) * 32, r0);
        pass1(r1);
        tc::tmem_ld_wait();
    }
    s2 = __fadd2_rn(s2, s3); q2 = __fadd2_rn(q2, q3);
    const float mean = (s2.x + s2.y) * (1.0f / 256.f);
    const float var = fmaxf((q2.x + q2.y) * (1.0f / 256.f) - mean * mean, 0.f);
    const float rstd = rsqrtf(var + NERF_LN_EPS);
    const float2 rstd2 = make_float2(rstd, rstd), nmean2 = make_float2(-mean, -mean);
    const __nv_bfloat162 zero_bf = __floats2bfloat162_rn(0.f, 0.f);
    auto pass2 = [&](const uint32_t (&rc)[32], int ch) {
        uint32_t packed[16];
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float4 g = tc::lds128(gamma_addr + (ch * 32 + j) * 4), b = tc::lds128(gamma_addr + (256 + ch * 32 + j) * 4);
            const float2 a0 = __fmul2_rn(make_float2(g.x, g.y), rstd2), a1 = __fmul2_rn(make_float2(g.z, g.w), rstd2);
            const float2 c0 = __ffma2_rn(a0, nmean2, make_float2(b.x, b.y)), c1 = __ffma2_rn(a1, nmean2, make_float2(b.z, b.w));
            const float2 y0 = __ffma2_rn(hf2(rc[j], rc[j + 1]), a0, c0), y1 = __ffma2_rn(hf2(rc[j + 2], rc[j + 3]), a1, c1);
            __nv_bfloat162 p0 = __hmax2(__float22bfloat162_rn(y0), zero_bf), p1 = __hmax2(__float22bfloat162_rn(y1), zero_bf);
            packed[j / 2] = *reinterpret_cast<uint32_t*>(&p0); packed[j / 2 + 1] = *reinterpret_cast<uint32_t*>(&p1);
        }
        const uint32_t slab = act_addr + (ch >> 1) * TC_SLAB_BYTES;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            ht_store_chunk(slab, row, (ch & 1) * 4 + i, make_uint4(packed[i * 4], packed[i * 4 + 1], packed[i * 4 + 2], packed[i * 4 + 3]));
    };
    tc::tmem_ld32(taddr, r0); tc::tmem_ld_wait();
#pragma unroll 1
    for (int c2 = 0; c2 < 4; ++c2) {
        tc::tmem_ld32(taddr + (2 * c2 + 1) * 32, r1);
        pass2(r0, 2 * c2);
        tc::tmem_ld_wait();
        if (c2 < 3) tc::tmem_ld32(taddr + (2 * c2 + 2) * 32, r0);
        pass2(r1, 2 * c2 + 1);
        tc::tmem_ld_wait();
    }
}

// output head: the accumulator row already holds W5 n + b5; n_out (<= 64) floats per interval.  With
// log_softmax_seg the n_out - 4 segmentation logits leave as log_softmax (nerf/model.py:662): the row is in
// this thread's registers, so the normaliser costs ~2 instructions per class here and saves the compositing
// kernel two warp reductions per sample.
__device__ __noinline__ void ht_epilogue_out(const HeadTcArgs& a, uint32_t taddr, int64_t p, bool valid) {
    uint32_t r0[32], r1[32];
    tc::tmem_ld32(taddr, r0);
    tc::tmem_ld32(taddr + 32, r1);
    tc::tmem_ld_wait();
    if (!valid) return;
    const int n1 = a.n_out - 32;                              // columns held in r1
    float lse = 0.f;
    if (a.log_softmax_seg) {
        float m = -INFINITY;
#pragma unroll
        for (int j = 4; j < 32; ++j) if (j < a.n_out) m = fmaxf(m, __uint_as_float(r0[j]));
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j < n1) m = fmaxf(m, __uint_as_float(r1[j]));
        float sum = 0.f;
#pragma unroll
        for (int j = 4; j < 32; ++j) if (j < a.n_out) sum += __expf(__uint_as_float(r0[j]) - m);
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j < n1) sum += __expf(__uint_as_float(r1[j]) - m);
        lse = m + __logf(sum);
    }
    float* o = a.out + p * a.n_out;
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
        const float v0 = __uint_as_float(r0[j]) - (j >= 4 ? lse : 0.f), v1 = __uint_as_float(r0[j + 1]) - (j >= 4 ? lse : 0.f);
        if (j + 1 < a.n_out) *reinterpret_cast<float2*>(o + j) = make_float2(v0, v1);
        else if (j < a.n_out) o[j] = v0;
    }
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
        const float v0 = __uint_as_float(r1[j]) - lse, v1 = __uint_as_float(r1[j + 1]) - lse;
        if (j + 1 < n1) *reinterpret_cast<float2*>(o + 32 + j) = make_float2(v0, v1);
        else if (j < n1) o[32 + j] = v0;
    }
}

__global__ void __launch_bounds__(HT_THREADS, 1) head_tc_kernel(const HeadTcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (base - raw);
    const uint32_t bar0 = base + HT_SM_BAR;
    auto ring_full = [&](int s) { return bar0 + 8u * s; };
    auto ring_empty = [&](int s) { return bar0 + 8u * (HT_RING + s); };
    auto act_ready = [&](int x) { return bar0 + 8u * (2 * HT_RING + x); };
    auto acc_ready = [&](int x) { return bar0 + 8u * (2 * HT_RING + 2 + x); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + HT_SM_BAR + 8 * (2 * HT_RING + 4));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int G = gridDim.x / 2, me = blockIdx.x / 2;
    const int n_units = (a.n_tiles + 1) / 2;                         // tile pairs: this CTA takes tile 2*pair + rank
    const int n_my = (me < n_units) ? (n_units - me + G - 1) / G : 0;
    const int cnt0 = (n_my + 1) >> 1, cnt1 = n_my >> 1;

    for (int i = threadIdx.x; i < 10 * 256; i += HT_THREADS) reinterpret_cast<float*>(sm + HT_SM_VEC)[i] = a.vec[i];
    for (int i = threadIdx.x; i < 256; i += HT_THREADS) {            // ones block: 16-byte rows, k-core 0 = (1,1,1,0,0,0,0,0)
        const int kc = (i >> 3) & 1;
        reinterpret_cast<uint4*>(sm + HT_SM_ONES)[i] = kc ? make_uint4(0u, 0u, 0u, 0u) : make_uint4(0x3F803F80u, 0x00003F80u, 0u, 0u);
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < HT_RING; ++s) { tc::mbar_init(ring_full(s), leader ? 2 : 1); tc::mbar_init(ring_empty(s), 1); }
        for (int x = 0; x < 2; ++x) { tc::mbar_init(act_ready(x), 8); tc::mbar_init(acc_ready(x), 1); }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc2(tc::smem_u32((const void*)tmem_slot), 512);
    tc::fence_proxy_async_smem();
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        uint32_t it = 0;
        for (int q = 0; q < cnt0; ++q)
            for (int l = 0; l < HT_UNITS; ++l)
                for (int x = 0; x < 2; ++x) {
                    if (x == 1 && q >= cnt1) continue;
                    const uint32_t kcb = a.kcore_bytes[l];
                    const uint8_t* src = a.stages + a.unit_off[l] + (size_t)rank * a.half_kcores[l] * kcb;
                    for (int j = 0; j < a.nk[l]; ++j, ++it) {
                        const int slot = it % HT_RING; const uint32_t ph = (it / HT_RING) & 1;
                        const uint32_t bytes = (TF_STAGE_KCORES + (j == 0 ? TF_BIAS_KCORES : 0)) * kcb;
                        tc::mbar_wait_spin(ring_empty(slot), ph ^ 1, a.wait_hint);
                        if (tc::elect_one()) {
                            tc::mbar_arrive_expect_tx(ring_full(slot), bytes);
                            tc::tma_load_1d(base + HT_SM_RING + slot * TF_SLOT_BYTES, src, bytes, ring_full(slot));
                        }
                        __syncwarp();
                        src += bytes;
                    }
                }
    } else if (warp == 1) {
        if (leader) {
            // ===================== MMA issuer =====================
            constexpr uint32_t idesc_hid = tc::umma_idesc_bf16(256, 256), idesc_out = tc::umma_idesc_bf16(256, HT_OUT_N);
            constexpr uint32_t hi_a = (1024u >> 4) | (1u << 14) | (2u << 29);             // SW128 K-major slabs
            constexpr uint32_t hi_b = (TF_B_SBO >> 4) | (1u << 14);                       // no-swizzle k-cores
            constexpr uint32_t hi_ones = (HT_ONES_SBO >> 4) | (1u << 14);
            const uint32_t ring_addr = ((base + HT_SM_RING) & 0x3FFFFu) >> 4;
            const uint32_t ones_lo = (((base + HT_SM_ONES) & 0x3FFFFu) >> 4) | ((HT_ONES_LBO >> 4) << 16);
            uint32_t idesc = idesc_hid;
            auto mma = [&](uint32_t d, uint32_t ahi, uint32_t alo, uint32_t blo, uint32_t acc) {
                tc::umma_bf16_2cta(d, ((uint64_t)ahi << 32) | alo, ((uint64_t)hi_b << 32) | blo, idesc, acc);
            };
            uint32_t it = 0, act_it[2] = {0, 0};
            for (int q = 0; q < cnt0; ++q)
                for (int l = 0; l < HT_UNITS; ++l)
                    for (int x = 0; x < 2; ++x) {
                        if (x == 1 && q >= cnt1) continue;
                        tc::mbar_wait_spin(act_ready(x), act_it[x] & 1, a.wait_hint); ++act_it[x];
                        tc::tcgen05_fence_after();
                        const uint32_t act_lo = (((base + HT_SM_ACT + x * 65536) & 0x3FFFFu) >> 4) | (1u << 16);
                        const uint32_t d_tmem = tmem_base + x * 256;
                        const int nk = a.nk[l];
                        const uint32_t kstep = (2u * a.kcore_bytes[l]) >> 4;                 // one K=16 step = two k-cores
                        const uint32_t ring_lo = ring_addr | ((uint32_t)(a.kcore_bytes[l] >> 4) << 16);   // LBO = k-core stride
                        idesc = l < 5 ? idesc_hid : idesc_out;
                        for (int j = 0; j < nk; ++j, ++it) {
                            const int slot = it % HT_RING; const uint32_t ph = (it / HT_RING) & 1;
                            const uint32_t alo = act_lo + j * (TC_SLAB_BYTES >> 4);
                            uint32_t blo = ring_lo + slot * (TF_SLOT_BYTES >> 4);
                            tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                            if (tc::elect_one()) {
                                if (j == 0) {                                // acc = bias: ones block x bias block
                                    mma(d_tmem, hi_ones, ones_lo, blo, 0);
                                    blo += kstep;
                                }
                                mma(d_tmem, hi_a, alo, blo, 1);
                                mma(d_tmem, hi_a, alo + 2, blo + kstep, 1);
                                mma(d_tmem, hi_a, alo + 4, blo + 2 * kstep, 1);
                                mma(d_tmem, hi_a, alo + 6, blo + 3 * kstep, 1);
                                tc::umma_commit_2cta(ring_empty(slot));
                                if (j == nk - 1) tc::umma_commit_2cta(acc_ready(x));
                            }
                            __syncwarp();
                        }
                    }
        } else if (lane == 0) {
            // ===================== peer relay of weight arrivals =====================
            uint32_t it = 0;
            for (int q = 0; q < cnt0; ++q)
                for (int l = 0; l < HT_UNITS; ++l)
                    for (int x = 0; x < 2; ++x) {
                        if (x == 1 && q >= cnt1) continue;
                        for (int j = 0; j < a.nk[l]; ++j, ++it) {
                            const int slot = it % HT_RING; const uint32_t ph = (it / HT_RING) & 1;
                            tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                            tc::mbar_arrive_cluster(tc::mapa(ring_full(slot), 0));
                        }
                    }
        }
    } else {
        // ===================== encoder + epilogue groups =====================
        const int x = (warp - 2) >> 2;
        const uint32_t row = (warp & 3) * 32 + lane;
        const uint32_t taddr = tmem_base + (((warp & 3) * 32u) << 16) + x * 256;
        const uint32_t act_addr = base + HT_SM_ACT + x * 65536;
        const int cnt = x == 0 ? cnt0 : cnt1;
        const uint32_t act_bar_remote = tc::mapa(act_ready(x), 0);
        auto signal_act = [&]() {
            tc::fence_proxy_async_smem();
            tc::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive_cluster(act_bar_remote);
        };
        uint32_t acc_it = 0;
        for (int q = 0; q < cnt; ++q) {
            const int unit = me + (2 * q + x) * G;
            const int tile = 2 * unit + (int)rank;
            int64_t p; bool valid;
            ht_encode_row(a, tile, row, act_addr, p, valid);
            signal_act();
            for (int l = 0; l < HT_UNITS; ++l) {
                tc::mbar_wait(acc_ready(x), acc_it & 1, a.wait_hint); ++acc_it;
                tc::tcgen05_fence_after();
                if (l < 5) { ht_epilogue_ln(taddr, act_addr, row, base + HT_SM_VEC + l * 2048); signal_act(); }
                else ht_epilogue_out(a, taddr, p, valid);
            }
        }
    }
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    if (warp == 1) { tc::tcgen05_fence_after(); tc::tmem_dealloc2(tmem_base, 512); }
}

// ------------------------------------------------------------------ host side ----
// Encoder + trunk + output head for B rays x (S-1) intervals: out [B (S-1), n_out] fp32 rows
// (density | colour | segmentation) and the Gaussian means [B (S-1), 3] that nerf_head_composite takes as points.
extern "C" int nerf_head_fwd_tc(const void* packed, const float* rays_o, const float* rays_d, const float* t, int64_t B, int S,
                                float base_radius, int min_deg, int n_deg, int n_out, int log_softmax_seg, float* out, float* mean_out,
                                void* stream) {
    if (B == 0 && packed && S >= 2) return NERF_OK;
    if (!packed || !rays_o || !rays_d || !t || !out || !mean_out || B < 0 || S < 2) return NERF_ERR_BAD_ARG;
    if (n_deg != 16 || n_out <= 0 || n_out > HT_OUT_N) return NERF_ERR_UNSUPPORTED;     // encoding_size = 32, 1+3+<=60 outputs
    if (B == 0) return NERF_OK;
    HeadTcLayout L; make_head_tc_layout(L);
    HeadTcArgs a{};
    a.stages = (const uint8_t*)packed; a.vec = reinterpret_cast<const float*>((const uint8_t*)packed + L.vec_off);
    for (int l = 0; l < HT_UNITS; ++l) { a.nk[l] = L.nk[l]; a.half_kcores[l] = L.half_kcores[l]; a.kcore_bytes[l] = L.kcore_bytes[l]; a.unit_off[l] = L.unit_off[l]; }
    a.rays_o = rays_o; a.rays_d = rays_d; a.t = t; a.B = B; a.S = S; a.P = B * (int64_t)(S - 1); a.n_tiles = (int)((a.P + 127) / 128);
    a.base_radius = base_radius; a.min_deg = min_deg; a.out = out; a.n_out = n_out; a.mean_out = mean_out; a.log_softmax_seg = log_softmax_seg;
    static const uint32_t wait_hint = nerf_env_u32("NERF_TC_WAIT_HINT", 0x989680u);
    a.wait_hint = wait_hint;
    NERF_CUDA(nerf_smem_opt_in<head_tc_kernel>(HT_SMEM_BYTES));
    const int sms = nerf_num_sms();
    const int pairs = (a.n_tiles + 1) / 2, clusters = pairs < sms / 2 ? pairs : sms / 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(HT_THREADS); cfg.dynamicSmemBytes = HT_SMEM_BYTES; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    NERF_CUDA(cudaLaunchKernelEx(&cfg, head_tc_kernel, a));
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
