// Fused ray-gen + sample + encode + G1 MLP on tcgen05 (bf16 in, fp32 accumulate in TMEM).
//
// One persistent CTA per SM, 320 threads:
//   warp 0      TMA producer: streams the packed weight stages (16 KB: 128 n x 64 k, + 4 KB bias block) through a ring
//   warp 1      MMA issuer (one elected lane): tcgen05.mma K=16, accumulators in TMEM
//   warps 2-5   epilogue/encoder group of tile slot 0 (thread = point row = TMEM lane)
//   warps 6-9   epilogue/encoder group of tile slot 1
// Two 128-point tiles are in flight (slots), each with its own 256 TMEM columns, activation
// buffer (4 K-slabs, SW128 K-major, written by the epilogue in exactly the layout the next
// layer's A descriptor reads) and encoding slab.  While the tensor pipe runs unit l of one
// slot, the other slot's epilogue (bias, ReLU, LayerNorm statistics, normalise, bf16, heads)
// drains unit l of its tile.  Activations never leave the SM; HBM sees rays in and
// (t, sigma, rgb) out.
#pragma once
#include "tc_layout.cuh"

#define TC_THREADS 352                            // 10 role warps + the training forward's activation-store warp
#define TC_RING 3
#define SM_ACT 0                                  // 2 x 64 KB
#define SM_PE (2 * 65536)                         // 2 x 16 KB
#define SM_RING (SM_PE + 2 * TC_SLAB_BYTES)       // TC_RING x 20 KB
#define SM_VEC (SM_RING + TC_RING * TF_SLOT_BYTES)         // head weights [4][256] + head biases [4], fp32
#define TC_SMVEC_FLOATS (TC_VEC_FLOATS - TC_VEC_HEADW)
#define SM_BAR (SM_VEC + ((TC_SMVEC_FLOATS * 4 + 15) & ~15))
#define SM_HSUM (SM_BAR + 192)                    // sum_k w_k of the 4 heads (fp32), inside the barrier block's slack
#define SM_TOTAL (SM_BAR + 256)
#define TC_SMEM_BYTES (SM_TOTAL + 1024)           // + slack for the manual 1024 B alignment

struct FusedArgs {
    const uint8_t* stages; const float* vec;
    int nk[10]; int half_kcores[10]; int64_t unit_off[10];
    const float* rays_o; const float* rays_d;     // [B,3] or null -> camera
    const float* cam_o; const float* cam_r; int H, W; float focal; int64_t ray_begin;
    float* ws_rays_o; float* ws_rays_d;           // written when rays come from the camera
    int64_t B; int S; int64_t P; int n_tiles;
    const float* t_in; const float* t_base; const float* u; int jitter; uint64_t seed; int64_t ray_id_base;
    float near_plane, far_plane;
    float norm; int normalize_dir;
    float* out_t; float* out_sigma; float* out_rgb;
    float* dbg; int dbg_layer;
    // training forward (TRAIN = true): slab images of every unit's normalised output and of the two
    // encodings, plus (rstd, -mean*rstd) per point and unit, for the backward and wgrad kernels
    uint8_t* tr_act; uint8_t* tr_pex; uint8_t* tr_ped; float2* tr_stats; int64_t P_pad; int tr_tiles;
    uint8_t* tr_act_lo; uint8_t* tr_pex_lo; uint8_t* tr_ped_lo;   // fp32-grade training forward (fused_tc_x3.cu): the lo parts' images
    uint32_t* tr_mask;                                            //   and the ReLU masks, [10][P_pad][8] words (bit = column active)
    const float* ray_bias;                        // conditional inputs: [2 (units 0, 4)][rb_rows][256] fp32 bias rows, or null;
    int64_t rb_rows, rb_ray0, rb_div;             //   ray r of this launch uses row (rb_ray0 + r) / rb_div (one row per ray, or per camera)
    long long* prof;                              // NERF_TC_PROF: clock64 stamps of CTA 0, second work unit
    uint32_t wait_hint;                           // mbarrier try_wait suspend-time hint (0 = plain spin)
    uint32_t store_gap;                           // TRAIN: cycles between the 16 KB pieces of an activation store
};

__device__ __forceinline__ void store_row_chunk(uint32_t slab_addr, uint32_t row, uint32_t c16, uint4 v) {
    uint32_t a = slab_addr + tc::sw128_offset(row, c16);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// sin(pi w), cos(pi w) with an exact reduction of w to [-1, 1]
__device__ __forceinline__ void sincospi_fast(float w, float& s, float& c) {
    float r = w - 2.0f * rintf(0.5f * w);
    s = __sinf(3.14159265358979f * r); c = __cosf(3.14159265358979f * r);
}

// positional encoding of NV values -> one [128 x 64] bf16 slab row (features beyond NV*2*L are zero).
// X3 (fused_tc_x3.cu): the row is written twice, hi = bf16(f) here and lo = bf16(f - hi) one slab further.
template <int NV, int L, bool X3 = false>
__device__ __forceinline__ void write_pe_row(uint32_t slab_addr, uint32_t row, const float (&v)[NV]) {
    float f[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) f[i] = 0.f;
#pragma unroll
    for (int c = 0; c < NV; ++c) {
        float w = v[c];
#pragma unroll
        for (int j = 0; j < L; ++j) {
            float s, co; sincospi_fast(w, s, co);
            f[c * 2 * L + j] = s; f[c * 2 * L + L + j] = co;
            w = w + w;                                       // exact octave doubling
        }
    }
    static_assert(NV * 2 * L <= TC_ONES_COL, "the slab's ones columns must be free");
    f[TC_ONES_COL] = f[TC_ONES_COL + 1] = f[TC_ONES_COL + 2] = 1.0f;     // x (bias hi, mid, lo) in the weights
#pragma unroll
    for (int c16 = 0; c16 < 8; ++c16) {
        uint4 q = make_uint4(tc::pack_bf16x2(f[c16 * 8 + 0], f[c16 * 8 + 1]), tc::pack_bf16x2(f[c16 * 8 + 2], f[c16 * 8 + 3]),
                             tc::pack_bf16x2(f[c16 * 8 + 4], f[c16 * 8 + 5]), tc::pack_bf16x2(f[c16 * 8 + 6], f[c16 * 8 + 7]));
        store_row_chunk(slab_addr, row, c16, q);
        if (X3) {
            uint32_t lo[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t h = (&q.x)[i];
                lo[i] = tc::pack_bf16x2(f[c16 * 8 + 2 * i] - __uint_as_float(h << 16), f[c16 * 8 + 2 * i + 1] - __uint_as_float(h & 0xffff0000u));
            }
            store_row_chunk(slab_addr + TC_SLAB_BYTES, row, c16, make_uint4(lo[0], lo[1], lo[2], lo[3]));
        }
    }
}

struct RowState { float d[3]; int64_t p; bool valid; };

// ray-gen + stratified sample + position encoding for one point row of a tile
template <bool X3 = false>
__device__ __forceinline__ void encode_row(const FusedArgs& a, int tile, uint32_t row, uint32_t pe_addr, RowState& st) {
    int64_t p = (int64_t)tile * 128 + row;
    st.p = p; st.valid = p < a.P;
    float v[3] = {0.f, 0.f, 0.f};
    st.d[0] = st.d[1] = st.d[2] = 0.f;
    if (st.valid) {
        int64_t ray = p / a.S; int k = (int)(p - ray * a.S);
        float o[3];
        if (a.rays_o) {
#pragma unroll
            for (int c = 0; c < 3; ++c) { o[c] = a.rays_o[ray * 3 + c]; st.d[c] = a.rays_d[ray * 3 + c]; }
        } else {
            int64_t id = a.ray_begin + ray;
            int col = (int)(id % a.W); int64_t q = id / a.W; int r_ = (int)(q % a.H); int cam = (int)(q / a.H);
            float cr[3]; camera_ray(r_, col, a.H, a.W, a.focal, cr);
            rotate_ray(a.cam_r + (size_t)cam * 9, cr, st.d);
#pragma unroll
            for (int c = 0; c < 3; ++c) o[c] = a.cam_o[cam * 3 + c];
            if (k == 0) {
#pragma unroll
                for (int c = 0; c < 3; ++c) { a.ws_rays_o[ray * 3 + c] = o[c]; a.ws_rays_d[ray * 3 + c] = st.d[c]; }
            }
        }
        float t;
        if (a.t_in) t = a.t_in[p];
        else {
            auto lad = [&](int kk) { return a.t_base ? a.t_base[kk] : linspace_at(a.near_plane, a.far_plane, a.S, kk); };
            float tk = lad(k); t = tk;
            if (a.jitter) {                                    // nerf/model.py:425-433
                float lo = tk, hi = tk;
                if (k > 0) lo = __fmul_rn(0.5f, __fadd_rn(tk, lad(k - 1)));
                if (k < a.S - 1) hi = __fmul_rn(0.5f, __fadd_rn(lad(k + 1), tk));
                float uu;
                if (a.u) uu = a.u[p];
                else { uint32_t r4[4]; philox4x32(a.seed, (uint64_t)(a.ray_id_base + ray), (uint32_t)(k >> 2), r4); uu = u01_from_bits(r4[k & 3]); }
                t = __fadd_rn(lo, __fmul_rn(__fsub_rn(hi, lo), uu));
            }
        }
        a.out_t[p] = t;
#pragma unroll
        for (int c = 0; c < 3; ++c) v[c] = __fdiv_rn(__fadd_rn(o[c], __fmul_rn(st.d[c], t)), a.norm);
    }
    write_pe_row<3, NERF_PE_X_OCTAVES, X3>(pe_addr, row, v);
}

template <bool X3 = false>
__device__ __forceinline__ void encode_dir_row(const FusedArgs& a, uint32_t row, uint32_t pe_addr, const RowState& st) {
    float v[3] = {st.d[0], st.d[1], st.d[2]};
    if (a.normalize_dir) {
        float n = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        float inv = n > 0.f ? 1.0f / n : 0.f;
        v[0] *= inv; v[1] *= inv; v[2] *= inv;
    }
    write_pe_row<3, NERF_PE_D_OCTAVES, X3>(pe_addr, row, v);
}

// ReLU + LayerNorm (no affine: folded into the consumer) for one unit of one tile slot; the
// accumulator already holds z = W'x + b' (the bias comes through the slab's ones columns).
// Two passes over the TMEM accumulator with packed f32x2 math and the next tcgen05.ld in
// flight while the current 32 columns are processed:
//   pass 1: v = max(z, 0), sum v, sum v^2                          (FMNMX, FADD2, FFMA2)
//   pass 2: n = max(z*rstd + shift, shift), shift = -mean*rstd  (rstd > 0, so this equals
//           relu(z)*rstd + shift); the max is taken after the bf16 conversion
//           (monotone rounding => identical result)              (FFMA2, F2FP, HMNMX2)
__device__ __forceinline__ float2 f2(uint32_t a, uint32_t b) { return make_float2(__uint_as_float(a), __uint_as_float(b)); }

// MODE: 0 = plain unit, 1 = unit 7 (+ density head), 2 = unit 9 (+ colour head, no activation
// store), 3 = debug dump of the normalised row.  Chunk loops are unrolled by two only (static
// register double-buffer) to keep the kernel inside the instruction cache.
// RB (units 0 and 4 of a conditional model): the row's ray contributes a bias vector rb[256] (the states'
// share of the unit, tc_layout.cuh), added to z in both passes.
template <int MODE, bool TRAIN, bool RB = false>
__device__ __noinline__ void epilogue_unit(const FusedArgs& a, int l, uint32_t taddr, uint32_t act_addr,
                                           uint32_t row, uint32_t vec_addr, const RowState& st, const float* rb = nullptr) {
    const uint32_t hsum_addr = vec_addr + (SM_HSUM - SM_VEC);
    uint32_t r0[32], r1[32];
    float2 s2 = make_float2(0.f, 0.f), q2 = make_float2(0.f, 0.f), s3 = s2, q3 = q2;   // two chains each: the adds are latency-bound
    float2 hp0 = make_float2(0.f, 0.f), hp1 = hp0, hp2 = hp0, hp0b = hp0, hp1b = hp0, hp2b = hp0;   // head partial sums (two chains each)
    auto pass1 = [&](const uint32_t (&rc)[32], int ch) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            float2 z0 = f2(rc[j], rc[j + 1]), z1 = f2(rc[j + 2], rc[j + 3]);
            if (RB) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(rb + ch * 32 + j));
                z0 = __fadd2_rn(z0, make_float2(b.x, b.y)); z1 = __fadd2_rn(z1, make_float2(b.z, b.w));
            }
            z0.x = fmaxf(z0.x, 0.f); z0.y = fmaxf(z0.y, 0.f); z1.x = fmaxf(z1.x, 0.f); z1.y = fmaxf(z1.y, 0.f);
            s2 = __fadd2_rn(s2, z0); s3 = __fadd2_rn(s3, z1);
            q2 = __ffma2_rn(z0, z0, q2); q3 = __ffma2_rn(z1, z1, q3);
            // heads: sum_k n_k w_k with n_k = relu(z_k) rstd + shift  =  rstd sum_k relu(z_k) w_k + shift sum_k w_k: the dot product
            // rides on the relu(z) pairs of this pass (packed FMAs), the affine part is applied once after it
            if (MODE == 1) {
                const float4 w = tc::lds128(vec_addr + (ch * 32 + j) * 4);
                hp0 = __ffma2_rn(z0, make_float2(w.x, w.y), hp0); hp0b = __ffma2_rn(z1, make_float2(w.z, w.w), hp0b);
            } else if (MODE == 2) {
                const float4 w0 = tc::lds128(vec_addr + (256 + ch * 32 + j) * 4);
                const float4 w1 = tc::lds128(vec_addr + (512 + ch * 32 + j) * 4);
                const float4 w2 = tc::lds128(vec_addr + (768 + ch * 32 + j) * 4);
                hp0 = __ffma2_rn(z0, make_float2(w0.x, w0.y), hp0); hp0b = __ffma2_rn(z1, make_float2(w0.z, w0.w), hp0b);
                hp1 = __ffma2_rn(z0, make_float2(w1.x, w1.y), hp1); hp1b = __ffma2_rn(z1, make_float2(w1.z, w1.w), hp1b);
                hp2 = __ffma2_rn(z0, make_float2(w2.x, w2.y), hp2); hp2b = __ffma2_rn(z1, make_float2(w2.z, w2.w), hp2b);
            }
        }
    };
    tc::tmem_ld32(taddr, r0); tc::tmem_ld_wait();
#pragma unroll 1
    for (int c2 = 0; c2 < 4; ++c2) {
        tc::tmem_ld32(taddr + (2 * c2 + 1) * 32, r1);
        pass1(r0, 2 * c2);
        tc::tmem_ld_wait();
        if (c2 < 3) tc::tmem_ld32(taddr + (2 * c2 + 2) * 32, r0);
        pass1(r1, 2 * c2 + 1);
        tc::tmem_ld_wait();
    }
    s2 = __fadd2_rn(s2, s3); q2 = __fadd2_rn(q2, q3);
    const float mean = (s2.x + s2.y) * (1.0f / 256.f);
    const float var = fmaxf((q2.x + q2.y) * (1.0f / 256.f) - mean * mean, 0.f);
    const float rstd = rsqrtf(var + NERF_LN_EPS);
    const float shift = -mean * rstd;
    const float2 rstd2 = make_float2(rstd, rstd), shift2 = make_float2(shift, shift);
    const __nv_bfloat162 shift_bf = __floats2bfloat162_rn(shift, shift);
    float h0 = 0.f, h1 = 0.f, h2 = 0.f;                      // head outputs
    if (MODE == 1 || MODE == 2) {
        const float4 hbias = tc::lds128(vec_addr + (TC_VEC_HEADB - TC_VEC_HEADW) * 4);   // head biases and sum_k w_k from shared memory
        const float4 hwsum = tc::lds128(hsum_addr);
        hp0 = __fadd2_rn(hp0, hp0b); hp1 = __fadd2_rn(hp1, hp1b); hp2 = __fadd2_rn(hp2, hp2b);
        if (MODE == 1) h0 = fmaf(rstd, hp0.x + hp0.y, fmaf(shift, hwsum.x, hbias.x));
        else {
            h0 = fmaf(rstd, hp0.x + hp0.y, fmaf(shift, hwsum.y, hbias.y));
            h1 = fmaf(rstd, hp1.x + hp1.y, fmaf(shift, hwsum.z, hbias.z));
            h2 = fmaf(rstd, hp2.x + hp2.y, fmaf(shift, hwsum.w, hbias.w));
        }
    }
    auto pass2 = [&](const uint32_t (&rc)[32], int ch) {
        uint32_t packed[16];
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            float2 c0 = shift2, c1 = shift2;
            if (RB) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(rb + ch * 32 + j));
                c0 = __ffma2_rn(make_float2(b.x, b.y), rstd2, shift2); c1 = __ffma2_rn(make_float2(b.z, b.w), rstd2, shift2);
            }
            const float2 y0 = __ffma2_rn(f2(rc[j], rc[j + 1]), rstd2, c0);
            const float2 y1 = __ffma2_rn(f2(rc[j + 2], rc[j + 3]), rstd2, c1);
            __nv_bfloat162 p0 = __hmax2(__float22bfloat162_rn(y0), shift_bf);
            __nv_bfloat162 p1 = __hmax2(__float22bfloat162_rn(y1), shift_bf);
            packed[j / 2] = *reinterpret_cast<uint32_t*>(&p0); packed[j / 2 + 1] = *reinterpret_cast<uint32_t*>(&p1);
            if (MODE == 3 && st.valid) {
                const float n0 = fmaxf(y0.x, shift), n1 = fmaxf(y0.y, shift), n2 = fmaxf(y1.x, shift), n3 = fmaxf(y1.y, shift);
                float* o = a.dbg + st.p * 256 + ch * 32 + j;
                o[0] = n0; o[1] = n1; o[2] = n2; o[3] = n3;
            }
        }
        if (MODE != 2 || TRAIN) {
            const uint32_t slab = act_addr + (ch >> 1) * TC_SLAB_BYTES;
#pragma unroll
            for (int i = 0; i < 4; ++i)
                store_row_chunk(slab, row, (ch & 1) * 4 + i, make_uint4(packed[i * 4], packed[i * 4 + 1], packed[i * 4 + 2], packed[i * 4 + 3]));
        }
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
    if (TRAIN) a.tr_stats[(int64_t)l * a.P_pad + st.p] = make_float2(rstd, shift);
    if (st.valid) {
        if (MODE == 1) a.out_sigma[st.p] = h0;
        if (MODE == 2) { a.out_rgb[st.p * 3] = h0; a.out_rgb[st.p * 3 + 1] = h1; a.out_rgb[st.p * 3 + 2] = h2; }
    }
}
