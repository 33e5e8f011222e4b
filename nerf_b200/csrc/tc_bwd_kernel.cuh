// Backward of the G1 MLP on tcgen05 (bf16 operands, fp32 accumulate in TMEM): the data-gradient
// chain of one 128-point tile stays on the SM, like the forward's activations.
//
// With the previous LayerNorm's affine folded into the weights (tc_pack.cu) unit l computes
//     z_l = W'_l [n_{l-1} | enc] + b'_l,   r_l = relu(z_l),   n_l = (r_l - mean) * rstd
// and the heads read n_7 (density) and n_9 (colour).  Per tile, going down from l = 9:
//     g_9 = W'_c^T d_rgb                               (CUDA cores, 3 FMAs per element)
//     dz_l = 1[n_l > shift] * rstd * (g_l - mean(g_l) - n_l * mean(g_l n_l))     (epilogue)
//     g_{l-1} = dz_l W'_l[:, :256]  (+ d_sigma * w'_sigma for l-1 = 7)           (tcgen05, N=256, K=256)
// n_l comes from the slab image the training forward stored (TMA bulk load into the slot's
// activation buffer); dz_l overwrites it IN PLACE, which is exactly the K-major A operand of the
// next MMA, and is also bulk-stored to HBM for the weight-gradient GEMM (tc_wgrad.cu).  As soon as
// the MMA has consumed a K-slab (tcgen05.commit per slab) the slab is refilled with n_{l-1}.
//
// Roles (320 threads, one persistent CTA per SM, CTA pairs with cta_group::2 like the forward):
//   warp 0      TMA producer of the transposed weight stages (tc_pack.cu, pack_bwd_stage_kernel)
//   warp 1      MMA issuer (leader CTA) / weight-arrival relay (peer CTA)
//   warps 2-5   LayerNorm-backward epilogue of tile slot 0 (thread = point row = TMEM lane);
//               thread 0 of the group also issues the slot's bulk loads and stores
//   warps 6-9   same for tile slot 1
#pragma once
#include "tc_layout.cuh"

#define TB_THREADS 320
#define TB_RING 4
#define TB_SM_ACT 0                                   // 2 x 64 KB
#define TB_SM_RING (2 * 65536)                        // TB_RING x 16 KB
#define TB_SM_VEC (TB_SM_RING + TB_RING * TC_STAGE_BYTES)      // head weights [4][256] fp32
#define TB_SM_BAR (TB_SM_VEC + 4 * 256 * 4)
#define TB_SM_TOTAL (TB_SM_BAR + 320)
#define TB_SMEM_BYTES (TB_SM_TOTAL + 1024)

struct BwdArgs {
    const uint8_t* stages;            // transposed weight stages: unit l in 1..9 -> 8 stages at (l-1)*8
    const float* vec;                 // forward epilogue vectors (head weights with gamma folded)
    const uint8_t* act;               // [10][tr_tiles][4 slabs] normalised unit outputs (forward)
    const float2* stats;              // [10][P_pad] (rstd, -mean*rstd)
    const float* d_sigma;             // [P]
    const float* d_rgb;               // [P,3]
    uint8_t* dz;                      // [10][tr_tiles][4 slabs] out
    uint8_t* dh;                      // [tr_tiles][4 KB] out: (d_sigma, d_rgb) small-operand image
    float* hsum;                      // [4] += sum over points of (d_sigma, d_rgb)
    int64_t P, P_pad; int n_tiles, tr_tiles;
    uint32_t wait_hint;
    long long* prof;                  // NERF_TC_PROF: clock64 stamps of CTA 0, second tile of each slot
    int pf_dist, pf_next;             // L2 prefetch distance in units (0 = off), and whether to reach into the slot's next tile
};

__device__ __forceinline__ float bf_lo(uint32_t x) { return __uint_as_float(x << 16); }
__device__ __forceinline__ float bf_hi(uint32_t x) { return __uint_as_float(x & 0xffff0000u); }
__device__ __forceinline__ uint4 lds128u(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// LayerNorm + ReLU backward of one unit for one row.
// SRC: 0 = g from TMEM, 1 = g from TMEM + d_sigma * w'_sigma (unit 7), 2 = g = W'_c^T d_rgb (unit 9)
template <int SRC>
__device__ __noinline__ void bwd_epilogue_unit(uint32_t taddr, uint32_t act_addr, uint32_t row, uint32_t hw_addr,
                                               float rstd, float shift, float ds, float dr, float dg, float db,
                                               uint32_t nfull_base, uint32_t nf_parity, uint32_t wait_hint, long long* pf = nullptr) {
    float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f), s1b = s1, s2b = s2;
    uint32_t r[32];
    // g for columns [ch*32 + j, +4)
    auto grad4 = [&](const uint32_t (&rc)[32], int ch, int j, float2& g0, float2& g1) {
        if (SRC == 2) {
            const float4 w0 = tc::lds128(hw_addr + (256 + ch * 32 + j) * 4);
            const float4 w1 = tc::lds128(hw_addr + (512 + ch * 32 + j) * 4);
            const float4 w2 = tc::lds128(hw_addr + (768 + ch * 32 + j) * 4);
            g0.x = fmaf(dr, w0.x, fmaf(dg, w1.x, db * w2.x)); g0.y = fmaf(dr, w0.y, fmaf(dg, w1.y, db * w2.y));
            g1.x = fmaf(dr, w0.z, fmaf(dg, w1.z, db * w2.z)); g1.y = fmaf(dr, w0.w, fmaf(dg, w1.w, db * w2.w));
        } else {
            g0 = make_float2(__uint_as_float(rc[j]), __uint_as_float(rc[j + 1]));
            g1 = make_float2(__uint_as_float(rc[j + 2]), __uint_as_float(rc[j + 3]));
            if (SRC == 1) {
                const float4 w = tc::lds128(hw_addr + (ch * 32 + j) * 4);
                g0 = __ffma2_rn(make_float2(ds, ds), make_float2(w.x, w.y), g0);
                g1 = __ffma2_rn(make_float2(ds, ds), make_float2(w.z, w.w), g1);
            }
        }
    };
    // ---- pass 1: sum g, sum g*n
#pragma unroll 1
    for (int ch = 0; ch < 8; ++ch) {
        if (SRC != 2) tc::tmem_ld32(taddr + ch * 32, r);
        if ((ch & 1) == 0) { tc::mbar_wait(nfull_base + 8u * (ch >> 1), nf_parity, wait_hint); if (pf) pf[ch >> 1] = clock64(); }
        if (SRC != 2) tc::tmem_ld_wait();
        const uint32_t slab = act_addr + (ch >> 1) * TC_SLAB_BYTES;
        // the four 16-byte row pieces of the chunk are requested back to back (the asm loads keep program order:
        // issued one per use they would expose the shared-memory latency 32 times per pass), two sum chains each
        uint4 nq[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) nq[i] = lds128u(slab + tc::sw128_offset(row, (ch & 1) * 4 + i));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint4 nv = nq[i];
            float2 g0, g1, g2, g3;
            grad4(r, ch, i * 8, g0, g1); grad4(r, ch, i * 8 + 4, g2, g3);
            s1 = __fadd2_rn(s1, g0); s1b = __fadd2_rn(s1b, g1); s1 = __fadd2_rn(s1, g2); s1b = __fadd2_rn(s1b, g3);
            s2 = __ffma2_rn(g0, make_float2(bf_lo(nv.x), bf_hi(nv.x)), s2);
            s2b = __ffma2_rn(g1, make_float2(bf_lo(nv.y), bf_hi(nv.y)), s2b);
            s2 = __ffma2_rn(g2, make_float2(bf_lo(nv.z), bf_hi(nv.z)), s2);
            s2b = __ffma2_rn(g3, make_float2(bf_lo(nv.w), bf_hi(nv.w)), s2b);
        }
    }
    s1 = __fadd2_rn(s1, s1b); s2 = __fadd2_rn(s2, s2b);
    if (pf) pf[4] = clock64();
    const float m1 = (s1.x + s1.y) * (1.0f / 256.f), m2 = (s2.x + s2.y) * (1.0f / 256.f);
    const float2 rs2 = make_float2(rstd, rstd);
    const float2 c0 = make_float2(-rstd * m1, -rstd * m1), c1 = make_float2(-rstd * m2, -rstd * m2);
    const __nv_bfloat162 shift_bf = __floats2bfloat162_rn(shift, shift);
    // ---- pass 2: dz = mask * (rstd g - rstd m1 - n rstd m2), in place over n
    auto dz2 = [&](float2 g, uint32_t nbits) -> uint32_t {
        const float2 n = make_float2(bf_lo(nbits), bf_hi(nbits));
        const float2 d = __ffma2_rn(n, c1, __ffma2_rn(g, rs2, c0));
        __nv_bfloat162 dbf = __float22bfloat162_rn(d);
        const uint32_t mask = __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&nbits), shift_bf);
        return *reinterpret_cast<uint32_t*>(&dbf) & mask;
    };
    auto pass2 = [&](const uint32_t (&rc)[32], int ch) {
        const uint32_t slab = act_addr + (ch >> 1) * TC_SLAB_BYTES;
        uint4 nq[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) nq[i] = lds128u(slab + tc::sw128_offset(row, (ch & 1) * 4 + i));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t addr = slab + tc::sw128_offset(row, (ch & 1) * 4 + i);
            const uint4 nv = nq[i];
            float2 g0, g1, g2, g3;
            grad4(rc, ch, i * 8, g0, g1); grad4(rc, ch, i * 8 + 4, g2, g3);
            const uint4 o = make_uint4(dz2(g0, nv.x), dz2(g1, nv.y), dz2(g2, nv.z), dz2(g3, nv.w));
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
        }
    };
    if (SRC == 2) {
#pragma unroll 1
        for (int ch = 0; ch < 8; ++ch) pass2(r, ch);
    } else {
        // the next 32 columns of g are in flight while the current ones are turned into dz (static register double buffer)
        uint32_t rb[32];
        tc::tmem_ld32(taddr, r); tc::tmem_ld_wait();
#pragma unroll 1
        for (int c2 = 0; c2 < 4; ++c2) {
            tc::tmem_ld32(taddr + (2 * c2 + 1) * 32, rb);
            pass2(r, 2 * c2);
            tc::tmem_ld_wait();
            if (c2 < 3) tc::tmem_ld32(taddr + (2 * c2 + 2) * 32, r);
            pass2(rb, 2 * c2 + 1);
            tc::tmem_ld_wait();
        }
    }
}

template <int CG>
__global__ void __launch_bounds__(TB_THREADS, 1) g1_bwd_tc_kernel(const BwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (base - raw);
    float* hw_s = reinterpret_cast<float*>(sm + TB_SM_VEC);
    const uint32_t bar0 = base + TB_SM_BAR;
    auto ring_full = [&](int s) { return bar0 + 8u * s; };
    auto ring_empty = [&](int s) { return bar0 + 8u * (TB_RING + s); };
    auto act_ready = [&](int x) { return bar0 + 8u * (2 * TB_RING + x); };
    auto acc_ready = [&](int x) { return bar0 + 8u * (2 * TB_RING + 2 + x); };
    auto n_full = [&](int x, int ks) { return bar0 + 8u * (2 * TB_RING + 4 + x * 4 + ks); };
    auto slab_free = [&](int x, int ks) { return bar0 + 8u * (2 * TB_RING + 12 + x * 4 + ks); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + TB_SM_BAR + 8 * (2 * TB_RING + 20));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = (CG == 2) ? tc::cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    const int G = gridDim.x / CG, me = blockIdx.x / CG;
    const int n_units = (CG == 2) ? (a.n_tiles + 1) / 2 : a.n_tiles;
    const int n_my = (me < n_units) ? (n_units - me + G - 1) / G : 0;
    const int cnt0 = (n_my + 1) >> 1, cnt1 = n_my >> 1;
    const int stages_per_slab = (CG == 2) ? 1 : 2;

    for (int i = threadIdx.x; i < 4 * 256; i += TB_THREADS) hw_s[i] = a.vec[TC_VEC_HEADW + i];
    if (threadIdx.x == 0) {
        for (int s = 0; s < TB_RING; ++s) { tc::mbar_init(ring_full(s), (CG == 2 && leader) ? 2 : 1); tc::mbar_init(ring_empty(s), 1); }
        for (int x = 0; x < 2; ++x) {
            tc::mbar_init(act_ready(x), CG == 2 ? 8 : 128); tc::mbar_init(acc_ready(x), 1);
            for (int ks = 0; ks < 4; ++ks) { tc::mbar_init(n_full(x, ks), 1); tc::mbar_init(slab_free(x, ks), 1); }
        }
        tc::fence_barrier_init();
    }
    if (warp == 1) { if (CG == 2) tc::tmem_alloc2(tc::smem_u32((const void*)tmem_slot), 512); else tc::tmem_alloc(tc::smem_u32((const void*)tmem_slot), 512); }
    tc::tcgen05_fence_before();
    if (CG == 2) tc::cluster_sync_all(); else __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer: transposed weight stages of units 9..1 =====================
        // (whole warp converged; one elected lane issues)
        uint32_t it = 0;
        for (int q = 0; q < cnt0; ++q)
            for (int l = 9; l >= 1; --l)
                for (int x = 0; x < 2; ++x) {
                    if (x == 1 && q >= cnt1) continue;
                    const uint8_t* src = a.stages + (size_t)(l - 1) * 8 * TC_STAGE_BYTES;
                    for (int ks = 0; ks < 4; ++ks)
                        for (int h = 0; h < stages_per_slab; ++h, ++it) {
                            const int nh = (CG == 2) ? (int)rank : h;
                            const int slot = it % TB_RING; const uint32_t ph = (it / TB_RING) & 1;
                            tc::mbar_wait_spin(ring_empty(slot), ph ^ 1, a.wait_hint);
                            if (tc::elect_one()) {
                                tc::mbar_arrive_expect_tx(ring_full(slot), TC_STAGE_BYTES);
                                tc::tma_load_1d(base + TB_SM_RING + slot * TC_STAGE_BYTES, src + (size_t)(ks * 2 + nh) * TC_STAGE_BYTES,
                                                TC_STAGE_BYTES, ring_full(slot));
                            }
                            __syncwarp();
                        }
                }
    } else if (warp == 1) {
        if (leader) {
            // ===================== MMA issuer: g_{l-1} = dz_l W'_l (whole warp converged; elected lane issues) =====================
            // (per-stage work kept to a few 32-bit adds on precomputed descriptor words: the loop is a latency
            // chain of one warp and the tensor pipe hides only ~256 cycles of issue bubble, see fused_tc.cu)
            const uint32_t idesc = (CG == 2) ? tc::umma_idesc_bf16(256, 256) : tc::umma_idesc_bf16(128, 128);
            constexpr uint32_t hi_sw = (1024u >> 4) | (1u << 14) | (2u << 29);            // SW128 K-major: SBO, version, layout
            auto lo_sw = [](uint32_t addr) { return ((addr & 0x3FFFFu) >> 4) | (1u << 16); };
            const uint32_t ring_lo = lo_sw(base + TB_SM_RING);
            auto mma = [&](uint32_t d, uint32_t alo, uint32_t blo, uint32_t acc) {
                const uint64_t ad = ((uint64_t)hi_sw << 32) | alo, bd = ((uint64_t)hi_sw << 32) | blo;
                if (CG == 2) tc::umma_bf16_2cta(d, ad, bd, idesc, acc); else tc::umma_bf16(d, ad, bd, idesc, acc);
            };
            uint32_t it = 0, act_it[2] = {0, 0};
            for (int q = 0; q < cnt0; ++q)
                for (int l = 9; l >= 1; --l)
                    for (int x = 0; x < 2; ++x) {
                        if (x == 1 && q >= cnt1) continue;
                        const bool pfm = a.prof && blockIdx.x == 0 && q == 1 && lane == 0;
                        if (pfm) a.prof[(l * 2 + x) * 8 + 0] = clock64();
                        tc::mbar_wait_spin(act_ready(x), act_it[x] & 1, a.wait_hint); ++act_it[x];
                        tc::tcgen05_fence_after();
                        if (pfm) a.prof[(l * 2 + x) * 8 + 1] = clock64();
                        const uint32_t act_lo = lo_sw(base + TB_SM_ACT + x * 65536);
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint32_t alo = act_lo + ks * (TC_SLAB_BYTES >> 4);
                            for (int h = 0; h < stages_per_slab; ++h, ++it) {
                                const int slot = it % TB_RING; const uint32_t ph = (it / TB_RING) & 1;
                                const uint32_t blo = ring_lo + slot * (TC_STAGE_BYTES >> 4);
                                const uint32_t d_tmem = tmem_base + x * 256 + ((CG == 2) ? 0 : h * 128);
                                tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);   // TMA data -> MMA: both async proxy, no tcgen05 fence
                                if (tc::elect_one()) {
                                    mma(d_tmem, alo, blo, ks != 0);
                                    mma(d_tmem, alo + 2, blo + 2, 1);
                                    mma(d_tmem, alo + 4, blo + 4, 1);
                                    mma(d_tmem, alo + 6, blo + 6, 1);
                                    if (CG == 2) tc::umma_commit_2cta(ring_empty(slot)); else tc::umma_commit(ring_empty(slot));
                                    if (h == stages_per_slab - 1) {
                                        if (CG == 2) tc::umma_commit_2cta(slab_free(x, ks)); else tc::umma_commit(slab_free(x, ks));
                                        if (ks == 3) { if (CG == 2) tc::umma_commit_2cta(acc_ready(x)); else tc::umma_commit(acc_ready(x)); }
                                    }
                                }
                                __syncwarp();
                            }
                        }
                        if (pfm) a.prof[(l * 2 + x) * 8 + 2] = clock64();
                    }
        } else if (CG == 2 && lane == 0 && !leader) {
            // ===================== peer relay of weight arrivals =====================
            uint32_t it = 0;
            for (int q = 0; q < cnt0; ++q)
                for (int l = 9; l >= 1; --l)
                    for (int x = 0; x < 2; ++x) {
                        if (x == 1 && q >= cnt1) continue;
                        for (int ks = 0; ks < 4; ++ks, ++it) {
                            const int slot = it % TB_RING; const uint32_t ph = (it / TB_RING) & 1;
                            tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                            tc::mbar_arrive_cluster(tc::mapa(ring_full(slot), 0));
                        }
                    }
        }
    } else {
        // ===================== LayerNorm-backward epilogue groups =====================
        const int x = (warp - 2) >> 2;
        const uint32_t row = (warp & 3) * 32 + lane;
        const uint32_t taddr = tmem_base + (((warp & 3) * 32u) << 16) + x * 256;
        const uint32_t act_addr = base + TB_SM_ACT + x * 65536;
        const uint32_t hw_addr = base + TB_SM_VEC;
        const int cnt = x == 0 ? cnt0 : cnt1;
        const bool io = row == 0;                                    // the group's bulk-copy thread
        const uint32_t act_bar_remote = (CG == 2) ? tc::mapa(act_ready(x), 0) : 0u;
        uint32_t acc_it = 0, nf_it = 0, sf_it = 0;
        float hs0 = 0.f, hs1 = 0.f, hs2 = 0.f, hs3 = 0.f;
        for (int q = 0; q < cnt; ++q) {
            const int unit = me + (2 * q + x) * G;
            const int tile = (CG == 2) ? 2 * unit + (int)rank : unit;
            const int64_t p = (int64_t)tile * 128 + row;
            if (io) {                                                // n_9 of this tile -> the slot's buffer
                tc::tma_store_wait_read0();                          // (dz_0 of the previous tile has been read out)
                for (int ks = 0; ks < 4; ++ks) {
                    tc::mbar_arrive_expect_tx(n_full(x, ks), TC_SLAB_BYTES);
                    tc::tma_load_1d(act_addr + ks * TC_SLAB_BYTES, a.act + (((size_t)9 * a.tr_tiles + tile) * 4 + ks) * TC_SLAB_BYTES,
                                    TC_SLAB_BYTES, n_full(x, ks));
                }
            }
            float ds = 0.f, dr = 0.f, dg = 0.f, db = 0.f;
            if (p < a.P) { ds = a.d_sigma[p]; dr = a.d_rgb[p * 3]; dg = a.d_rgb[p * 3 + 1]; db = a.d_rgb[p * 3 + 2]; }
            hs0 += ds; hs1 += dr; hs2 += dg; hs3 += db;
            {   // small-operand image of (d_sigma, d_rgb) for the head weight gradients:
                // [16 point groups][2 column halves][8 points][8 columns] bf16
                uint8_t* o = a.dh + (size_t)tile * 4096 + (row >> 3) * 256 + (row & 7) * 16;
                *reinterpret_cast<uint4*>(o) = make_uint4(tc::pack_bf16x2(ds, dr), tc::pack_bf16x2(dg, db), 0u, 0u);
                *reinterpret_cast<uint4*>(o + 128) = make_uint4(0u, 0u, 0u, 0u);
            }
            for (int l = 9; l >= 0; --l) {
                // n_{l-1} can only land in the slot once the MMA has consumed dz_l; pull it into L2 a whole unit ahead
                if (io && a.pf_dist > 0) {
                    const int d = a.pf_dist;
                    if (l == 9) for (int e = 1; e < d; ++e) tc::tma_prefetch_l2(a.act + ((size_t)(9 - e) * a.tr_tiles + tile) * (4 * TC_SLAB_BYTES), 4 * TC_SLAB_BYTES);
                    if (l >= d) tc::tma_prefetch_l2(a.act + ((size_t)(l - d) * a.tr_tiles + tile) * (4 * TC_SLAB_BYTES), 4 * TC_SLAB_BYTES);
                    else if (a.pf_next && q + 1 < cnt) {             // n_9.. of this slot's next tile
                        const int nunit = me + (2 * (q + 1) + x) * G;
                        const int ntile = (CG == 2) ? 2 * nunit + (int)rank : nunit;
                        tc::tma_prefetch_l2(a.act + ((size_t)(10 - d + l) * a.tr_tiles + ntile) * (4 * TC_SLAB_BYTES), 4 * TC_SLAB_BYTES);
                    }
                }
                const float2 stt = a.stats[(int64_t)l * a.P_pad + p];
                const bool pfe = a.prof && blockIdx.x == 0 && q == 1 && row == 1;
                if (pfe) a.prof[(l * 2 + x) * 8 + 3] = clock64();
                if (l < 9) { tc::mbar_wait(acc_ready(x), acc_it & 1, a.wait_hint); ++acc_it; tc::tcgen05_fence_after(); }
                if (pfe) a.prof[(l * 2 + x) * 8 + 4] = clock64();
                if (l == 9) bwd_epilogue_unit<2>(taddr, act_addr, row, hw_addr, stt.x, stt.y, ds, dr, dg, db, n_full(x, 0), nf_it & 1, a.wait_hint);
                else if (l == 7) bwd_epilogue_unit<1>(taddr, act_addr, row, hw_addr, stt.x, stt.y, ds, dr, dg, db, n_full(x, 0), nf_it & 1, a.wait_hint);
                else bwd_epilogue_unit<0>(taddr, act_addr, row, hw_addr, stt.x, stt.y, ds, dr, dg, db, n_full(x, 0), nf_it & 1, a.wait_hint,
                                          (pfe && l == 5) ? a.prof + 160 + x * 8 : nullptr);
                ++nf_it;
                if (pfe) a.prof[(l * 2 + x) * 8 + 5] = clock64();
                tc::fence_proxy_async_smem();
                tc::named_bar_sync(1 + x, 128);                      // dz_l complete in shared memory
                if (pfe) a.prof[(l * 2 + x) * 8 + 6] = clock64();
                if (l > 0) {
                    tc::tcgen05_fence_before();
                    if (CG == 2) { __syncwarp(); if (lane == 0) tc::mbar_arrive_cluster(act_bar_remote); }
                    else tc::mbar_arrive(act_ready(x));
                }
                if (io) {
                    // dz_l -> HBM slab by slab, each K-slab refilled with n_{l-1} as soon as its own store has been read out
                    // and the MMA has consumed it: the refills do not queue behind the whole 64 KB store
                    uint8_t* gdz = a.dz + ((size_t)l * a.tr_tiles + tile) * (4 * TC_SLAB_BYTES);
                    const bool pfi = a.prof && blockIdx.x == 0 && q == 1 && l == 6;
                    for (int ks = 0; ks < 4; ++ks) {
                        tc::tma_store_1d(gdz + (size_t)ks * TC_SLAB_BYTES, act_addr + ks * TC_SLAB_BYTES, TC_SLAB_BYTES);
                        tc::tma_store_commit();
                        if (l > 0) {
                            if (pfi && ks == 0) a.prof[176 + x * 4] = clock64();
                            tc::mbar_wait(slab_free(x, ks), sf_it & 1, a.wait_hint);
                            if (pfi && ks == 0) a.prof[177 + x * 4] = clock64();
                            tc::tma_store_wait_read0();
                            if (pfi && ks == 0) a.prof[178 + x * 4] = clock64();
                            if (pfi && ks == 3) a.prof[179 + x * 4] = clock64();
                            tc::mbar_arrive_expect_tx(n_full(x, ks), TC_SLAB_BYTES);
                            tc::tma_load_1d(act_addr + ks * TC_SLAB_BYTES,
                                            a.act + (((size_t)(l - 1) * a.tr_tiles + tile) * 4 + ks) * TC_SLAB_BYTES, TC_SLAB_BYTES, n_full(x, ks));
                        }
                    }
                    if (l > 0) ++sf_it;
                }
            }
        }
        if (io) tc::tma_store_wait_all0();
        hs0 = warp_sum(hs0); hs1 = warp_sum(hs1); hs2 = warp_sum(hs2); hs3 = warp_sum(hs3);
        if (lane == 0 && cnt > 0) { atomicAdd(a.hsum, hs0); atomicAdd(a.hsum + 1, hs1); atomicAdd(a.hsum + 2, hs2); atomicAdd(a.hsum + 3, hs3); }
    }
    tc::tcgen05_fence_before();
    if (CG == 2) tc::cluster_sync_all(); else __syncthreads();
    if (warp == 1) { tc::tcgen05_fence_after(); if (CG == 2) tc::tmem_dealloc2(tmem_base, 512); else tc::tmem_dealloc(tmem_base, 512); }
}
