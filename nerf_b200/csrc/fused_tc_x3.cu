// fp32-grade forward on the tensor cores ("bf16x3"): ray-gen + sample + encode + G1 MLP, the same persistent
// CTA-pair kernel as fused_tc.cu, with every GEMM operand split into two bf16 parts,
//     a = a_hi + a_lo,  w = w_hi + w_lo,   a w  ~=  a_hi w_hi + a_lo w_hi + a_hi w_lo      (fp32 accumulate in TMEM),
// i.e. three tcgen05.mma (kind::f16) per K = 16 step.  The dropped a_lo w_lo term is ~2^-18 of the product, so the
// contraction carries ~16-17 significant bits -- the reference's arithmetic is fp32 `addmm`
// (/root/reference/nerf/model.py:525-542) and tcgen05 has no fp32 MMA.  LayerNorm statistics, heads and the
// encodings stay fp32 as in the bf16 kernel; bias' still rides in the accumulator as three bf16 parts.
//
// Shared memory is the bf16 kernel's layout, used by ONE 128-point tile per CTA instead of two:
//   SM_ACT  [hi: 4 slabs | lo: 4 slabs]   (the second tile slot's 64 KB hold the lo parts)
//   SM_PE   [hi slab | lo slab]
//   ring    3 x 16 KB [4 k-cores hi | 4 k-cores lo], then ONE 4 KB bias block (of the unit in flight; units without an
//           encoding slab) and the 8 KB exchange buffer of the epilogue's thread pairs -- the bf16 kernel's 3 x 20 KB
// so a ring stage covers K = 32 (two K-steps, six MMAs).  One tile per CTA (the price of fp32-grade operands in 227 KB is
// ~3x the MMAs and no tile ping-pong), but TWO accumulators (units alternate between TMEM columns 0 and 256): pass 2 of the
// epilogue of unit l writes the activation slabs in K order and releases the MMAs of unit l+1 slab by slab, so they run
// under the rest of the pass instead of after it.
//   warp 0    TMA producer      warp 1    MMA issuer (leader) / weight-arrival relay (peer)
//   warps 2-9 encoder + epilogue, TWO threads per point row (warps w and w + 4 share a TMEM lane quarter; each takes 128
//             of the row's 256 columns and the row sums are exchanged through shared memory).  With one tile slot and one
//             thread per row there was one epilogue warp per SMSP and every dependent instruction's latency was exposed.
// TRAIN: the slab images go to HBM as coalesced st.global copies by the same 256 threads while the MMAs of the next
// unit run.  As bulk stores they shared the SM's one TMA queue (~32 B/cycle) with the weight stages, which then arrived
// late: the measured unit time was 19k cycles against 6k of MMAs.
#include "tc_fused_kernel.cuh"
#include <stdio.h>

#define X3_THREADS 320
#define X3_SM_BIAS (SM_RING + TC_RING * X3_GROUP_BYTES)                 // 4 KB
#define X3_SM_XCH (X3_SM_BIAS + TF_BIAS_KCORES * TF_KCORE_BYTES)       // [128 rows][2 halves][8 floats]
static_assert(X3_SM_XCH + 128 * 2 * 32 <= SM_VEC, "the x3 ring, bias block and exchange buffer take the bf16 kernel's ring bytes");

__device__ __forceinline__ uint4 lds128v(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

// ReLU + LayerNorm epilogue of one unit (the arithmetic of epilogue_unit in tc_fused_kernel.cuh); pass 2 writes the
// normalised row twice: hi = bf16(n), lo = bf16(n - hi).   MODE: 0 plain, 1 unit 7 (+ density), 2 unit 9 (+ colour, no store)
// TRAIN: unit 9 is normalised and stored too, (rstd, shift) go to tr_stats, and the ReLU mask of the row goes to tr_mask as
// 8 words (bit = column active): the backward needs the mask EXACTLY (a fraction f of wrongly masked units is a relative
// gradient error of sqrt(f)) and a reconstructed hi + lo > shift is not an exact test.  The MMA operands (hi, lo) are those
// of the inference kernel, bit for bit.  (A third residual part n - hi - lo for the LayerNorm backward was built and
// measured: no change in any gradient -- the differences to the fp32 path are the gradient's conditioning)
// This thread takes columns 128 * half .. + 127; both threads of a row form the same (rstd, shift) from the exchanged sums
// (a + b = b + a in fp32), half 0 writes the row's outputs.
template <int MODE, bool TRAIN>
__device__ __noinline__ void epilogue_unit_x3(const FusedArgs& a, int l, uint32_t taddr, uint32_t act_addr, uint32_t row,
                                              uint32_t vec_addr, const RowState& st, int half, uint32_t xch_addr,
                                              uint32_t slab_bar_remote, bool signal) {
    const uint32_t hsum_addr = vec_addr + (SM_HSUM - SM_VEC);
    uint32_t r0[32], r1[32];
    float2 s2 = make_float2(0.f, 0.f), q2 = s2, s3 = s2, q3 = s2;
    float2 hp0 = s2, hp1 = s2, hp2 = s2, hp0b = s2, hp1b = s2, hp2b = s2;
    auto pass1 = [&](const uint32_t (&rc)[32], int ch) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            float2 z0 = f2(rc[j], rc[j + 1]), z1 = f2(rc[j + 2], rc[j + 3]);
            z0.x = fmaxf(z0.x, 0.f); z0.y = fmaxf(z0.y, 0.f); z1.x = fmaxf(z1.x, 0.f); z1.y = fmaxf(z1.y, 0.f);
            s2 = __fadd2_rn(s2, z0); s3 = __fadd2_rn(s3, z1);
            q2 = __ffma2_rn(z0, z0, q2); q3 = __ffma2_rn(z1, z1, q3);
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
    // this thread's chunks: 2k + half, k = 0..3 -- one 32-column chunk of every 64-column slab, so that slab k is complete
    // when both threads of the row have done their k-th chunk of pass 2 and the next unit's MMAs over it can start
    tc::tmem_ld32(taddr + half * 32, r0); tc::tmem_ld_wait();
#pragma unroll 1
    for (int c2 = 0; c2 < 2; ++c2) {
        tc::tmem_ld32(taddr + (4 * c2 + 2 + half) * 32, r1);
        pass1(r0, 4 * c2 + half);
        tc::tmem_ld_wait();
        if (c2 < 1) tc::tmem_ld32(taddr + (4 * c2 + 4 + half) * 32, r0);
        pass1(r1, 4 * c2 + 2 + half);
        tc::tmem_ld_wait();
    }
    s2 = __fadd2_rn(s2, s3); q2 = __fadd2_rn(q2, q3);
    hp0 = __fadd2_rn(hp0, hp0b); hp1 = __fadd2_rn(hp1, hp1b); hp2 = __fadd2_rn(hp2, hp2b);
    float ps = s2.x + s2.y, pq = q2.x + q2.y, ph0 = hp0.x + hp0.y, ph1 = hp1.x + hp1.y, ph2 = hp2.x + hp2.y;
    {   // the other half's sums (named barrier 2: the 256 epilogue threads)
        const uint32_t mine = xch_addr + (row * 2 + half) * 32, other = xch_addr + (row * 2 + (half ^ 1)) * 32;
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(mine), "f"(ps), "f"(pq), "f"(ph0), "f"(ph1) : "memory");
        if (MODE == 2) asm volatile("st.shared.f32 [%0], %1;" ::"r"(mine + 16), "f"(ph2) : "memory");
        tc::named_bar_sync(2, 256);
        float os, oq, oh0, oh1, oh2 = 0.f;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(os), "=f"(oq), "=f"(oh0), "=f"(oh1) : "r"(other) : "memory");
        if (MODE == 2) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(oh2) : "r"(other + 16) : "memory");
        ps += os; pq += oq; ph0 += oh0; ph1 += oh1; ph2 += oh2;
    }
    const float mean = ps * (1.0f / 256.f);
    const float var = fmaxf(pq * (1.0f / 256.f) - mean * mean, 0.f);
    const float rstd = rsqrtf(var + NERF_LN_EPS);
    const float shift = -mean * rstd;
    const float2 rstd2 = make_float2(rstd, rstd), shift2 = make_float2(shift, shift);
    if (MODE == 1 || MODE == 2) {
        const float4 hbias = tc::lds128(vec_addr + (TC_VEC_HEADB - TC_VEC_HEADW) * 4);
        const float4 hwsum = tc::lds128(hsum_addr);
        if (st.valid && half == 0) {
            if (MODE == 1) a.out_sigma[st.p] = fmaf(rstd, ph0, fmaf(shift, hwsum.x, hbias.x));
            else {
                a.out_rgb[st.p * 3] = fmaf(rstd, ph0, fmaf(shift, hwsum.y, hbias.y));
                a.out_rgb[st.p * 3 + 1] = fmaf(rstd, ph1, fmaf(shift, hwsum.z, hbias.z));
                a.out_rgb[st.p * 3 + 2] = fmaf(rstd, ph2, fmaf(shift, hwsum.w, hbias.w));
            }
        }
    }
    if (TRAIN && half == 0) a.tr_stats[(int64_t)l * a.P_pad + st.p] = make_float2(rstd, shift);
    if (MODE == 2 && !TRAIN) return;                               // unit 9 feeds only the colour head
    auto pass2 = [&](const uint32_t (&rc)[32], int ch) {
        uint32_t hi[16], lo[16], mask = 0;
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
            float2 y = __ffma2_rn(f2(rc[j], rc[j + 1]), rstd2, shift2);
            if (TRAIN) mask |= (y.x > shift ? 1u << j : 0u) | (y.y > shift ? 2u << j : 0u);
            y.x = fmaxf(y.x, shift); y.y = fmaxf(y.y, shift);      // = relu(z) rstd + shift  (rstd > 0)
            const __nv_bfloat162 h = __float22bfloat162_rn(y);
            const float2 hf = __bfloat1622float2(h);
            const __nv_bfloat162 l2 = __float22bfloat162_rn(make_float2(y.x - hf.x, y.y - hf.y));
            hi[j / 2] = *reinterpret_cast<const uint32_t*>(&h); lo[j / 2] = *reinterpret_cast<const uint32_t*>(&l2);

        }
        const uint32_t slab = act_addr + (ch >> 1) * TC_SLAB_BYTES;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            store_row_chunk(slab, row, (ch & 1) * 4 + i, make_uint4(hi[i * 4], hi[i * 4 + 1], hi[i * 4 + 2], hi[i * 4 + 3]));
            store_row_chunk(slab + 65536, row, (ch & 1) * 4 + i, make_uint4(lo[i * 4], lo[i * 4 + 1], lo[i * 4 + 2], lo[i * 4 + 3]));
        }
        if (TRAIN) a.tr_mask[((size_t)l * a.P_pad + st.p) * 8 + ch] = mask;
    };
    // slab k of the activations is written: release the next unit's MMAs over it (one arrival per warp and slab; the leader's
    // barrier k counts the 16 epilogue warps of the pair)
    auto slab_done = [&](int k) {
        if (!signal) return;
        tc::fence_proxy_async_smem();
        tc::tcgen05_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) tc::mbar_arrive_cluster(slab_bar_remote + 8u * k);
    };
    tc::tmem_ld32(taddr + half * 32, r0); tc::tmem_ld_wait();
#pragma unroll 1
    for (int c2 = 0; c2 < 2; ++c2) {
        tc::tmem_ld32(taddr + (4 * c2 + 2 + half) * 32, r1);
        pass2(r0, 4 * c2 + half);
        tc::tmem_ld_wait();
        slab_done(2 * c2);
        if (c2 < 1) tc::tmem_ld32(taddr + (4 * c2 + 4 + half) * 32, r0);
        pass2(r1, 4 * c2 + 2 + half);
        tc::tmem_ld_wait();
        slab_done(2 * c2 + 1);
    }
}

template <bool TRAIN>
__global__ void __launch_bounds__(X3_THREADS, 1) g1_fused_x3_kernel(const FusedArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (base - raw);
    float* vec_s = reinterpret_cast<float*>(sm + SM_VEC);
    const uint32_t bar0 = base + SM_BAR;
    auto ring_full = [&](int s) { return bar0 + 8u * s; };
    auto ring_empty = [&](int s) { return bar0 + 8u * (TC_RING + s); };
    // pe_ready: the tile's encoding slab is written (unit 0's MMAs); slab_ready[k]: activation slab k of the unit in the epilogue
    // is written (the next unit's MMAs over K-slab k); acc_ready: a unit's MMAs are complete
    const uint32_t pe_ready = bar0 + 8u * (2 * TC_RING), acc_ready = bar0 + 8u * (2 * TC_RING + 1);
    auto slab_ready = [&](int k) { return bar0 + 8u * (2 * TC_RING + 2 + k); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + SM_BAR + 8 * (2 * TC_RING + 6));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int G = gridDim.x / 2, me = blockIdx.x / 2;
    const int n_units = (a.n_tiles + 1) / 2;                       // tile pairs; this CTA takes tile 2 * unit + rank
    const int n_my = (me < n_units) ? (n_units - me + G - 1) / G : 0;

    for (int i = threadIdx.x; i < TC_SMVEC_FLOATS; i += X3_THREADS) vec_s[i] = a.vec[TC_VEC_HEADW + i];
    if (warp < 4) {
        float sw = 0.f;
        for (int i = lane; i < 256; i += 32) sw += a.vec[TC_VEC_HEADW + warp * 256 + i];
        sw = warp_sum(sw);
        if (lane == 0) reinterpret_cast<float*>(sm + SM_HSUM)[warp] = sw;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_RING; ++s) { tc::mbar_init(ring_full(s), leader ? 2 : 1); tc::mbar_init(ring_empty(s), 1); }
        tc::mbar_init(pe_ready, 16); tc::mbar_init(acc_ready, 1);
        for (int k = 0; k < 4; ++k) tc::mbar_init(slab_ready(k), 16);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc2(tc::smem_u32((const void*)tmem_slot), 512);   // two accumulators: units alternate
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer: this CTA's n-half of every K = 32 group, hi and lo together =====================
        uint32_t it = 0;
        for (int q = 0; q < n_my; ++q)
            for (int l = 0; l < 10; ++l) {
                const uint32_t blk_bytes = tc_unit_has_bias_block(l) ? TF_BIAS_KCORES * TF_KCORE_BYTES : 0;
                const uint8_t* src = a.stages + a.unit_off[l] + (size_t)rank * a.half_kcores[l] * TF_KCORE_BYTES;
                for (int j = 0; j < 2 * a.nk[l]; ++j, ++it) {
                    const int slot = it % TC_RING; const uint32_t ph = (it / TC_RING) & 1;
                    const uint32_t front = j == 0 ? blk_bytes : 0;     // the unit's bias block precedes its first group in the image
                    tc::mbar_wait_spin(ring_empty(slot), ph ^ 1, a.wait_hint);
                    if (tc::elect_one()) {
                        tc::mbar_arrive_expect_tx(ring_full(slot), X3_GROUP_BYTES + front);
                        // One bias area for all units: bias blocks are >= 8 stages apart and the wait above covers the MMAs of
                        // every stage up to it - TC_RING (a commit completes after all earlier MMAs), the last bias MMA included.
                        if (front) tc::tma_load_1d(base + X3_SM_BIAS, src, front, ring_full(slot));
                        tc::tma_load_1d(base + SM_RING + slot * X3_GROUP_BYTES, src + front, X3_GROUP_BYTES, ring_full(slot));
                    }
                    __syncwarp();
                    src += X3_GROUP_BYTES + front;
                }
            }
    } else if (warp == 1) {
        if (leader) {
            // ===================== MMA issuer: three products per K-step =====================
            const uint32_t idesc = tc::umma_idesc_bf16(256, 256);
            constexpr uint32_t hi_a = (1024u >> 4) | (1u << 14) | (2u << 29);             // SW128 K-major slabs
            constexpr uint32_t hi_b = (TF_B_SBO >> 4) | (1u << 14);                       // no-swizzle k-cores
            auto lo_a = [](uint32_t addr) { return ((addr & 0x3FFFFu) >> 4) | (1u << 16); };
            const uint32_t ring_lo = (((base + SM_RING) & 0x3FFFFu) >> 4) | ((TF_B_LBO >> 4) << 16);
            const uint32_t bias_lo = (((base + X3_SM_BIAS) & 0x3FFFFu) >> 4) | ((TF_B_LBO >> 4) << 16);
            auto mma = [&](uint32_t d, uint32_t alo, uint32_t blo, uint32_t acc) {
                tc::umma_bf16_2cta(d, ((uint64_t)hi_a << 32) | alo, ((uint64_t)hi_b << 32) | blo, idesc, acc);
            };
            const uint32_t act_lo = lo_a(base + SM_ACT), pe_lo = lo_a(base + SM_PE);
            constexpr uint32_t ACT_LO_PART = 65536u >> 4, PE_LO_PART = TC_SLAB_BYTES >> 4;   // hi -> lo distance, in descriptor units
            // Unit l accumulates into TMEM columns (l & 1) * 256: the epilogue of unit l-1 is still reading the other accumulator when
            // these MMAs start on the first slabs it has written.
            uint32_t it = 0, pe_it = 0, slab_it = 0;
            for (int q = 0; q < n_my; ++q)
                for (int l = 0; l < 10; ++l) {
                    const bool pfm = a.prof && blockIdx.x == 0 && q == 1 && lane == 0;
                    if (pfm) a.prof[l * 8 + 0] = clock64();
                    if (l == 0) { tc::mbar_wait_spin(pe_ready, pe_it & 1, a.wait_hint); ++pe_it; tc::tcgen05_fence_after(); }
                    const uint32_t d_tmem = tmem_base + (uint32_t)(l & 1) * 256u;
                    const bool blk = tc_unit_has_bias_block(l);
                    const int ng = 2 * a.nk[l];
                    for (int j = 0; j < ng; ++j, ++it) {
                        const int slot = it % TC_RING; const uint32_t ph = (it / TC_RING) & 1;
                        const int slab = j >> 1;
                        const bool enc = (l == 0 || slab >= 4);
                        if (!enc && (j & 1) == 0) {                    // K-slab `slab` of n_{l-1}
                            tc::mbar_wait_spin(slab_ready(slab), slab_it & 1, a.wait_hint);
                            tc::tcgen05_fence_after();
                            if (pfm && slab == 0) a.prof[l * 8 + 1] = clock64();
                        }
                        const uint32_t a_hi = (enc ? pe_lo : act_lo + slab * (TC_SLAB_BYTES >> 4)) + (j & 1) * 4;   // 32 columns = 64 B = 4 units
                        const uint32_t a_lo = a_hi + (enc ? PE_LO_PART : ACT_LO_PART);
                        const uint32_t blo = ring_lo + slot * (X3_GROUP_BYTES >> 4);
                        tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                        if (tc::elect_one()) {
                            uint32_t acc = (j != 0) ? 1u : 0u;
                            if (blk && j == 0) {                   // acc = b': hi slab columns 48..63 (ones at 60..62) x bias block
                                mma(d_tmem, pe_lo + 3 * 2, bias_lo, 0); acc = 1u;
                            }
#pragma unroll
                            for (int kk = 0; kk < 2; ++kk) {
                                const uint32_t bh = blo + kk * (2 * TF_KCORE_BYTES >> 4), bl = bh + (4 * TF_KCORE_BYTES >> 4);
                                mma(d_tmem, a_hi + kk * 2, bh, acc);
                                mma(d_tmem, a_lo + kk * 2, bh, 1u);
                                mma(d_tmem, a_hi + kk * 2, bl, 1u);
                                acc = 1u;
                            }
                            tc::umma_commit_2cta(ring_empty(slot));
                            if (j == ng - 1) tc::umma_commit_2cta(acc_ready);
                        }
                        __syncwarp();
                    }
                    if (l >= 1) ++slab_it;
                    if (pfm) a.prof[l * 8 + 2] = clock64();
                }
        } else if (lane == 0) {
            // ===================== peer relay: "my weight half has landed" -> leader's full barrier =====================
            uint32_t it = 0;
            for (int q = 0; q < n_my; ++q)
                for (int l = 0; l < 10; ++l)
                    for (int j = 0; j < 2 * a.nk[l]; ++j, ++it) {
                        const int slot = it % TC_RING; const uint32_t ph = (it / TC_RING) & 1;
                        tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                        tc::mbar_arrive_cluster(tc::mapa(ring_full(slot), 0));
                    }
        }
    } else {
        // ===================== encoder + epilogue =====================
        const uint32_t row = (warp & 3) * 32 + lane;                 // TMEM lane quarter = warp % 4: warps w and w + 4 share rows
        const int half = (warp - 2) >> 2;                            // which 128 columns of the row this thread takes
        const uint32_t et = threadIdx.x - 64;                        // 0..255 within the group
        const uint32_t taddr = tmem_base + (((warp & 3) * 32u) << 16);
        const uint32_t act_addr = base + SM_ACT, pe_addr = base + SM_PE, xch_addr = base + X3_SM_XCH;
        const uint32_t pe_bar_remote = tc::mapa(pe_ready, 0), slab_bar_remote = tc::mapa(slab_ready(0), 0);
        auto signal_pe = [&]() {
            tc::fence_proxy_async_smem();
            tc::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive_cluster(pe_bar_remote);
        };
        uint32_t acc_it = 0;
        // TRAIN: the slab images (hi and lo) the backward needs, shared memory -> HBM, 16 bytes per thread and step, four in
        // flight.  Issued after the slabs are complete (`slabs_complete`) and the MMAs have been released; the MMAs only read
        // the slabs.  The next writer of a slab is a pass 2 or an encoder at least one exchange barrier later, and a thread
        // reaches that barrier only after its own copies have been read out (the loads feed its stores).
        auto slabs_complete = [&]() { tc::fence_proxy_async_smem(); tc::named_bar_sync(1, 256); };
        // (the HBM images interleave hi and lo in 4 KB pieces of 32 points: tc_layout.cuh, "slab images of the bf16x3 step")
        auto copy_hidden = [&](uint8_t* img) {                     // the 4 activation slabs, both parts
#pragma unroll 1
            for (uint32_t g = 0; g < 8; ++g) {
                const uint32_t part = g >> 2, o = (g & 3u) * 4u * 4096u + et * 16u;
                uint4 v[4];
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) v[k] = lds128v(act_addr + part * 65536u + o + k * 4096u);
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) *reinterpret_cast<uint4*>(img + x3_img_offset(o + k * 4096u, part)) = v[k];
            }
        };
        auto copy_enc = [&](uint8_t* img) {                        // the encoding slab, both parts
#pragma unroll
            for (uint32_t part = 0; part < 2; ++part) {
                uint4 v[4];
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) v[k] = lds128v(pe_addr + part * TC_SLAB_BYTES + et * 16u + k * 4096u);
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) *reinterpret_cast<uint4*>(img + x3_enc_img_offset(et * 16u + k * 4096u, part)) = v[k];
            }
        };
        for (int q = 0; q < n_my; ++q) {
            const int unit = me + q * G;
            const int tile = 2 * unit + (int)rank;
            RowState st;
            if (half == 0) encode_row<true>(a, tile, row, pe_addr, st);
            else { st.p = (int64_t)tile * 128 + row; st.valid = false; st.d[0] = st.d[1] = st.d[2] = 0.f; }
            if (TRAIN) {
                st.p = (int64_t)tile * 128 + row;                      // rows beyond P still own a row of the padded images
                slabs_complete();
            }
            signal_pe();
            if (TRAIN) copy_enc(a.tr_pex + (size_t)tile * X3_ENC_IMG_BYTES);
            for (int l = 0; l < 10; ++l) {
                const bool pfe = a.prof && blockIdx.x == 0 && q == 1 && row == 1 && half == 0;
                if (pfe) a.prof[l * 8 + 3] = clock64();
                tc::mbar_wait(acc_ready, acc_it & 1, a.wait_hint); ++acc_it;
                tc::tcgen05_fence_after();
                if (pfe) a.prof[l * 8 + 4] = clock64();
                // unit 4 was the last reader of the position slabs: the direction encoding goes in before this unit's slab signals,
                // which release MMAs that read the slab's ones columns (the bias product of unit 5)
                if (l == 4 && half == 0) encode_dir_row<true>(a, row, pe_addr, st);
                const uint32_t tacc = taddr + (uint32_t)(l & 1) * 256u;
                if (l == 7) epilogue_unit_x3<1, TRAIN>(a, l, tacc, act_addr, row, base + SM_VEC, st, half, xch_addr, slab_bar_remote, true);
                else if (l == 9) epilogue_unit_x3<2, TRAIN>(a, l, tacc, act_addr, row, base + SM_VEC, st, half, xch_addr, slab_bar_remote, false);
                else epilogue_unit_x3<0, TRAIN>(a, l, tacc, act_addr, row, base + SM_VEC, st, half, xch_addr, slab_bar_remote, true);
                if (pfe) a.prof[l * 8 + 5] = clock64();
                if (TRAIN) slabs_complete();
                if (pfe) a.prof[l * 8 + 6] = clock64();
                if (TRAIN) {
                    copy_hidden(a.tr_act + ((size_t)l * a.tr_tiles + tile) * X3_IMG_BYTES);
                    if (l == 4) copy_enc(a.tr_ped + (size_t)tile * X3_ENC_IMG_BYTES);
                }
            }
        }
    }
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    if (warp == 1) { tc::tcgen05_fence_after(); tc::tmem_dealloc2(tmem_base, 512); }
}

// launches the x3 kernel on a.n_tiles tiles; `packed` = the image of nerf_g1_pack_weights_x3
int launch_fused_x3(const void* packed, FusedArgs& a, cudaStream_t st, bool train) {
    X3Layout X; make_x3_layout(X);
    for (int l = 0; l < 10; ++l) { a.nk[l] = X.nk[l]; a.unit_off[l] = X.unit_off[l]; a.half_kcores[l] = (int)(X.half_bytes[l] / TF_KCORE_BYTES); }
    a.stages = (const uint8_t*)packed;
    a.vec = reinterpret_cast<const float*>((const uint8_t*)packed + X.weight_bytes);
    static const uint32_t wait_hint = nerf_env_u32("NERF_TC_WAIT_HINT", 0x989680u);
    a.wait_hint = wait_hint; a.prof = nullptr; a.ray_bias = nullptr;
    static const bool want_prof = nerf_env_set("NERF_X3_PROF");   // debug timeline of CTA 0's second tile: allocates and synchronises, never on by default
    if (want_prof) { NERF_CUDA(cudaMalloc(&a.prof, 80 * sizeof(long long))); NERF_CUDA(cudaMemset(a.prof, 0, 80 * sizeof(long long))); }
    NERF_CUDA(train ? nerf_smem_opt_in<g1_fused_x3_kernel<true>>(TC_SMEM_BYTES) : nerf_smem_opt_in<g1_fused_x3_kernel<false>>(TC_SMEM_BYTES));
    const int sms = nerf_num_sms();
    const int pairs = (a.n_tiles + 1) / 2, clusters = pairs < sms / 2 ? pairs : sms / 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(X3_THREADS); cfg.dynamicSmemBytes = TC_SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension; attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (train) NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_fused_x3_kernel<true>, a));
    else NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_fused_x3_kernel<false>, a));
    NERF_CUDA_CHECK_LAUNCH();
    if (a.prof) {
        long long h[80];
        NERF_CUDA(cudaStreamSynchronize(st));
        NERF_CUDA(cudaMemcpy(h, a.prof, sizeof(h), cudaMemcpyDeviceToHost));
        const long long t0 = h[3];
        for (int l = 0; l < 10; ++l) {
            const long long* r = h + l * 8;
            fprintf(stderr, "x3 fwd%s unit %d: mma wait act %lld..%lld issued %lld | epi: start %lld acc_ready %lld math done %lld signalled %lld\n",
                    train ? " (train)" : "", l, r[0] - t0, r[1] - t0, r[2] - t0, r[3] - t0, r[4] - t0, r[5] - t0, r[6] - t0);
        }
        cudaFree(a.prof); a.prof = nullptr;
    }
    return NERF_OK;
}
