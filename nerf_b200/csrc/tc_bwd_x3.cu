// fp32-grade data-gradient kernel of the G1 MLP ("bf16x3" training): the chain of tc_bwd_kernel.cuh with every GEMM
// operand split into two bf16 parts (dz = dz_hi + dz_lo, W' = W_hi + W_lo) and each product formed as
//     dz W'  ~=  dz_hi W_hi + dz_lo W_hi + dz_hi W_lo            (fp32 accumulate in TMEM),
// i.e. the autograd of the reference's fp32 nn.Linear / LayerNorm / ReLU stack (/root/reference/nerf/model.py:525-542 in
// the HEAD tree; the G1 stack of examples/nerf.pth) with tensor-core contractions.
//
// One 128-point tile per CTA (CTA pairs, cta_group::2, M = 256 = the pair's two tiles), ONE accumulator (256 TMEM columns):
// the second tile slot's shared memory of the bf16 kernel holds the lo parts here.
//   warp 0     TMA producer of the transposed weight stages: per K-slab the hi stage, then the lo stage
//   warp 1     MMA issuer (leader CTA) / weight-arrival relay (peer CTA)
//   warps 2-9  LayerNorm + ReLU backward in fp32, TWO threads per point row (warps w and w + 4 share a TMEM lane quarter;
//              each takes 128 of the row's 256 columns, the two row sums are exchanged through shared memory): with one
//              tile slot there is one such warp pair per SMSP, and a single warp per SMSP left every dependent
//              instruction's latency exposed (ncu: 1.5 warps per SMSP, issue slots 20 % busy).
//                  dz = mask * rstd * (g - mean(g) - n * mean(g n)),   n = n_hi + n_lo (forward's slab images),
//              the mask is the forward's exact bitmask (8 words per row and unit, fused_tc_x3.cu);
//              dz overwrites n in place as (hi, lo) = the A operands of the next MMAs; the same 256 threads bring n in
//              (cp.async) and copy dz out (st.global) for the weight-gradient GEMM -- see the note at cp_async16.
// Per unit the steps are serial (MMAs -> refill -> epilogue): the bf16 kernel's two-slot overlap does not fit next to the
// lo parts in 227 KB.  n_{l-1} is prefetched into L2 while the MMAs of unit l run.  (Releasing the next MMAs slab by slab
// from inside pass 2, into a second accumulator -- what the forward does -- was measured here and is not kept: pass 2 reads
// and writes both parts of the image, ~73 B/cycle of shared-memory traffic, the MMAs need ~64 of the same 128, and the two
// only slow each other down: 1.36 ms against 1.30.)
#include "tc_bwd_kernel.cuh"
#include "tc_bwd_x3.cuh"
#include <stdio.h>

#define BX_THREADS 320
#define BX_RING 4
#define BX_SM_HI 0                                    // 4 slabs: n_l hi -> dz_l hi
#define BX_SM_LO 65536                                // 4 slabs: n_l lo -> dz_l lo
#define BX_SM_RING (2 * 65536)                        // BX_RING x 16 KB weight stages
#define BX_SM_VEC (BX_SM_RING + BX_RING * TC_STAGE_BYTES)      // head weights [4][256] fp32
#define BX_SM_XCH (BX_SM_VEC + 4 * 256 * 4)            // [128 rows][2 halves] float2: partial (sum g, sum g n)
#define BX_SM_BAR (BX_SM_XCH + 128 * 2 * 8)
#define BX_SM_TOTAL (BX_SM_BAR + 256)
#define BX_SMEM_BYTES (BX_SM_TOTAL + 1024)


__device__ __forceinline__ void sts128u(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// The tile images move on the LSU path (LDGSTS in, st.global out), not through TMA: the SM's one TMA queue drains at
// ~32 B/cycle, and with n in, dz out and the L2 prefetch on it (384 KB a unit) the 128 KB of weight stages queued behind
// them and the MMAs ran at half rate (clock64 timeline, NERF_X3_PROF).
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(__cvta_generic_to_global(src)) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void prefetch_l2_line(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(__cvta_generic_to_global(p))); }
__device__ __forceinline__ uint4 lds128v(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

// (hi, lo) bf16 pair of two fp32 values: hi = bf16(x), lo = bf16(x - hi)
__device__ __forceinline__ void split2(float x, float y, uint32_t& hi, uint32_t& lo) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
    const float2 hf = __bfloat1622float2(h);
    const __nv_bfloat162 l = __floats2bfloat162_rn(x - hf.x, y - hf.y);
    hi = *reinterpret_cast<const uint32_t*>(&h); lo = *reinterpret_cast<const uint32_t*>(&l);
}

// LayerNorm + ReLU backward of one unit for one row, fp32 throughout; dz replaces n in place as (hi, lo).
// SRC: 0 = g from TMEM, 1 = g from TMEM + d_sigma * w'_sigma (unit 7), 2 = g = W'_c^T d_rgb (unit 9).
// Two passes over this thread's HALF of the row (columns 128 * half .. + 127): sums, exchanged with the other half's
// thread, then dz; TMEM is read in both.
template <int SRC>
__device__ __noinline__ void bwd_x3_epilogue_unit(uint32_t taddr, uint32_t hi_addr, uint32_t lo_addr, uint32_t row, uint32_t hw_addr,
                                                  const uint32_t (&mw)[8], float rstd, float ds, float dr, float dg, float db,
                                                  int half, uint32_t xch_addr) {
    uint32_t r0[32], r1[32];
    const float2 ds2 = make_float2(ds, ds), dr2 = make_float2(dr, dr), dg2 = make_float2(dg, dg), db2 = make_float2(db, db);
    // g of the 8 columns ch*32 + i*8 .. +7 as four column pairs (rc = the chunk's 32 TMEM columns); packed f32x2 arithmetic
    auto grad8 = [&](const uint32_t (&rc)[32], int ch, int i, float2 (&g)[4]) {
        const int c0 = ch * 32 + i * 8;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (SRC == 2) {
                const float4 w0 = tc::lds128(hw_addr + (256 + c0 + 4 * h) * 4);
                const float4 w1 = tc::lds128(hw_addr + (512 + c0 + 4 * h) * 4);
                const float4 w2 = tc::lds128(hw_addr + (768 + c0 + 4 * h) * 4);
                g[2 * h] = __ffma2_rn(dr2, make_float2(w0.x, w0.y), __ffma2_rn(dg2, make_float2(w1.x, w1.y), __fmul2_rn(db2, make_float2(w2.x, w2.y))));
                g[2 * h + 1] = __ffma2_rn(dr2, make_float2(w0.z, w0.w), __ffma2_rn(dg2, make_float2(w1.z, w1.w), __fmul2_rn(db2, make_float2(w2.z, w2.w))));
            } else {
                g[2 * h] = make_float2(__uint_as_float(rc[i * 8 + 4 * h]), __uint_as_float(rc[i * 8 + 4 * h + 1]));
                g[2 * h + 1] = make_float2(__uint_as_float(rc[i * 8 + 4 * h + 2]), __uint_as_float(rc[i * 8 + 4 * h + 3]));
                if (SRC == 1) {
                    const float4 w = tc::lds128(hw_addr + (c0 + 4 * h) * 4);
                    g[2 * h] = __ffma2_rn(ds2, make_float2(w.x, w.y), g[2 * h]);
                    g[2 * h + 1] = __ffma2_rn(ds2, make_float2(w.z, w.w), g[2 * h + 1]);
                }
            }
        }
    };
    // n = hi + lo of the same 8 columns (four pairs) from the two 16-byte row pieces
    auto sum_n = [&](const uint4& h, const uint4& l, float2 (&n)[4]) {
        const uint32_t hv[4] = {h.x, h.y, h.z, h.w}, lv[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) n[q] = __fadd2_rn(make_float2(bf_lo(hv[q]), bf_hi(hv[q])), make_float2(bf_lo(lv[q]), bf_hi(lv[q])));
    };
    auto piece = [&](int ch, int i) { return (uint32_t)(ch >> 1) * TC_SLAB_BYTES + tc::sw128_offset(row, (ch & 1) * 4 + i); };
    // ---- pass 1: sum g, sum g n   (two chains each); the chunk's eight row pieces are requested back to back
    float2 s1a = make_float2(0.f, 0.f), s1b = s1a, s2a = s1a, s2b = s1a;
    auto pass1 = [&](const uint32_t (&rc)[32], int ch) {
        uint4 nh[4], nl[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { const uint32_t off = piece(ch, i); nh[i] = lds128u(hi_addr + off); nl[i] = lds128u(lo_addr + off); }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 g[4], n[4];
            grad8(rc, ch, i, g); sum_n(nh[i], nl[i], n);
            s1a = __fadd2_rn(s1a, g[0]); s1b = __fadd2_rn(s1b, g[1]); s1a = __fadd2_rn(s1a, g[2]); s1b = __fadd2_rn(s1b, g[3]);
            s2a = __ffma2_rn(g[0], n[0], s2a); s2b = __ffma2_rn(g[1], n[1], s2b);
            s2a = __ffma2_rn(g[2], n[2], s2a); s2b = __ffma2_rn(g[3], n[3], s2b);
        }
    };
    // ---- pass 2: dz = mask * (rstd g - rstd m1 - n rstd m2), (hi, lo) in place over (n_hi, n_lo)
    float2 c0, c1;
    const float2 rs2 = make_float2(rstd, rstd);
    auto pass2 = [&](const uint32_t (&rc)[32], int ch) {
        uint4 nh[4], nl[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { const uint32_t off = piece(ch, i); nh[i] = lds128u(hi_addr + off); nl[i] = lds128u(lo_addr + off); }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 g[4], n[4];
            grad8(rc, ch, i, g); sum_n(nh[i], nl[i], n);
            const uint32_t mask = mw[ch] >> (i * 8);
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float2 d = __ffma2_rn(n[q], c1, __ffma2_rn(g[q], rs2, c0));
                d.x = (mask & (1u << (2 * q))) ? d.x : 0.f;
                d.y = (mask & (2u << (2 * q))) ? d.y : 0.f;
                const __nv_bfloat162 h = __float22bfloat162_rn(d);
                const uint32_t hb = *reinterpret_cast<const uint32_t*>(&h);
                const float2 res = __fadd2_rn(d, make_float2(-bf_lo(hb), -bf_hi(hb)));
                const __nv_bfloat162 l = __float22bfloat162_rn(res);
                hi[q] = hb; lo[q] = *reinterpret_cast<const uint32_t*>(&l);
            }
            const uint32_t off = piece(ch, i);
            sts128u(hi_addr + off, make_uint4(hi[0], hi[1], hi[2], hi[3]));
            sts128u(lo_addr + off, make_uint4(lo[0], lo[1], lo[2], lo[3]));
        }
    };
    // both passes walk this thread's four chunks with the next chunk's TMEM load in flight (static register double buffer)
    const int cb = 4 * half;
    auto walk = [&](auto&& pass) {
        if (SRC == 2) {
#pragma unroll 1
            for (int ch = cb; ch < cb + 4; ++ch) pass(r0, ch);
        } else {
            tc::tmem_ld32(taddr + cb * 32, r0); tc::tmem_ld_wait();
#pragma unroll 1
            for (int c2 = 0; c2 < 2; ++c2) {
                tc::tmem_ld32(taddr + (cb + 2 * c2 + 1) * 32, r1);
                pass(r0, cb + 2 * c2);
                tc::tmem_ld_wait();
                if (c2 < 1) tc::tmem_ld32(taddr + (cb + 2 * c2 + 2) * 32, r0);
                pass(r1, cb + 2 * c2 + 1);
                tc::tmem_ld_wait();
            }
        }
    };
    walk(pass1);
    // the other half's sums (named barrier 2: the 256 epilogue threads)
    const float2 s1 = __fadd2_rn(s1a, s1b), s2 = __fadd2_rn(s2a, s2b);
    float p1 = s1.x + s1.y, p2 = s2.x + s2.y;
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(xch_addr + (row * 2 + half) * 8), "f"(p1), "f"(p2) : "memory");
    tc::named_bar_sync(2, 256);
    {
        float o1, o2;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(o1), "=f"(o2) : "r"(xch_addr + (row * 2 + (half ^ 1)) * 8) : "memory");
        p1 += o1; p2 += o2;
    }
    const float m1 = p1 * (1.0f / 256.f), m2 = p2 * (1.0f / 256.f);
    c0 = make_float2(-rstd * m1, -rstd * m1); c1 = make_float2(-rstd * m2, -rstd * m2);
    walk(pass2);
}

__global__ void __launch_bounds__(BX_THREADS, 1) g1_bwd_x3_kernel(const BwdX3Args a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (base - raw);
    float* hw_s = reinterpret_cast<float*>(sm + BX_SM_VEC);
    const uint32_t bar0 = base + BX_SM_BAR;
    auto ring_full = [&](int s) { return bar0 + 8u * s; };
    auto ring_empty = [&](int s) { return bar0 + 8u * (BX_RING + s); };
    const uint32_t act_ready = bar0 + 8u * (2 * BX_RING), acc_ready = bar0 + 8u * (2 * BX_RING + 1);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + BX_SM_BAR + 8 * (2 * BX_RING + 3));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int G = gridDim.x / 2, me = blockIdx.x / 2;
    const int n_units = (a.n_tiles + 1) / 2;                       // tile pairs; this CTA takes tile 2 * unit + rank
    const int n_my = (me < n_units) ? (n_units - me + G - 1) / G : 0;

    for (int i = threadIdx.x; i < 4 * 256; i += BX_THREADS) hw_s[i] = a.vec[TC_VEC_HEADW + i];
    if (threadIdx.x == 0) {
        for (int s = 0; s < BX_RING; ++s) { tc::mbar_init(ring_full(s), leader ? 2 : 1); tc::mbar_init(ring_empty(s), 1); }
        tc::mbar_init(act_ready, 16); tc::mbar_init(acc_ready, 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc2(tc::smem_u32((const void*)tmem_slot), 256);
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer: per unit 9..1 and K-slab, this CTA's half of the hi stage, then of the lo stage =====================
        uint32_t it = 0;
        for (int q = 0; q < n_my; ++q)
            for (int l = 9; l >= 1; --l)
                for (int ks = 0; ks < 4; ++ks)
                    for (int part = 0; part < 2; ++part, ++it) {
                        const uint8_t* src = (part ? a.stages_lo : a.stages_hi) + ((size_t)(l - 1) * 8 + ks * 2 + rank) * TC_STAGE_BYTES;
                        const int slot = it % BX_RING; const uint32_t ph = (it / BX_RING) & 1;
                        tc::mbar_wait_spin(ring_empty(slot), ph ^ 1, a.wait_hint);
                        if (tc::elect_one()) {
                            tc::mbar_arrive_expect_tx(ring_full(slot), TC_STAGE_BYTES);
                            tc::tma_load_1d(base + BX_SM_RING + slot * TC_STAGE_BYTES, src, TC_STAGE_BYTES, ring_full(slot));
                        }
                        __syncwarp();
                    }
    } else if (warp == 1) {
        if (leader) {
            // ===================== MMA issuer: g_{l-1} = dz_hi W_hi + dz_lo W_hi + dz_hi W_lo =====================
            const uint32_t idesc = tc::umma_idesc_bf16(256, 256);
            constexpr uint32_t hi_sw = (1024u >> 4) | (1u << 14) | (2u << 29);            // SW128 K-major: SBO, version, layout
            auto lo_sw = [](uint32_t addr) { return ((addr & 0x3FFFFu) >> 4) | (1u << 16); };
            const uint32_t ring_lo = lo_sw(base + BX_SM_RING);
            const uint32_t dzh_lo = lo_sw(base + BX_SM_HI), dzl_lo = lo_sw(base + BX_SM_LO);
            auto mma = [&](uint32_t alo, uint32_t blo, uint32_t acc) {
                tc::umma_bf16_2cta(tmem_base, ((uint64_t)hi_sw << 32) | alo, ((uint64_t)hi_sw << 32) | blo, idesc, acc);
            };
            uint32_t it = 0, act_it = 0;
            for (int q = 0; q < n_my; ++q)
                for (int l = 9; l >= 1; --l) {
                    const bool pfm = a.prof && blockIdx.x == 0 && q == 1 && lane == 0;
                    if (pfm) a.prof[l * 8 + 0] = clock64();
                    tc::mbar_wait_spin(act_ready, act_it & 1, a.wait_hint); ++act_it;
                    tc::tcgen05_fence_after();
                    if (pfm) a.prof[l * 8 + 1] = clock64();
                    for (int ks = 0; ks < 4; ++ks)
                        for (int part = 0; part < 2; ++part, ++it) {
                            const int slot = it % BX_RING; const uint32_t ph = (it / BX_RING) & 1;
                            const uint32_t blo = ring_lo + slot * (TC_STAGE_BYTES >> 4);
                            const uint32_t ah = dzh_lo + ks * (TC_SLAB_BYTES >> 4), al = dzl_lo + ks * (TC_SLAB_BYTES >> 4);
                            tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                            if (pfm && l == 5) a.prof[80 + ks * 2 + part] = clock64();
                            if (tc::elect_one()) {
                                if (part == 0) {                       // W_hi: with dz_hi, then with dz_lo
#pragma unroll
                                    for (int kk = 0; kk < 4; ++kk) mma(ah + 2 * kk, blo + 2 * kk, (ks | kk) ? 1u : 0u);
#pragma unroll
                                    for (int kk = 0; kk < 4; ++kk) mma(al + 2 * kk, blo + 2 * kk, 1u);
                                } else {                               // W_lo: with dz_hi
#pragma unroll
                                    for (int kk = 0; kk < 4; ++kk) mma(ah + 2 * kk, blo + 2 * kk, 1u);
                                }
                                tc::umma_commit_2cta(ring_empty(slot));
                                if (ks == 3 && part == 1) tc::umma_commit_2cta(acc_ready);
                            }
                            __syncwarp();
                        }
                    if (pfm) a.prof[l * 8 + 2] = clock64();
                }
        } else if (lane == 0) {
            // ===================== peer relay of weight arrivals =====================
            uint32_t it = 0;
            for (int q = 0; q < n_my; ++q)
                for (int l = 9; l >= 1; --l)
                    for (int j = 0; j < 8; ++j, ++it) {
                        const int slot = it % BX_RING; const uint32_t ph = (it / BX_RING) & 1;
                        tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                        tc::mbar_arrive_cluster(tc::mapa(ring_full(slot), 0));
                    }
        }
    } else {
        // ===================== LayerNorm-backward epilogue group =====================
        const uint32_t row = (warp & 3) * 32 + lane;                     // TMEM lane quarter = warp % 4: warps w and w + 4 share rows
        const int half = (warp - 2) >> 2;                                // which 128 columns of the row this thread takes
        const uint32_t taddr = tmem_base + (((warp & 3) * 32u) << 16);
        const uint32_t hi_addr = base + BX_SM_HI, lo_addr = base + BX_SM_LO, hw_addr = base + BX_SM_VEC, xch_addr = base + BX_SM_XCH;
        const uint32_t et = threadIdx.x - 64;                            // 0..255 within the group: owner of the 16-byte chunks et + 256 j
        const uint32_t act_bar_remote = tc::mapa(act_ready, 0);
        uint32_t acc_it = 0;
        float hs0 = 0.f, hs1 = 0.f, hs2 = 0.f, hs3 = 0.f;
        // n_l (hi, lo) of this tile -> the two buffers.  A thread writes only the chunks it read out in store_unit, so the
        // refill needs no wait for the other threads' copies.
        // (HBM images: hi and lo interleaved in 4 KB pieces of 32 points, tc_layout.cuh)
        auto load_unit = [&](int l, int tile) {
            const uint8_t* img = a.act + ((size_t)l * a.tr_tiles + tile) * X3_IMG_BYTES;
#pragma unroll 4
            for (int j = 0; j < 16; ++j) {
                const uint32_t o = ((uint32_t)j * 256u + et) * 16u;
                cp_async16(hi_addr + o, img + x3_img_offset(o, 0)); cp_async16(lo_addr + o, img + x3_img_offset(o, 1));
            }
            cp_async_wait_all();
            tc::named_bar_sync(3, 256);
        };
        // dz_l -> HBM for the weight-gradient GEMM (while the MMAs read the same shared memory); n_{l-1} -> L2
        auto store_unit = [&](int l, int tile) {
            uint8_t* img = a.dz + ((size_t)l * a.tr_tiles + tile) * X3_IMG_BYTES;
#pragma unroll 1
            for (int j0 = 0; j0 < 16; j0 += 4) {
                uint4 vh[4], vl[4];
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const uint32_t o = ((uint32_t)(j0 + jj) * 256u + et) * 16u;
                    vh[jj] = lds128v(hi_addr + o); vl[jj] = lds128v(lo_addr + o);
                }
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const uint32_t o = ((uint32_t)(j0 + jj) * 256u + et) * 16u;
                    *reinterpret_cast<uint4*>(img + x3_img_offset(o, 0)) = vh[jj]; *reinterpret_cast<uint4*>(img + x3_img_offset(o, 1)) = vl[jj];
                }
            }
            if (a.prof && blockIdx.x == 0 && et == 1 && l == 5 && tile == (int)gridDim.x) a.prof[88] = clock64();
            if (l > 0) {
                const uint8_t* nx = a.act + ((size_t)(l - 1) * a.tr_tiles + tile) * X3_IMG_BYTES;
                prefetch_l2_line(nx + et * 128u); prefetch_l2_line(nx + (et + 256u) * 128u);
                prefetch_l2_line(nx + (et + 512u) * 128u); prefetch_l2_line(nx + (et + 768u) * 128u);
            }
            if (a.prof && blockIdx.x == 0 && et == 1 && l == 5 && tile == (int)gridDim.x) a.prof[89] = clock64();
        };
        for (int q = 0; q < n_my; ++q) {
            const int unit = me + q * G;
            const int tile = 2 * unit + (int)rank;
            const int64_t p = (int64_t)tile * 128 + row;
            float ds = 0.f, dr = 0.f, dg = 0.f, db = 0.f;
            if (p < a.P) { ds = a.d_sigma[p]; dr = a.d_rgb[p * 3]; dg = a.d_rgb[p * 3 + 1]; db = a.d_rgb[p * 3 + 2]; }
            if (half == 0) { hs0 += ds; hs1 += dr; hs2 += dg; hs3 += db; }
            if (half == 0) {   // small-operand images of (d_sigma, d_rgb), hi and lo: [16 point groups][2 column halves][8 points][8 columns] bf16
                uint32_t h0, l0, h1, l1;
                split2(ds, dr, h0, l0); split2(dg, db, h1, l1);
                uint8_t* dh = a.dh + (size_t)tile * X3_DH_IMG_BYTES;
                *reinterpret_cast<uint4*>(dh + x3_dh_img_offset(row, 0, 0)) = make_uint4(h0, h1, 0u, 0u);
                *reinterpret_cast<uint4*>(dh + x3_dh_img_offset(row, 1, 0)) = make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4*>(dh + x3_dh_img_offset(row, 0, 1)) = make_uint4(l0, l1, 0u, 0u);
                *reinterpret_cast<uint4*>(dh + x3_dh_img_offset(row, 1, 1)) = make_uint4(0u, 0u, 0u, 0u);
            }
            // (rstd, shift) and the ReLU mask of this row for unit l: requested a whole copy-out ahead of their use and held in
            // registers -- asked for after store_unit they waited ~4 k cycles on a memory system busy with that copy
            float2 stt_n; uint4 m0_n, m1_n;
            auto fetch_row = [&](int l) {
                stt_n = a.stats[(int64_t)l * a.P_pad + p];
                const uint4* mp = reinterpret_cast<const uint4*>(a.mask + ((size_t)l * a.P_pad + p) * 8);
                m0_n = mp[0]; m1_n = mp[1];
            };
            fetch_row(9);
            for (int l = 9; l >= 0; --l) {
                const float2 stt = stt_n;
                uint32_t mw[8];
                mw[0] = m0_n.x; mw[1] = m0_n.y; mw[2] = m0_n.z; mw[3] = m0_n.w; mw[4] = m1_n.x; mw[5] = m1_n.y; mw[6] = m1_n.z; mw[7] = m1_n.w;
                const bool pfe = a.prof && blockIdx.x == 0 && q == 1 && row == 1 && half == 0;
                if (pfe) a.prof[l * 8 + 3] = clock64();
                if (l < 9) {
                    tc::mbar_wait(acc_ready, acc_it & 1, a.wait_hint); ++acc_it;     // g_l is in TMEM, the MMAs have consumed dz_{l+1}
                    tc::tcgen05_fence_after();
                }
                if (pfe) a.prof[l * 8 + 4] = clock64();
                load_unit(l, tile);
                if (pfe) a.prof[l * 8 + 5] = clock64();
                if (l == 9) bwd_x3_epilogue_unit<2>(taddr, hi_addr, lo_addr, row, hw_addr, mw, stt.x, ds, dr, dg, db, half, xch_addr);
                else if (l == 7) bwd_x3_epilogue_unit<1>(taddr, hi_addr, lo_addr, row, hw_addr, mw, stt.x, ds, dr, dg, db, half, xch_addr);
                else bwd_x3_epilogue_unit<0>(taddr, hi_addr, lo_addr, row, hw_addr, mw, stt.x, ds, dr, dg, db, half, xch_addr);
                if (pfe) a.prof[l * 8 + 6] = clock64();
                tc::fence_proxy_async_smem();
                tc::named_bar_sync(1, 256);                                       // dz_l complete in shared memory
                if (pfe) a.prof[l * 8 + 7] = clock64();
                if (l > 0) {
                    tc::tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive_cluster(act_bar_remote);
                }
                if (l > 0) fetch_row(l - 1);
                store_unit(l, tile);
            }
        }
        hs0 = warp_sum(hs0); hs1 = warp_sum(hs1); hs2 = warp_sum(hs2); hs3 = warp_sum(hs3);
        if (lane == 0 && n_my > 0) { atomicAdd(a.hsum, hs0); atomicAdd(a.hsum + 1, hs1); atomicAdd(a.hsum + 2, hs2); atomicAdd(a.hsum + 3, hs3); }
    }
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    if (warp == 1) { tc::tcgen05_fence_after(); tc::tmem_dealloc2(tmem_base, 256); }
}

int launch_bwd_x3(BwdX3Args& a, cudaStream_t st) {
    static const uint32_t wait_hint = nerf_env_u32("NERF_TC_WAIT_HINT", 0x989680u);
    a.wait_hint = wait_hint;
    NERF_CUDA(nerf_smem_opt_in<g1_bwd_x3_kernel>(BX_SMEM_BYTES));
    const int sms = nerf_num_sms();
    const int pairs = (a.n_tiles + 1) / 2, clusters = pairs < sms / 2 ? pairs : sms / 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(BX_THREADS); cfg.dynamicSmemBytes = BX_SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension; attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    static const bool want_prof = nerf_env_set("NERF_X3_PROF");   // debug timeline: allocates and synchronises, never on by default
    a.prof = nullptr;
    if (want_prof) { NERF_CUDA(cudaMalloc(&a.prof, 96 * sizeof(long long))); NERF_CUDA(cudaMemset(a.prof, 0, 96 * sizeof(long long))); }
    NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_bwd_x3_kernel, a));
    NERF_CUDA_CHECK_LAUNCH();
    if (a.prof) {
        long long h[96];
        NERF_CUDA(cudaStreamSynchronize(st));
        NERF_CUDA(cudaMemcpy(h, a.prof, sizeof(h), cudaMemcpyDeviceToHost));
        const long long t0 = h[9 * 8 + 3];
        for (int l = 9; l >= 0; --l) {
            const long long* r = h + l * 8;
            fprintf(stderr, "x3 bwd unit %d: mma wait act %lld..%lld issued %lld | epi: start %lld acc_ready %lld n_full %lld math done %lld synced %lld\n",
                    l, r[0] - t0, r[1] - t0, r[2] - t0, r[3] - t0, r[4] - t0, r[5] - t0, r[6] - t0, r[7] - t0);
        }
        fprintf(stderr, "x3 bwd unit 5 stage waits passed at:");
        for (int j = 0; j < 8; ++j) fprintf(stderr, " %lld", h[80 + j] - t0);
        fprintf(stderr, "; store_unit(5): copy done %lld, prefetch issued %lld\n", h[88] - t0, h[89] - t0);
        cudaFree(a.prof);
    }
    return NERF_OK;
}
