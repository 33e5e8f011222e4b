// Weight gradients of the G1 MLP on tcgen05 (bf16 operands, fp32 accumulate in TMEM).
//
//   dW'_l [256 out x K_l in] = sum over points p of  dZ_l[p, out] * X_l[p, in]
//
// (the autograd of the 10 nn.Linear layers of examples/nerf.pth).  Both
// operands are read straight from the slab images the fused forward / backward kernels left in
// HBM: a slab is [128 points x 64 features] bf16, 128-byte rows, 128B-swizzled.  The points are
// the CONTRACTION index here, so both operands are "MN-major" UMMA operands (feature index
// contiguous): the same bytes the forward consumed K-major feed this GEMM without a transpose.
//
// Work = 12 (unit | head) x n_tiles point tiles, cut into 74 equal-cost contiguous segments, one
// per CTA PAIR (cluster of 2, cta_group::2): the pair issues M=256 MMAs, CTA r holding output
// features [128r, 128r+128) (its half of dZ) and input features [128r, 128r+128) (its half of the
// B operand), so every byte is fetched from HBM/L2 exactly once per pair.  A pair accumulates its
// segment's 256 x K_l block in TMEM over all its tiles and flushes it once with vector fp32
// reductions into dW'_l.
//   warp 0   TMA producer (64-point stages; lanes 0-4 issue the stage's five 8 KB bulk copies together)
//   warp 1   MMA issuer (leader; N=256 hidden + N=128 [encoding | ones] or N=32 ones, K=16 points
//            per MMA, descriptors advanced with one 32-bit add) / arrival relay (peer)
//   warps 2-5 flush       (tcgen05.ld -> red.global.add.v4.f32)
// The kernel is HBM-bound: 64 KB of dZ + 64..80 KB of activations per 128-point tile and unit.
// (v1 ran one CTA per out-half with M=128 MMAs on 64-point stages: its single issuing thread needed
//  ~2.6 K cycles per stage for 12 small MMAs and 7 copies -- instruction-latency bound at 0.95 ms.)
#include "tc_layout.cuh"
#include <stdlib.h>
#include <stdio.h>

#define WG_THREADS 192
#define WG_STAGES 5
#define WG_HALF 8192                      // half a slab: 64 points x 128 B; one pipeline stage covers 64 points
#define WG_OFF_B (2 * WG_HALF)            // stage: [A: 2 dZ half slabs | B: 2 hidden half slabs (or 2 KB of the small operand) | E: encoding half slab]
#define WG_OFF_E (4 * WG_HALF)
#define WG_STAGE_BYTES (5 * WG_HALF)
#define WG_MAX_SEGS 24
#define WG_SM_ONES (WG_STAGES * WG_STAGE_BYTES)      // 1 KB of bf16 1.0: the "ones" operand that yields column sums
#define WG_SM_BAR (WG_SM_ONES + 1024)
#define WG_NU 12                          // 10 units + 2 head pseudo-units
#define WG_SM_SEGS (WG_SM_BAR + 128)
#define WG_SM_TOTAL (WG_SM_SEGS + WG_MAX_SEGS * 16)
#define WG_SMEM_BYTES (WG_SM_TOTAL + 1024)

// Output columns (= TMEM columns): [0,256) hidden inputs, [256,320) encoding inputs, [320,336) the
// column sums of dz (bias gradient, 16 identical copies) when `ones` is set.  Head pseudo-units put
// a normalised activation in `dz` and the (d_sigma, d_rgb) small-operand image in `xsmall`:
// out[feature][0..3] = sum_p n[p][feature] * (d_sigma, d_rgb)[p].
struct WgradArgs {
    const uint8_t* dz[WG_NU];        // [n_tiles][4 slabs]: the M operand (null = absent)
    const uint8_t* xmain[WG_NU];     // [n_tiles][4 slabs] hidden input of unit l (null = none)
    const uint8_t* xenc[WG_NU];      // [n_tiles][1 slab] encoding input of unit l (null = none)
    const uint8_t* xsmall[WG_NU];    // [n_tiles][4 KB] 16-column operand (null = none; excludes xmain)
    float* out[WG_NU];               // [256][ld] fp32, accumulated with atomics
    int ld[WG_NU];
    int ones[WG_NU];
    int n_tiles;
    uint32_t wait_hint;
    int dbg;                         // experiments: bit 0 = no MMAs, bit 1 = no loads
    long long* prof;                 // NERF_WG_PROF: per-CTA cycle counters [8]
};
// Relative time per tile of a unit, for the equal-cost cut of the work list over the CTA pairs.  Bytes alone
// (8 quarter-slabs of dZ + 8 of X + 4 of the encoding + 1 small) mispredict: a stage costs its producer thread
// ~250 cycles per bulk copy whatever the size, so the units with small stages (unit 0, the two head pseudo-units)
// run at the producer's rate, not at HBM's.  Measured on B200 (clock64 span of every pair, NERF_WG_PROF):
// cycles per tile 2254 (hidden), 2361 (hidden + encoding), 1838 (unit 0), 1875 (heads) -- with the byte costs the
// head pairs finished 45 % after the average pair.
__host__ __device__ inline int wg_cost(const WgradArgs& a, int l) {
    if (a.xmain[l]) return a.xenc[l] ? 126 : 120;
    return a.xenc[l] ? 98 : 100;
}

namespace tc {
// MN-major operand, 128-byte swizzle: atoms of 64 MN-elements (128 B) x 8 K-rows (1024 B);
// LBO = byte distance between atoms along MN, SBO = between 8-row groups along K.
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// MN-major operand without swizzle: 8 x 8 core matrices (8 K-rows of 16 B); SBO = byte distance between
// core matrices along MN, LBO = along K.
__device__ __forceinline__ uint64_t umma_desc_noswz_mn(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
}
__host__ __device__ constexpr uint32_t umma_idesc_bf16_mn(int M, int N) {
    return umma_idesc_bf16(M, N) | (1u << 15) | (1u << 16);      // a_major = b_major = MN
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
}  // namespace tc

struct WgSeg { int l, tb, te, pad; };

__device__ __forceinline__ uint64_t wg_desc(uint32_t hi, uint32_t lo) { return ((uint64_t)hi << 32) | lo; }

__global__ void __launch_bounds__(WG_THREADS, 1) g1_wgrad_tc_kernel(const WgradArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (base - raw);
    const uint32_t bar0 = base + WG_SM_BAR;
    auto full = [&](int s) { return bar0 + 8u * s; };
    auto empty = [&](int s) { return bar0 + 8u * (WG_STAGES + s); };
    const uint32_t acc_done = bar0 + 8u * (2 * WG_STAGES), acc_free = bar0 + 8u * (2 * WG_STAGES + 1);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + WG_SM_BAR + 8 * (2 * WG_STAGES + 2));
    volatile int* n_segs_s = reinterpret_cast<volatile int*>(sm + WG_SM_BAR + 8 * (2 * WG_STAGES + 2) + 4);
    WgSeg* segs = reinterpret_cast<WgSeg*>(sm + WG_SM_SEGS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int n_pairs = gridDim.x >> 1, pair = blockIdx.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < WG_STAGES; ++s) { tc::mbar_init(full(s), leader ? 2 : 1); tc::mbar_init(empty(s), 1); }
        tc::mbar_init(acc_done, 1); tc::mbar_init(acc_free, 8);
        tc::fence_barrier_init();
        // equal-cost contiguous cut of the (unit, tile) work list over the CTA pairs
        int64_t total = 0;
        for (int l = 0; l < WG_NU; ++l) if (a.dz[l]) total += (int64_t)a.n_tiles * wg_cost(a, l);
        const int64_t lo = total * pair / n_pairs, hi = total * (pair + 1) / n_pairs;
        int n = 0; int64_t pre = 0;
        for (int l = 0; l < WG_NU; ++l) {
            if (!a.dz[l]) continue;
            const int64_t c = wg_cost(a, l);
            const int64_t end = pre + c * a.n_tiles;
            const int64_t s0 = lo > pre ? lo : pre, s1 = hi < end ? hi : end;
            if (s1 > s0) {
                const int tb = (int)((s0 - pre + c - 1) / c), te = (int)((s1 - pre + c - 1) / c);
                if (te > tb && n < WG_MAX_SEGS) { segs[n].l = l; segs[n].tb = tb; segs[n].te = te; ++n; }
            }
            pre = end;
        }
        *n_segs_s = n;
    }
    // constant "ones" operands: a 1 KB no-swizzle block, and (peer CTA) the encoding position of every stage
    for (int i = threadIdx.x; i < 256; i += WG_THREADS) reinterpret_cast<uint32_t*>(sm + WG_SM_ONES)[i] = 0x3F803F80u;
    if (!leader)
        for (int s = 0; s < WG_STAGES; ++s)
            for (int i = threadIdx.x; i < WG_HALF / 4; i += WG_THREADS)
                reinterpret_cast<uint32_t*>(sm + s * WG_STAGE_BYTES + WG_OFF_E)[i] = 0x3F803F80u;
    tc::fence_proxy_async_smem();
    if (warp == 1) tc::tmem_alloc2(tc::smem_u32((const void*)tmem_slot), 512);
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_segs = *n_segs_s;

    if (warp == 0) {
        // ===================== TMA producer: this CTA's halves of both operands =====================
        // (whole warp, warp-uniform addresses; one elected lane issues the stage's bulk copies)
        uint32_t it = 0;
        for (int si = 0; si < n_segs; ++si) {
            const WgSeg sg = segs[si];
            const uint8_t* xm = a.xmain[sg.l]; const uint8_t* xs = a.xsmall[sg.l];
            const uint8_t* xe = leader ? a.xenc[sg.l] : nullptr;
            const uint32_t bytes = 2 * WG_HALF + (xm ? 2 * WG_HALF : 0) + (xe ? WG_HALF : 0) + (xs ? 2048 : 0);
            const uint8_t* pa = a.dz[sg.l] + ((size_t)sg.tb * 4 + 2 * rank) * TC_SLAB_BYTES;
            const uint8_t* pm = xm ? xm + ((size_t)sg.tb * 4 + 2 * rank) * TC_SLAB_BYTES : nullptr;
            const uint8_t* pe = xe ? xe + (size_t)sg.tb * TC_SLAB_BYTES : nullptr;
            const uint8_t* ps = xs ? xs + (size_t)sg.tb * 4096 : nullptr;
            for (int n = 2 * (sg.te - sg.tb), hf = 0; n > 0; --n, ++it, hf ^= 1) {
                const int slot = it % WG_STAGES; const uint32_t ph = (it / WG_STAGES) & 1;
                const long long c0 = a.prof ? clock64() : 0;
                tc::mbar_wait(empty(slot), ph ^ 1, a.wait_hint);
                if (a.prof && lane == 0) a.prof[blockIdx.x * 8 + 0] += clock64() - c0;
                const uint32_t st = base + slot * WG_STAGE_BYTES, fb = full(slot);
                if (tc::elect_one()) {
                    if (a.dbg & 2) tc::mbar_arrive(fb);
                    else {
                        tc::mbar_arrive_expect_tx(fb, bytes);
                        tc::tma_load_1d(st, pa, WG_HALF, fb);
                        tc::tma_load_1d(st + WG_HALF, pa + TC_SLAB_BYTES, WG_HALF, fb);
                        if (pm) { tc::tma_load_1d(st + WG_OFF_B, pm, WG_HALF, fb); tc::tma_load_1d(st + WG_OFF_B + WG_HALF, pm + TC_SLAB_BYTES, WG_HALF, fb); }
                        if (pe) tc::tma_load_1d(st + WG_OFF_E, pe, WG_HALF, fb);
                        if (ps) tc::tma_load_1d(st + WG_OFF_B, ps, 2048, fb);
                    }
                }
                __syncwarp();
                const size_t adv = hf ? (size_t)(4 * TC_SLAB_BYTES - WG_HALF) : (size_t)WG_HALF;
                pa += adv; if (pm) pm += adv; if (pe) pe += WG_HALF; if (ps) ps += 2048;
            }
        }
    } else if (warp == 1) {
        if (leader) {
            // ===================== MMA issuer =====================
            constexpr uint32_t idesc_main = tc::umma_idesc_bf16_mn(256, 256), idesc_enc = tc::umma_idesc_bf16_mn(256, 128);
            constexpr uint32_t idesc_32 = tc::umma_idesc_bf16_mn(256, 32);
            // descriptor words: hi = SBO | version | layout, lo = start address | LBO; K advances by 2048 B (+128 in lo)
            constexpr uint32_t hi_sw = (1024u >> 4) | (1u << 14) | (2u << 29), hi_ns = (128u >> 4) | (1u << 14);
            constexpr uint32_t lbo_half = (uint32_t)(WG_HALF >> 4) << 16;
            const uint64_t ones_desc = wg_desc(hi_ns, (((base + WG_SM_ONES) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16));
            uint32_t it = 0;
            for (int si = 0; si < n_segs; ++si) {
                const WgSeg sg = segs[si];
                const bool has_main = a.xmain[sg.l] != nullptr, has_enc = a.xenc[sg.l] != nullptr;
                const bool has_small = a.xsmall[sg.l] != nullptr, has_ones = a.ones[sg.l] != 0;
                if (si > 0) { const long long cf = a.prof ? clock64() : 0; tc::mbar_wait(acc_free, (si - 1) & 1, a.wait_hint); tc::tcgen05_fence_after();
                              if (a.prof && lane == 0) a.prof[blockIdx.x * 8 + 4] += clock64() - cf; }
                uint32_t acc = 0;
                for (int n = 2 * (sg.te - sg.tb); n > 0; --n, ++it) {
                    const int slot = it % WG_STAGES; const uint32_t ph = (it / WG_STAGES) & 1;
                    const long long c0 = a.prof ? clock64() : 0;
                    tc::mbar_wait(full(slot), ph, a.wait_hint);
                    const long long c1 = a.prof ? clock64() : 0;
                    if (a.prof && lane == 0) { a.prof[blockIdx.x * 8 + 1] += c1 - c0; a.prof[blockIdx.x * 8 + 3] += 1; }
                    tc::tcgen05_fence_after();
                    if (a.dbg & 1) { if (lane == 0) { tc::mbar_arrive_cluster(tc::mapa(empty(slot), 0)); tc::mbar_arrive_cluster(tc::mapa(empty(slot), 1)); } continue; }
                    const uint32_t st = ((base + slot * WG_STAGE_BYTES) & 0x3FFFFu) >> 4;
                    uint32_t a_lo = st | lbo_half, m_lo = (st + (WG_OFF_B >> 4)) | lbo_half, e_lo = (st + (WG_OFF_E >> 4)) | lbo_half;
                    uint32_t s_lo = (st + (WG_OFF_B >> 4)) | ((256u >> 4) << 16);
                    if (tc::elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) {
                            const uint64_t ad = wg_desc(hi_sw, a_lo + kk * 128);
                            const uint32_t ac = (kk == 0) ? acc : 1u;
                            if (has_main) tc::umma_bf16_2cta(tmem_base, ad, wg_desc(hi_sw, m_lo + kk * 128), idesc_main, ac);
                            if (has_enc) tc::umma_bf16_2cta(tmem_base + 256, ad, wg_desc(hi_sw, e_lo + kk * 128), idesc_enc, ac);
                            else if (has_ones) tc::umma_bf16_2cta(tmem_base + 320, ad, ones_desc, idesc_32, ac);
                            if (has_small) tc::umma_bf16_2cta(tmem_base, ad, wg_desc(hi_ns, s_lo + kk * 32), idesc_32, ac);
                        }
                        tc::umma_commit_2cta(empty(slot));
                    }
                    __syncwarp();
                    acc = 1;
                    if (a.prof && lane == 0) a.prof[blockIdx.x * 8 + 2] += clock64() - c1;
                }
                if (tc::elect_one()) tc::umma_commit_2cta(acc_done);
                __syncwarp();
            }
        } else if (lane == 0 && !leader) {
            // ===================== peer relay: "my halves have landed" -> leader's full barrier =====================
            uint32_t it = 0;
            for (int si = 0; si < n_segs; ++si)
                for (int n = 2 * (segs[si].te - segs[si].tb); n > 0; --n, ++it) {
                    const int slot = it % WG_STAGES; const uint32_t ph = (it / WG_STAGES) & 1;
                    tc::mbar_wait(full(slot), ph, a.wait_hint);
                    tc::mbar_arrive_cluster(tc::mapa(full(slot), 0));
                }
        }
    } else {
        // ===================== flush: this CTA's 128 output features =====================
        const uint32_t q = warp & 3;                                   // TMEM lane quarter of this warp
        const uint32_t row = q * 32 + lane;
        const uint32_t taddr = tmem_base + ((q * 32u) << 16);
        const uint32_t free_remote = tc::mapa(acc_free, 0);
        for (int si = 0; si < n_segs; ++si) {
            const WgSeg sg = segs[si];
            const bool has_main = a.xmain[sg.l] != nullptr, has_enc = a.xenc[sg.l] != nullptr;
            float* o = a.out[sg.l] + (size_t)(rank * 128 + row) * a.ld[sg.l];
            tc::mbar_wait(acc_done, si & 1, a.wait_hint);
            tc::tcgen05_fence_after();
            const long long cfl = a.prof ? clock64() : 0;
            for (int c = 0; c < 336; c += 16) {
                const bool live = c < 256 ? (has_main || (c == 0 && a.xsmall[sg.l])) : (c < 320 ? has_enc : a.ones[sg.l] != 0);
                if (!live) continue;
                uint32_t r[16];
                tc::tmem_ld16(taddr + c, r); tc::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; j += 4)
                    tc::red_add_v4(o + c + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
            }
            tc::tcgen05_fence_before();
            __syncwarp();
            if (a.prof && threadIdx.x == 64) a.prof[blockIdx.x * 8 + 5] += clock64() - cfl;
            if (lane == 0) tc::mbar_arrive_cluster(free_remote);
        }
    }
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    if (warp == 1) { tc::tcgen05_fence_after(); tc::tmem_dealloc2(tmem_base, 512); }
}

int launch_wgrad_tc(WgradArgs& a, cudaStream_t st) {
    NERF_CUDA(nerf_smem_opt_in<g1_wgrad_tc_kernel>(WG_SMEM_BYTES));
    static const uint32_t wait_hint = nerf_env_u32("NERF_TC_WAIT_HINT", 0x989680u);
    static const int dbg = (int)nerf_env_u32("NERF_WG_DBG", 0);
    static const bool want_prof = nerf_env_set("NERF_WG_PROF");
    a.wait_hint = wait_hint; a.dbg = dbg;
    int n_units = 0;
    for (int l = 0; l < WG_NU; ++l) n_units += a.dz[l] != nullptr;
    int pairs = nerf_num_sms() / 2;
    const int64_t work = (int64_t)n_units * a.n_tiles;
    if (work < pairs) pairs = (int)(work > 0 ? work : 1);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(WG_THREADS); cfg.dynamicSmemBytes = WG_SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension; attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    a.prof = nullptr;
    if (want_prof) { cudaMalloc(&a.prof, 2 * pairs * 8 * sizeof(long long)); cudaMemset(a.prof, 0, 2 * pairs * 8 * sizeof(long long)); }
    NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_wgrad_tc_kernel, a));
    NERF_CUDA_CHECK_LAUNCH();
    if (a.prof) {
        cudaDeviceSynchronize();
        static long long h[148 * 8];
        cudaMemcpy(h, a.prof, 2 * pairs * 8 * sizeof(long long), cudaMemcpyDeviceToHost);
        double s0 = 0, s1 = 0, s2 = 0, n = 0, p1 = 0, s4 = 0, s5 = 0;
        for (int i = 0; i < pairs; ++i) { s4 += h[(2 * i) * 8 + 4]; s5 += h[(2 * i) * 8 + 5]; s0 += h[(2 * i) * 8]; p1 += h[(2 * i + 1) * 8]; s1 += h[(2 * i) * 8 + 1]; s2 += h[(2 * i) * 8 + 2]; n += h[(2 * i) * 8 + 3]; }
        fprintf(stderr, "wgrad prof: stages/pair %.0f; per stage cycles: leader producer wait-empty %.0f, peer producer wait-empty %.0f, mma wait-full %.0f, mma issue %.0f\n",
                n / pairs, s0 / n, p1 / n, s1 / n, s2 / n);
        fprintf(stderr, "wgrad prof: per pair cycles: mma wait acc_free %.0f, leader flush (all segs) %.0f\n", s4 / pairs, s5 / pairs);
        cudaFree(a.prof);
    }
    return NERF_OK;
}

// test entry: one unit.  dz [n_tiles][4 slabs], xmain [n_tiles][4 slabs] or null, xenc [n_tiles][1 slab]
// or null, xsmall [n_tiles][4 KB] or null; out [256][ld] fp32 is ACCUMULATED into (zero it first); output
// columns as in WgradArgs.
extern "C" int nerf_tc_wgrad_debug(const void* dz, const void* xmain, const void* xenc, const void* xsmall, int ones,
                                   int n_tiles, float* out, int ld, void* stream) {
    if (!dz || (!xmain && !xenc && !xsmall) || (xmain && xsmall) || !out || n_tiles <= 0 || (ld & 3)) return NERF_ERR_BAD_ARG;
    if (ld < (ones ? 336 : (xenc ? 320 : (xmain ? 256 : 16)))) return NERF_ERR_BAD_ARG;
    WgradArgs a{};
    a.dz[0] = (const uint8_t*)dz; a.xmain[0] = (const uint8_t*)xmain; a.xenc[0] = (const uint8_t*)xenc;
    a.xsmall[0] = (const uint8_t*)xsmall; a.ones[0] = ones;
    a.out[0] = out; a.ld[0] = ld; a.n_tiles = n_tiles;
    return launch_wgrad_tc(a, (cudaStream_t)stream);
}

// all weight gradients of one training step from the workspace the forward / backward kernels filled
int launch_wgrad_tc_units(const uint8_t* ws, const TrainLayout& T, cudaStream_t st) {
    WgradArgs a{};
    const size_t unit_bytes = (size_t)T.tr_tiles * 4 * TC_SLAB_BYTES;
    uint8_t* w = const_cast<uint8_t*>(ws);
    for (int l = 0; l < 10; ++l) {
        a.dz[l] = ws + T.off_dz + l * unit_bytes;
        a.xmain[l] = (l == 0) ? nullptr : ws + T.off_act + (l - 1) * unit_bytes;
        a.xenc[l] = (l == 0 || l == 4) ? ws + T.off_pex : (l == 8 ? ws + T.off_ped : nullptr);
        a.out[l] = reinterpret_cast<float*>(w + T.off_gw) + (size_t)l * 256 * TC_GW_LD; a.ld[l] = TC_GW_LD; a.ones[l] = 1;
    }
    for (int hd = 0; hd < 2; ++hd) {
        a.dz[10 + hd] = ws + T.off_act + (hd == 0 ? 7 : 9) * unit_bytes;
        a.xsmall[10 + hd] = ws + T.off_dh;
        a.out[10 + hd] = reinterpret_cast<float*>(w + T.off_gh) + (size_t)hd * 256 * 16; a.ld[10 + hd] = 16;
    }
    a.n_tiles = T.n_tiles;
    return launch_wgrad_tc(a, st);
}

// ------------------------------------------------------------------------------------------------------------------
// fp32-grade weight gradients (bf16x3 training):   dZ^T X  ~=  dZ_hi^T X_hi + dZ_lo^T X_hi + dZ_hi^T X_lo
// in ONE pass over the interleaved (hi, lo) images (tc_layout.cuh): the four operand parts of a stage are fetched once and
// feed three MMAs per K-step into the same accumulators.  (Three launches of the bf16 GEMM, one per product, read every
// part twice -- 8.1 GB for 5.4 GB of images at 4096 x 64 -- and ran at the HBM roof.)
// A stage covers 32 points: [A: 2 slabs x (hi | lo) x 4 KB] [B: the same of the hidden input, or the small operand (hi | lo)
// x 1 KB] [E: encoding (hi | lo) x 4 KB, leader CTA; the peer's E_hi holds ones, its E_lo zeros] = 40 KB, two or three bulk
// copies; five stages in flight.  Column sums of dZ (bias gradients): the ones see dZ_hi and dZ_lo.
#define W3_STAGES 5
#define W3_PIECE 4096
#define W3_OFF_B (4 * W3_PIECE)
#define W3_OFF_E (8 * W3_PIECE)
#define W3_STAGE_BYTES (10 * W3_PIECE)
#define W3_SM_ONES (W3_STAGES * W3_STAGE_BYTES)
#define W3_SM_BAR (W3_SM_ONES + 1024)
#define W3_SM_SEGS (W3_SM_BAR + 128)
#define W3_SM_TOTAL (W3_SM_SEGS + WG_MAX_SEGS * 16)
#define W3_SMEM_BYTES (W3_SM_TOTAL + 1024)

struct WgradX3Args {
    const uint8_t* dz[WG_NU];        // [n_tiles] hidden images: the M operand (null = absent)
    const uint8_t* xmain[WG_NU];     // [n_tiles] hidden images: hidden input of unit l (null = none)
    const uint8_t* xenc[WG_NU];      // [n_tiles] encoding images (null = none)
    const uint8_t* xsmall[WG_NU];    // [n_tiles] small-operand images (null = none; excludes xmain)
    float* out[WG_NU];
    int ld[WG_NU];
    int ones[WG_NU];
    int n_tiles;
    uint32_t wait_hint;
};
// relative time per tile: the pass is HBM-bound, so the bytes a pair fetches (in KB) plus ~4 KB worth of issue time per copy
__host__ __device__ inline int w3_cost(const WgradX3Args& a, int l) {
    return 64 + 8 + (a.xmain[l] ? 64 + 8 : 0) + (a.xenc[l] ? 16 + 4 : 0) + (a.xsmall[l] ? 4 + 8 : 0);
}

__global__ void __launch_bounds__(WG_THREADS, 1) g1_wgrad_x3_kernel(const WgradX3Args a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* sm = smem_raw + (base - raw);
    const uint32_t bar0 = base + W3_SM_BAR;
    auto full = [&](int s) { return bar0 + 8u * s; };
    auto empty = [&](int s) { return bar0 + 8u * (W3_STAGES + s); };
    const uint32_t acc_done = bar0 + 8u * (2 * W3_STAGES), acc_free = bar0 + 8u * (2 * W3_STAGES + 1);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + W3_SM_BAR + 8 * (2 * W3_STAGES + 2));
    volatile int* n_segs_s = reinterpret_cast<volatile int*>(sm + W3_SM_BAR + 8 * (2 * W3_STAGES + 2) + 4);
    WgSeg* segs = reinterpret_cast<WgSeg*>(sm + W3_SM_SEGS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int n_pairs = gridDim.x >> 1, pair = blockIdx.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < W3_STAGES; ++s) { tc::mbar_init(full(s), leader ? 2 : 1); tc::mbar_init(empty(s), 1); }
        tc::mbar_init(acc_done, 1); tc::mbar_init(acc_free, 8);
        tc::fence_barrier_init();
        int64_t total = 0;                                          // equal-cost contiguous cut of the (unit, tile) list, as above
        for (int l = 0; l < WG_NU; ++l) if (a.dz[l]) total += (int64_t)a.n_tiles * w3_cost(a, l);
        const int64_t lo = total * pair / n_pairs, hi = total * (pair + 1) / n_pairs;
        int n = 0; int64_t pre = 0;
        for (int l = 0; l < WG_NU; ++l) {
            if (!a.dz[l]) continue;
            const int64_t c = w3_cost(a, l);
            const int64_t end = pre + c * a.n_tiles;
            const int64_t s0 = lo > pre ? lo : pre, s1 = hi < end ? hi : end;
            if (s1 > s0) {
                const int tb = (int)((s0 - pre + c - 1) / c), te = (int)((s1 - pre + c - 1) / c);
                if (te > tb && n < WG_MAX_SEGS) { segs[n].l = l; segs[n].tb = tb; segs[n].te = te; ++n; }
            }
            pre = end;
        }
        *n_segs_s = n;
    }
    for (int i = threadIdx.x; i < 256; i += WG_THREADS) reinterpret_cast<uint32_t*>(sm + W3_SM_ONES)[i] = 0x3F803F80u;
    if (!leader)                                                    // the peer's share of the [encoding | ones] operand: ones against hi, zeros against lo
        for (int s = 0; s < W3_STAGES; ++s)
            for (int i = threadIdx.x; i < 2 * W3_PIECE / 4; i += WG_THREADS)
                reinterpret_cast<uint32_t*>(sm + s * W3_STAGE_BYTES + W3_OFF_E)[i] = i < W3_PIECE / 4 ? 0x3F803F80u : 0u;
    tc::fence_proxy_async_smem();
    if (warp == 1) tc::tmem_alloc2(tc::smem_u32((const void*)tmem_slot), 512);
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_segs = *n_segs_s;

    if (warp == 0) {
        // ===================== TMA producer: this CTA's two slabs (hi, lo) of both operands, 32 points a stage =====================
        uint32_t it = 0;
        for (int si = 0; si < n_segs; ++si) {
            const WgSeg sg = segs[si];
            const uint8_t* xm = a.xmain[sg.l]; const uint8_t* xs = a.xsmall[sg.l];
            const uint8_t* xe = leader ? a.xenc[sg.l] : nullptr;
            const uint32_t bytes = 4 * W3_PIECE + (xm ? 4 * W3_PIECE : 0) + (xe ? 2 * W3_PIECE : 0) + (xs ? 2048 : 0);
            const size_t mine = (size_t)rank * 4 * W3_PIECE;                    // slabs 2 rank, 2 rank + 1 inside a quarter
            const uint8_t* pa = a.dz[sg.l] + (size_t)sg.tb * X3_IMG_BYTES + mine;
            const uint8_t* pm = xm ? xm + (size_t)sg.tb * X3_IMG_BYTES + mine : nullptr;
            const uint8_t* pe = xe ? xe + (size_t)sg.tb * X3_ENC_IMG_BYTES : nullptr;
            const uint8_t* ps = xs ? xs + (size_t)sg.tb * X3_DH_IMG_BYTES : nullptr;
            for (int n = 4 * (sg.te - sg.tb); n > 0; --n, ++it) {
                const int slot = it % W3_STAGES; const uint32_t ph = (it / W3_STAGES) & 1;
                tc::mbar_wait(empty(slot), ph ^ 1, a.wait_hint);
                const uint32_t st = base + slot * W3_STAGE_BYTES, fb = full(slot);
                if (tc::elect_one()) {
                    tc::mbar_arrive_expect_tx(fb, bytes);
                    tc::tma_load_1d(st, pa, 4 * W3_PIECE, fb);
                    if (pm) tc::tma_load_1d(st + W3_OFF_B, pm, 4 * W3_PIECE, fb);
                    if (pe) tc::tma_load_1d(st + W3_OFF_E, pe, 2 * W3_PIECE, fb);
                    if (ps) tc::tma_load_1d(st + W3_OFF_B, ps, 2048, fb);
                }
                __syncwarp();
                pa += 8 * W3_PIECE; if (pm) pm += 8 * W3_PIECE; if (pe) pe += 2 * W3_PIECE; if (ps) ps += 2048;   // next quarter (tiles are contiguous)
            }
        }
    } else if (warp == 1) {
        if (leader) {
            // ===================== MMA issuer: per K-step (16 points) three products per accumulator =====================
            constexpr uint32_t idesc_main = tc::umma_idesc_bf16_mn(256, 256), idesc_enc = tc::umma_idesc_bf16_mn(256, 128);
            constexpr uint32_t idesc_32 = tc::umma_idesc_bf16_mn(256, 32);
            constexpr uint32_t hi_sw = (1024u >> 4) | (1u << 14) | (2u << 29), hi_ns = (128u >> 4) | (1u << 14);
            constexpr uint32_t lbo_slab = (uint32_t)(2 * W3_PIECE >> 4) << 16;     // next 64-feature atom = next slab of the stage: hi|lo pieces apart
            constexpr uint32_t LO = W3_PIECE >> 4;                                  // hi -> lo piece, in descriptor units
            const uint64_t ones_desc = wg_desc(hi_ns, (((base + W3_SM_ONES) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16));
            uint32_t it = 0;
            for (int si = 0; si < n_segs; ++si) {
                const WgSeg sg = segs[si];
                const bool has_main = a.xmain[sg.l] != nullptr, has_enc = a.xenc[sg.l] != nullptr;
                const bool has_small = a.xsmall[sg.l] != nullptr, has_ones = a.ones[sg.l] != 0;
                if (si > 0) { tc::mbar_wait(acc_free, (si - 1) & 1, a.wait_hint); tc::tcgen05_fence_after(); }
                uint32_t acc = 0;
                for (int n = 4 * (sg.te - sg.tb); n > 0; --n, ++it) {
                    const int slot = it % W3_STAGES; const uint32_t ph = (it / W3_STAGES) & 1;
                    tc::mbar_wait(full(slot), ph, a.wait_hint);
                    tc::tcgen05_fence_after();
                    const uint32_t st = ((base + slot * W3_STAGE_BYTES) & 0x3FFFFu) >> 4;
                    const uint32_t a_lo = st | lbo_slab, m_lo = (st + (W3_OFF_B >> 4)) | lbo_slab, e_lo = (st + (W3_OFF_E >> 4)) | lbo_slab;
                    const uint32_t s_lo = (st + (W3_OFF_B >> 4)) | ((256u >> 4) << 16);
                    if (tc::elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < 2; ++kk) {
                            const uint64_t ah = wg_desc(hi_sw, a_lo + kk * 128), al = wg_desc(hi_sw, a_lo + LO + kk * 128);
                            const uint32_t ac = (kk == 0) ? acc : 1u;
                            if (has_main) {
                                const uint64_t bh = wg_desc(hi_sw, m_lo + kk * 128), bl = wg_desc(hi_sw, m_lo + LO + kk * 128);
                                tc::umma_bf16_2cta(tmem_base, ah, bh, idesc_main, ac);
                                tc::umma_bf16_2cta(tmem_base, al, bh, idesc_main, 1u);
                                tc::umma_bf16_2cta(tmem_base, ah, bl, idesc_main, 1u);
                            }
                            if (has_enc) {
                                const uint64_t eh = wg_desc(hi_sw, e_lo + kk * 128), el = wg_desc(hi_sw, e_lo + LO + kk * 128);
                                tc::umma_bf16_2cta(tmem_base + 256, ah, eh, idesc_enc, ac);
                                tc::umma_bf16_2cta(tmem_base + 256, al, eh, idesc_enc, 1u);
                                tc::umma_bf16_2cta(tmem_base + 256, ah, el, idesc_enc, 1u);
                            } else if (has_ones) {
                                tc::umma_bf16_2cta(tmem_base + 320, ah, ones_desc, idesc_32, ac);
                                tc::umma_bf16_2cta(tmem_base + 320, al, ones_desc, idesc_32, 1u);
                            }
                            if (has_small) {                             // 16 points = 2 groups of 256 B; the lo part 1 KB on
                                const uint64_t sh = wg_desc(hi_ns, s_lo + kk * 32), sl = wg_desc(hi_ns, s_lo + (1024u >> 4) + kk * 32);
                                tc::umma_bf16_2cta(tmem_base, ah, sh, idesc_32, ac);
                                tc::umma_bf16_2cta(tmem_base, al, sh, idesc_32, 1u);
                                tc::umma_bf16_2cta(tmem_base, ah, sl, idesc_32, 1u);
                            }
                        }
                        tc::umma_commit_2cta(empty(slot));
                    }
                    __syncwarp();
                    acc = 1;
                }
                if (tc::elect_one()) tc::umma_commit_2cta(acc_done);
                __syncwarp();
            }
        } else if (lane == 0) {
            // ===================== peer relay =====================
            uint32_t it = 0;
            for (int si = 0; si < n_segs; ++si)
                for (int n = 4 * (segs[si].te - segs[si].tb); n > 0; --n, ++it) {
                    const int slot = it % W3_STAGES; const uint32_t ph = (it / W3_STAGES) & 1;
                    tc::mbar_wait(full(slot), ph, a.wait_hint);
                    tc::mbar_arrive_cluster(tc::mapa(full(slot), 0));
                }
        }
    } else {
        // ===================== flush: this CTA's 128 output features =====================
        const uint32_t q = warp & 3;
        const uint32_t row = q * 32 + lane;
        const uint32_t taddr = tmem_base + ((q * 32u) << 16);
        const uint32_t free_remote = tc::mapa(acc_free, 0);
        for (int si = 0; si < n_segs; ++si) {
            const WgSeg sg = segs[si];
            const bool has_main = a.xmain[sg.l] != nullptr, has_enc = a.xenc[sg.l] != nullptr;
            float* o = a.out[sg.l] + (size_t)(rank * 128 + row) * a.ld[sg.l];
            tc::mbar_wait(acc_done, si & 1, a.wait_hint);
            tc::tcgen05_fence_after();
            for (int c = 0; c < 336; c += 16) {
                const bool live = c < 256 ? (has_main || (c == 0 && a.xsmall[sg.l])) : (c < 320 ? has_enc : a.ones[sg.l] != 0);
                if (!live) continue;
                uint32_t r[16];
                tc::tmem_ld16(taddr + c, r); tc::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; j += 4)
                    tc::red_add_v4(o + c + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
            }
            tc::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive_cluster(free_remote);
        }
    }
    tc::tcgen05_fence_before();
    tc::cluster_sync_all();
    if (warp == 1) { tc::tcgen05_fence_after(); tc::tmem_dealloc2(tmem_base, 512); }
}

// all weight gradients of one fp32-grade training step from the workspace the x3 forward / data-gradient kernels filled
int launch_wgrad_tc_units_x3(const uint8_t* ws, const X3TrainLayout& X, cudaStream_t st) {
    const TrainLayout& T = X.T;
    const size_t unit_bytes = (size_t)T.tr_tiles * X3_IMG_BYTES;
    uint8_t* w = const_cast<uint8_t*>(ws);
    WgradX3Args a{};
    for (int l = 0; l < 10; ++l) {
        a.dz[l] = ws + X.off_dz3 + l * unit_bytes;
        a.xmain[l] = (l == 0) ? nullptr : ws + X.off_act3 + (l - 1) * unit_bytes;
        a.xenc[l] = (l == 0 || l == 4) ? ws + X.off_pex3 : (l == 8 ? ws + X.off_ped3 : nullptr);
        a.out[l] = reinterpret_cast<float*>(w + T.off_gw) + (size_t)l * 256 * TC_GW_LD; a.ld[l] = TC_GW_LD; a.ones[l] = 1;
    }
    for (int hd = 0; hd < 2; ++hd) {                 // head pseudo-units: n_7 / n_9 against the (d_sigma, d_rgb) images
        a.dz[10 + hd] = ws + X.off_act3 + (hd == 0 ? 7 : 9) * unit_bytes;
        a.xsmall[10 + hd] = ws + X.off_dh3;
        a.out[10 + hd] = reinterpret_cast<float*>(w + T.off_gh) + (size_t)hd * 256 * 16; a.ld[10 + hd] = 16;
    }
    a.n_tiles = T.n_tiles;
    NERF_CUDA(nerf_smem_opt_in<g1_wgrad_x3_kernel>(W3_SMEM_BYTES));
    static const uint32_t wait_hint = nerf_env_u32("NERF_TC_WAIT_HINT", 0x989680u);
    a.wait_hint = wait_hint;
    int pairs = nerf_num_sms() / 2;
    const int64_t work = 12ll * a.n_tiles;
    if (work < pairs) pairs = (int)(work > 0 ? work : 1);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(WG_THREADS); cfg.dynamicSmemBytes = W3_SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension; attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_wgrad_x3_kernel, a));
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
