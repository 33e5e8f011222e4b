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
// Work = 20 (unit, out-half) combinations x n_tiles point tiles, cut into 148 equal-cost
// contiguous segments (one per persistent CTA).  A CTA accumulates its segment's 128 x K_l block
// in TMEM over all its tiles and flushes it once with vector fp32 reductions into dW'_l.
//   warp 0   TMA producer (1-D bulk copies of half slabs: 64 points per pipeline stage)
//   warp 1   MMA issuer   (M=128, N=256 hidden + N=64 encoding columns, K=16 points per MMA)
//   warps 2-5 flush       (tcgen05.ld -> red.global.add.v4.f32)
// The kernel is HBM-bound: 32 KB of dZ + 64..80 KB of activations per 128-point tile and half.
#include "tc_layout.cuh"
#include <stdlib.h>

#define WG_THREADS 192
#define WG_STAGES 4
#define WG_HALF 8192                      // half a slab: 64 points x 128 B
#define WG_STAGE_BYTES (7 * WG_HALF)      // 2 dZ + 4 hidden + 1 encoding half slabs
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
};
__host__ __device__ inline int wg_cost(const WgradArgs& a, int l) {      // quarter half-slabs loaded per tile half
    return 8 + (a.xmain[l] ? 16 : 0) + (a.xenc[l] ? 4 : 0) + (a.xsmall[l] ? 1 : 0);
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

struct WgSeg { int l, h, tb, te; };

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

    if (threadIdx.x == 0) {
        for (int s = 0; s < WG_STAGES; ++s) { tc::mbar_init(full(s), 1); tc::mbar_init(empty(s), 1); }
        tc::mbar_init(acc_done, 1); tc::mbar_init(acc_free, 128);
        tc::fence_barrier_init();
        // equal-cost contiguous cut of the (unit, half, tile) work list
        int64_t total = 0;
        for (int l = 0; l < WG_NU; ++l) if (a.dz[l]) total += 2ll * a.n_tiles * wg_cost(a, l);
        const int64_t lo = total * blockIdx.x / gridDim.x, hi = total * (blockIdx.x + 1) / gridDim.x;
        int n = 0; int64_t pre = 0;
        for (int l = 0; l < WG_NU; ++l) {
            if (!a.dz[l]) continue;
            const int64_t c = wg_cost(a, l);
            for (int h = 0; h < 2; ++h) {
                const int64_t end = pre + c * a.n_tiles;
                const int64_t s0 = lo > pre ? lo : pre, s1 = hi < end ? hi : end;
                if (s1 > s0) {
                    const int tb = (int)((s0 - pre + c - 1) / c), te = (int)((s1 - pre + c - 1) / c);
                    if (te > tb && n < WG_MAX_SEGS) { segs[n].l = l; segs[n].h = h; segs[n].tb = tb; segs[n].te = te; ++n; }
                }
                pre = end;
            }
        }
        *n_segs_s = n;
    }
    for (int i = threadIdx.x; i < 256; i += WG_THREADS) reinterpret_cast<uint32_t*>(sm + WG_SM_ONES)[i] = 0x3F803F80u;
    tc::fence_proxy_async_smem();
    if (warp == 1) tc::tmem_alloc(tc::smem_u32((const void*)tmem_slot), 512);
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_segs = *n_segs_s;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int si = 0; si < n_segs; ++si) {
                const WgSeg sg = segs[si];
                const uint8_t* dz = a.dz[sg.l]; const uint8_t* xm = a.xmain[sg.l]; const uint8_t* xe = a.xenc[sg.l];
                const uint8_t* xs = a.xsmall[sg.l];
                const uint32_t bytes = (2 + (xm ? 4 : 0) + (xe ? 1 : 0)) * WG_HALF + (xs ? 2048 : 0);
                for (int t = sg.tb; t < sg.te; ++t)
                    for (int hf = 0; hf < 2; ++hf, ++it) {
                        const int slot = it % WG_STAGES; const uint32_t ph = (it / WG_STAGES) & 1;
                        tc::mbar_wait(empty(slot), ph ^ 1, a.wait_hint);
                        tc::mbar_arrive_expect_tx(full(slot), bytes);
                        const uint32_t st = base + slot * WG_STAGE_BYTES;
                        const size_t toff = (size_t)t * 4 * TC_SLAB_BYTES + hf * WG_HALF;
                        tc::tma_load_1d(st, dz + toff + (size_t)(2 * sg.h) * TC_SLAB_BYTES, WG_HALF, full(slot));
                        tc::tma_load_1d(st + WG_HALF, dz + toff + (size_t)(2 * sg.h + 1) * TC_SLAB_BYTES, WG_HALF, full(slot));
                        if (xm)
                            for (int s = 0; s < 4; ++s)
                                tc::tma_load_1d(st + (2 + s) * WG_HALF, xm + toff + (size_t)s * TC_SLAB_BYTES, WG_HALF, full(slot));
                        if (xe) tc::tma_load_1d(st + 6 * WG_HALF, xe + (size_t)t * TC_SLAB_BYTES + hf * WG_HALF, WG_HALF, full(slot));
                        if (xs) tc::tma_load_1d(st + 2 * WG_HALF, xs + (size_t)t * 4096 + hf * 2048, 2048, full(slot));
                    }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_main = tc::umma_idesc_bf16_mn(128, 256), idesc_enc = tc::umma_idesc_bf16_mn(128, 64);
            constexpr uint32_t idesc_16 = tc::umma_idesc_bf16_mn(128, 16);
            const uint64_t ones_desc = tc::umma_desc_noswz_mn(base + WG_SM_ONES, 128, 128);
            uint32_t it = 0;
            for (int si = 0; si < n_segs; ++si) {
                const WgSeg sg = segs[si];
                const bool has_main = a.xmain[sg.l] != nullptr, has_enc = a.xenc[sg.l] != nullptr;
                const bool has_small = a.xsmall[sg.l] != nullptr, has_ones = a.ones[sg.l] != 0;
                if (si > 0) { tc::mbar_wait(acc_free, (si - 1) & 1, a.wait_hint); tc::tcgen05_fence_after(); }
                bool first = true;
                for (int t = sg.tb; t < sg.te; ++t)
                    for (int hf = 0; hf < 2; ++hf, ++it) {
                        const int slot = it % WG_STAGES; const uint32_t ph = (it / WG_STAGES) & 1;
                        tc::mbar_wait(full(slot), ph, a.wait_hint);
                        tc::tcgen05_fence_after();
                        const uint32_t st = base + slot * WG_STAGE_BYTES;
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) {
                            const uint64_t ad = tc::umma_desc_sw128_mn(st + kk * 2048, WG_HALF, 1024);
                            if (has_main)
                                tc::umma_bf16(tmem_base, ad, tc::umma_desc_sw128_mn(st + 2 * WG_HALF + kk * 2048, WG_HALF, 1024),
                                              idesc_main, first ? 0u : 1u);
                            if (has_enc)
                                tc::umma_bf16(tmem_base + 256, ad,
                                              tc::umma_desc_sw128_mn(st + 6 * WG_HALF + kk * 2048, WG_HALF, 1024), idesc_enc, first ? 0u : 1u);
                            if (has_ones) tc::umma_bf16(tmem_base + 320, ad, ones_desc, idesc_16, first ? 0u : 1u);
                            if (has_small)
                                tc::umma_bf16(tmem_base, ad, tc::umma_desc_noswz_mn(st + 2 * WG_HALF + kk * 512, 256, 128), idesc_16, first ? 0u : 1u);
                            first = false;
                        }
                        tc::umma_commit(empty(slot));
                    }
                tc::umma_commit(acc_done);
            }
        }
    } else {
        const uint32_t q = warp & 3;                                   // TMEM lane quarter of this warp
        const uint32_t row = q * 32 + lane;
        const uint32_t taddr = tmem_base + ((q * 32u) << 16);
        for (int si = 0; si < n_segs; ++si) {
            const WgSeg sg = segs[si];
            const bool has_main = a.xmain[sg.l] != nullptr, has_enc = a.xenc[sg.l] != nullptr;
            float* o = a.out[sg.l] + (size_t)(sg.h * 128 + row) * a.ld[sg.l];
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
            tc::mbar_arrive(acc_free);
        }
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) { tc::tcgen05_fence_after(); tc::tmem_dealloc(tmem_base, 512); }
}

int launch_wgrad_tc(WgradArgs& a, cudaStream_t st) {
    static bool init = false;
    if (!init) {
        NERF_CUDA(cudaFuncSetAttribute(g1_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_BYTES));
        init = true;
    }
    { const char* e = getenv("NERF_TC_WAIT_HINT"); a.wait_hint = e ? (uint32_t)strtoul(e, nullptr, 0) : 0x989680u; }
    int grid = nerf_num_sms();
    const int64_t work = 2ll * WG_NU * a.n_tiles;
    if (work < grid) grid = (int)(work > 0 ? work : 1);
    g1_wgrad_tc_kernel<<<grid, WG_THREADS, WG_SMEM_BYTES, st>>>(a);
    NERF_CUDA_CHECK_LAUNCH();
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
