// tcgen05 fused path: kernel body (roles) and the C entry points built on it.
#include "tc_fused_kernel.cuh"
#include <stdlib.h>
#include <stdio.h>

// CG = 1: one CTA per SM works alone (M=128 MMAs, both weight halves per K-slab).
// CG = 2: CTA pairs (cluster of 2, cta_group::2): M=256 MMAs issued by the leader over both CTAs'
//         tiles; each CTA streams only ITS half (128 output features) of every weight K-slab, so
//         the weight bytes per tile and SM halve -- the v1 profile showed the kernel bound by the
//         weight ring's latency x depth, not by the tensor pipe or the epilogue.
template <int CG, bool TRAIN>
__global__ void __launch_bounds__(TC_THREADS, 1) g1_fused_tc_kernel(const FusedArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = tc::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;                 // SW128 slabs need 1024 B alignment
    uint8_t* sm = smem_raw + (base - raw);
    float* vec_s = reinterpret_cast<float*>(sm + SM_VEC);
    const uint32_t bar0 = base + SM_BAR;
    // barriers: ring_full[R], ring_empty[R], act_ready[2], acc_ready[2]; then the TMEM base slot
    auto ring_full = [&](int s) { return bar0 + 8u * s; };
    auto ring_empty = [&](int s) { return bar0 + 8u * (TC_RING + s); };
    auto act_ready = [&](int x) { return bar0 + 8u * (2 * TC_RING + x); };
    auto acc_ready = [&](int x) { return bar0 + 8u * (2 * TC_RING + 2 + x); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sm + SM_BAR + 8 * (2 * TC_RING + 4));
    // TRAIN: act_local[2] (this CTA's slab of slot x is complete: 4 epilogue warps), store_done[2] (its bulk store has been read out)
    auto act_local = [&](int x) { return bar0 + 8u * (2 * TC_RING + 5 + x); };
    auto store_done = [&](int x) { return bar0 + 8u * (2 * TC_RING + 7 + x); };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = (CG == 2) ? tc::cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    // work units: CG=1 -> tiles; CG=2 -> tile pairs (this CTA takes tile 2*pair + rank)
    const int G = gridDim.x / CG, me = blockIdx.x / CG;
    const int n_units = (CG == 2) ? (a.n_tiles + 1) / 2 : a.n_tiles;
    const int n_my = (me < n_units) ? (n_units - me + G - 1) / G : 0;
    const int cnt0 = (n_my + 1) >> 1, cnt1 = n_my >> 1;           // units of slot 0 / slot 1
    const int stages_per_slab = (CG == 2) ? 1 : 2;

    for (int i = threadIdx.x; i < TC_SMVEC_FLOATS; i += TC_THREADS) vec_s[i] = a.vec[TC_VEC_HEADW + i];
    if (warp < 4) {                                               // sum_k w_k of head `warp` (the affine part of the head dot products)
        float sw = 0.f;
        for (int i = lane; i < 256; i += 32) sw += a.vec[TC_VEC_HEADW + warp * 256 + i];
        sw = warp_sum(sw);
        if (lane == 0) reinterpret_cast<float*>(sm + SM_HSUM)[warp] = sw;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_RING; ++s) { tc::mbar_init(ring_full(s), (CG == 2 && leader) ? 2 : 1); tc::mbar_init(ring_empty(s), 1); }
        for (int x = 0; x < 2; ++x) { tc::mbar_init(act_ready(x), CG == 2 ? 8 : 128); tc::mbar_init(acc_ready(x), 1); tc::mbar_init(act_local(x), 4); tc::mbar_init(store_done(x), 1); }
        tc::fence_barrier_init();
    }
    if (warp == 1) { if (CG == 2) tc::tmem_alloc2(tc::smem_u32((const void*)tmem_slot), 512); else tc::tmem_alloc(tc::smem_u32((const void*)tmem_slot), 512); }
    tc::tcgen05_fence_before();
    if (CG == 2) tc::cluster_sync_all(); else __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // Role loops are latency chains of one warp (~6 cycles per dependent instruction, ~110 per mbarrier
    // test, ~250 per bulk-copy issue, ~45 per MMA issue; scripts/mma_issue.cu, scripts/tma_bw.cu), and the
    // tensor pipe only hides ~256 cycles of issue bubble: everything per stage is kept to a handful of
    // 32-bit adds on precomputed descriptor words.
    if (warp == 0) {
        // ===================== TMA producer (whole warp converged; one elected lane issues) =====================
        uint32_t it = 0;
        for (int q = 0; q < cnt0; ++q)
            for (int l = 0; l < 10; ++l)
                for (int x = 0; x < 2; ++x) {
                    if (x == 1 && q >= cnt1) continue;
                    const int blk_kc = tc_unit_has_bias_block(l) ? TF_BIAS_KCORES : 0;
                    const int nk = a.nk[l];
                    for (int h = 0; h < stages_per_slab; ++h) {           // CG=1 streams the two n-halves one after the other
                        const int nh = (CG == 2) ? (int)rank : h;         // pair: this CTA owns one feature half
                        const uint8_t* src = a.stages + a.unit_off[l] + (size_t)nh * a.half_kcores[l] * TF_KCORE_BYTES;
                        for (int j = 0; j < nk; ++j, ++it) {
                            const int slot = it % TC_RING; const uint32_t ph = (it / TC_RING) & 1;
                            const uint32_t bytes = TF_STAGE_BYTES + (j == 0 ? blk_kc * TF_KCORE_BYTES : 0);   // the bias block rides in front of stage 0
                            tc::mbar_wait_spin(ring_empty(slot), ph ^ 1, a.wait_hint);
                            if (tc::elect_one()) {
                                tc::mbar_arrive_expect_tx(ring_full(slot), bytes);
                                tc::tma_load_1d(base + SM_RING + slot * TF_SLOT_BYTES, src, bytes, ring_full(slot));
                            }
                            __syncwarp();
                            src += bytes;
                        }
                    }
                }
    } else if (warp == 1) {
        if (leader) {
            // ===================== MMA issuer (whole warp converged; one elected lane issues) =====================
            const uint32_t idesc = (CG == 2) ? tc::umma_idesc_bf16(256, 256) : tc::umma_idesc_bf16(128, 128);
            // descriptor words: hi = SBO | version | layout, lo = start address | LBO
            constexpr uint32_t hi_a = (1024u >> 4) | (1u << 14) | (2u << 29);             // SW128 K-major slabs
            constexpr uint32_t hi_b = (TF_B_SBO >> 4) | (1u << 14);                       // no-swizzle k-cores
            auto lo_a = [](uint32_t addr) { return ((addr & 0x3FFFFu) >> 4) | (1u << 16); };
            const uint32_t ring_lo = (((base + SM_RING) & 0x3FFFFu) >> 4) | ((TF_B_LBO >> 4) << 16);
            auto mma = [&](uint32_t d, uint32_t alo, uint32_t blo, uint32_t acc) {
                const uint64_t ad = ((uint64_t)hi_a << 32) | alo, bd = ((uint64_t)hi_b << 32) | blo;
                if (CG == 2) tc::umma_bf16_2cta(d, ad, bd, idesc, acc); else tc::umma_bf16(d, ad, bd, idesc, acc);
            };
            uint32_t it = 0, act_it[2] = {0, 0};
            for (int q = 0; q < cnt0; ++q)
                for (int l = 0; l < 10; ++l)
                    for (int x = 0; x < 2; ++x) {
                        if (x == 1 && q >= cnt1) continue;
                        if (a.prof && blockIdx.x == 0 && q == 1 && lane == 0) a.prof[(l * 2 + x) * 8 + 0] = clock64();
                        tc::mbar_wait_spin(act_ready(x), act_it[x] & 1, a.wait_hint); ++act_it[x];
                        tc::tcgen05_fence_after();
                        if (a.prof && blockIdx.x == 0 && q == 1 && lane == 0) a.prof[(l * 2 + x) * 8 + 1] = clock64();
                        const uint32_t act_lo = lo_a(base + SM_ACT + x * 65536), pe_lo = lo_a(base + SM_PE + x * TC_SLAB_BYTES);
                        const bool blk = tc_unit_has_bias_block(l);
                        const int nk = a.nk[l];
                        for (int h = 0; h < stages_per_slab; ++h) {
                            const uint32_t d_tmem = tmem_base + x * 256 + ((CG == 2) ? 0 : h * 128);
                            for (int j = 0; j < nk; ++j, ++it) {
                                const int slot = it % TC_RING; const uint32_t ph = (it / TC_RING) & 1;
                                const uint32_t alo = (l == 0 || j >= 4) ? pe_lo : act_lo + j * (TC_SLAB_BYTES >> 4);
                                uint32_t blo = ring_lo + slot * (TF_SLOT_BYTES >> 4);
                                tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);   // TMA data -> MMA: both async proxy, no tcgen05 fence
                                if (tc::elect_one()) {
                                    if (blk && j == 0) {                  // acc = b': slab columns 48..63 (ones at 60..62) x bias block
                                        mma(d_tmem, pe_lo + 3 * 2, blo, 0);
                                        blo += (TF_BIAS_KCORES * TF_KCORE_BYTES) >> 4;
                                    }
                                    const uint32_t acc0 = (blk || j != 0) ? 1u : 0u;
                                    mma(d_tmem, alo, blo, acc0);
                                    mma(d_tmem, alo + 2, blo + 1 * (2 * TF_KCORE_BYTES >> 4), 1);
                                    mma(d_tmem, alo + 4, blo + 2 * (2 * TF_KCORE_BYTES >> 4), 1);
                                    mma(d_tmem, alo + 6, blo + 3 * (2 * TF_KCORE_BYTES >> 4), 1);
                                    if (CG == 2) tc::umma_commit_2cta(ring_empty(slot)); else tc::umma_commit(ring_empty(slot));
                                    if (j == nk - 1 && h == stages_per_slab - 1) {
                                        if (CG == 2) tc::umma_commit_2cta(acc_ready(x)); else tc::umma_commit(acc_ready(x));
                                    }
                                }
                                __syncwarp();
                            }
                        }
                        if (a.prof && blockIdx.x == 0 && q == 1 && lane == 0) a.prof[(l * 2 + x) * 8 + 2] = clock64();
                    }
        } else if (CG == 2 && lane == 0 && !leader) {
            // ===================== peer relay: "my weight half has landed" -> leader's full barrier =====================
            uint32_t it = 0;
            for (int q = 0; q < cnt0; ++q)
                for (int l = 0; l < 10; ++l)
                    for (int x = 0; x < 2; ++x) {
                        if (x == 1 && q >= cnt1) continue;
                        for (int j = 0; j < a.nk[l]; ++j, ++it) {
                            const int slot = it % TC_RING; const uint32_t ph = (it / TC_RING) & 1;
                            tc::mbar_wait_spin(ring_full(slot), ph, a.wait_hint);
                            tc::mbar_arrive_cluster(tc::mapa(ring_full(slot), 0));
                        }
                    }
        }
    } else if (warp == 10) {
        // ===================== TRAIN: activation-store thread =====================
        // The normalised output of unit l is the A operand of unit l+1's MMAs AND a slab image the backward needs.  One 64 KB
        // bulk store issued by the epilogue right before it releases the MMA sits in front of the weight-stage loads in the
        // SM's bulk-copy queue (forward timeline: the slot whose MMAs follow the stores took 5.5 k cycles per unit instead of
        // 2.2 k); this thread spreads the store as four 16 KB pieces over the MMA's duration instead and tells the epilogue
        // when the slab may be overwritten.
        if (TRAIN && lane == 0) {
            uint32_t al_it[2] = {0, 0};
            for (int q = 0; q < cnt0; ++q)
                // (every act_local phase is followed, in the epilogue, by a wait on the store_done it leads to before the next one can
                // complete: two phases can never pass unobserved, whatever the timing)
                for (int sgn = 1; sgn < 10; ++sgn)                     // signal l+1: n_l ready
                    for (int x = 0; x < 2; ++x) {
                        if (x == 1 && q >= cnt1) continue;
                        tc::mbar_wait(act_local(x), al_it[x] & 1, a.wait_hint); ++al_it[x];
                        const int unit = me + (2 * q + x) * G;
                        const int tile = (CG == 2) ? 2 * unit + (int)rank : unit;
                        uint8_t* dst = a.tr_act + ((size_t)(sgn - 1) * a.tr_tiles + tile) * (4 * TC_SLAB_BYTES);
                        const uint32_t src = base + SM_ACT + x * 65536;
                        for (int ks = 0; ks < 4; ++ks) {
                            tc::tma_store_1d(dst + (size_t)ks * TC_SLAB_BYTES, src + ks * TC_SLAB_BYTES, TC_SLAB_BYTES);
                            tc::tma_store_commit();
                            if (ks < 3) { const long long t0 = clock64(); while (clock64() - t0 < (long long)a.store_gap) { } }
                        }
                        tc::tma_store_wait_read0();
                        tc::mbar_arrive(store_done(x));
                    }
            tc::tma_store_wait_all0();
        }
    } else {
        // ===================== encoder + epilogue groups =====================
        const int x = (warp - 2) >> 2;                               // tile slot of this group
        const uint32_t row = (warp & 3) * 32 + lane;                 // TMEM lane quarter = warp % 4
        const uint32_t taddr = tmem_base + (((warp & 3) * 32u) << 16) + x * 256;
        const uint32_t act_addr = base + SM_ACT + x * 65536, pe_addr = base + SM_PE + x * TC_SLAB_BYTES;
        const int cnt = x == 0 ? cnt0 : cnt1;
        const uint32_t act_bar_remote = (CG == 2) ? tc::mapa(act_ready(x), 0) : 0u;
        auto signal_act = [&](bool slab_to_store = true) {
            tc::fence_proxy_async_smem();
            tc::tcgen05_fence_before();
            if (CG == 2 || TRAIN) __syncwarp();
            if (TRAIN && slab_to_store && lane == 0) tc::mbar_arrive(act_local(x));
            if (CG == 2) { if (lane == 0) tc::mbar_arrive_cluster(act_bar_remote); }
            else tc::mbar_arrive(act_ready(x));
        };
        uint32_t acc_it = 0, sd_it = 0;
        for (int q = 0; q < cnt; ++q) {
            const int unit = me + (2 * q + x) * G;
            const int tile = (CG == 2) ? 2 * unit + (int)rank : unit;
            RowState st;
            encode_row(a, tile, row, pe_addr, st);
            if (TRAIN) {                                             // position-encoding slab -> HBM (wgrad of units 0, 4)
                tc::fence_proxy_async_smem(); tc::named_bar_sync(1 + x, 128);
                if (row == 0) { tc::tma_store_1d(a.tr_pex + (size_t)tile * TC_SLAB_BYTES, pe_addr, TC_SLAB_BYTES); tc::tma_store_commit(); }
            }
            signal_act(false);                                       // the encoding slab: nothing for the store thread
            for (int l = 0; l < 10; ++l) {
                if (a.prof && blockIdx.x == 0 && q == 1 && threadIdx.x == 64 + x * 128) a.prof[(l * 2 + x) * 8 + 3] = clock64();
                tc::mbar_wait(acc_ready(x), acc_it & 1, a.wait_hint); ++acc_it;
                tc::tcgen05_fence_after();
                if (a.prof && blockIdx.x == 0 && q == 1 && threadIdx.x == 64 + x * 128) a.prof[(l * 2 + x) * 8 + 4] = clock64();
                if (TRAIN) {                                         // pending bulk stores must have read the slabs we overwrite
                    if (row == 0) tc::tma_store_wait_read0();            // (this thread's: encodings, n_9 of the previous tile)
                    tc::named_bar_sync(1 + x, 128);
                    if (l >= 1) { tc::mbar_wait(store_done(x), sd_it & 1, a.wait_hint); ++sd_it; }   // n_{l-1}: the store thread's
                }
                if (!TRAIN && a.dbg && a.dbg_layer == l) epilogue_unit<3, false>(a, l, taddr, act_addr, row, base + SM_VEC, st);
                else if (l == 7) epilogue_unit<1, TRAIN>(a, l, taddr, act_addr, row, base + SM_VEC, st);
                else if (l == 9) epilogue_unit<2, TRAIN>(a, l, taddr, act_addr, row, base + SM_VEC, st);
                else if (a.ray_bias && (l == 0 || l == 4))
                    epilogue_unit<0, TRAIN, true>(a, l, taddr, act_addr, row, base + SM_VEC, st,
                                                  a.ray_bias + ((size_t)(l ? 1 : 0) * a.rb_rows + (st.valid ? (a.rb_ray0 + st.p / a.S) / a.rb_div : 0)) * 256);
                else epilogue_unit<0, TRAIN>(a, l, taddr, act_addr, row, base + SM_VEC, st);
                if (l == 4) encode_dir_row(a, row, pe_addr, st);     // unit 4 was the last reader of the position slab
                if (TRAIN && (l == 4 || l == 9)) {                   // direction slab, and n_9 (no MMA follows it) -> HBM from here;
                    tc::fence_proxy_async_smem(); tc::named_bar_sync(1 + x, 128);   // n_0..n_8 leave through the store thread
                    if (row == 0) {
                        if (l == 9) tc::tma_store_1d(a.tr_act + ((size_t)l * a.tr_tiles + tile) * (4 * TC_SLAB_BYTES), act_addr, 4 * TC_SLAB_BYTES);
                        if (l == 4) tc::tma_store_1d(a.tr_ped + (size_t)tile * TC_SLAB_BYTES, pe_addr, TC_SLAB_BYTES);
                        tc::tma_store_commit();
                    }
                }
                if (a.prof && blockIdx.x == 0 && q == 1 && threadIdx.x == 64 + x * 128) a.prof[(l * 2 + x) * 8 + 5] = clock64();
                if (l < 9) signal_act();
            }
        }
        if (TRAIN && row == 0) tc::tma_store_wait_all0();
    }
    tc::tcgen05_fence_before();
    if (CG == 2) tc::cluster_sync_all(); else __syncthreads();
    if (warp == 1) { tc::tcgen05_fence_after(); if (CG == 2) tc::tmem_dealloc2(tmem_base, 512); else tc::tmem_dealloc(tmem_base, 512); }
}

// ------------------------------------------------------------------ host side ----
static int launch_fused(const void* packed, int Dx, int Dd, FusedArgs& a, cudaStream_t st, bool train = false) {
    PackLayout L; int rc = make_pack_layout(Dx, Dd, L); if (rc) return rc;
    if (Dd != 3 || (Dx != 3 && !a.ray_bias)) return NERF_ERR_UNSUPPORTED;   // states_x need the per-ray bias; states_d: fp32 path
    for (int l = 0; l < 10; ++l) { a.nk[l] = L.nk[l]; a.half_kcores[l] = L.half_kcores[l]; a.unit_off[l] = L.unit_off[l]; }
    a.stages = (const uint8_t*)packed;
    a.vec = reinterpret_cast<const float*>((const uint8_t*)packed + tc_packed_vec_offset(L));
    static const uint32_t wait_hint = nerf_env_u32("NERF_TC_WAIT_HINT", 0x989680u);
    static const uint32_t store_gap = nerf_env_u32("NERF_FWD_STORE_GAP", 600u);   // measured: 0 -> 0.410, 400-800 -> 0.389-0.393, 1200 -> 0.442 ms
    static const int cg = nerf_env_u32("NERF_TC_CG", 2) == 1 ? 1 : 2;
    static const bool prof = nerf_env_set("NERF_TC_PROF");
    a.wait_hint = wait_hint; a.store_gap = store_gap;
    if (cg == 2) NERF_CUDA(train ? (nerf_smem_opt_in<g1_fused_tc_kernel<2, true>>(TC_SMEM_BYTES)) : (nerf_smem_opt_in<g1_fused_tc_kernel<2, false>>(TC_SMEM_BYTES)));
    else NERF_CUDA(train ? (nerf_smem_opt_in<g1_fused_tc_kernel<1, true>>(TC_SMEM_BYTES)) : (nerf_smem_opt_in<g1_fused_tc_kernel<1, false>>(TC_SMEM_BYTES)));
    a.prof = nullptr;
    if (prof) { NERF_CUDA(cudaMalloc(&a.prof, 192 * sizeof(long long))); NERF_CUDA(cudaMemset(a.prof, 0, 192 * sizeof(long long))); }
    const int sms = nerf_num_sms();
    if (cg == 2) {
        int pairs = (a.n_tiles + 1) / 2, clusters = pairs < sms / 2 ? pairs : sms / 2;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(TC_THREADS); cfg.dynamicSmemBytes = TC_SMEM_BYTES; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension; attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        if (train) NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_fused_tc_kernel<2, true>, a));
        else NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_fused_tc_kernel<2, false>, a));
    } else {
        int grid = a.n_tiles < sms ? a.n_tiles : sms;
        if (train) g1_fused_tc_kernel<1, true><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(a);
        else g1_fused_tc_kernel<1, false><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(a);
    }
    NERF_CUDA_CHECK_LAUNCH();
    if (a.prof) {                                                 // debug: print the stamps relative to the first one
        long long h[192];
        NERF_CUDA(cudaStreamSynchronize(st));
        NERF_CUDA(cudaMemcpy(h, a.prof, sizeof(h), cudaMemcpyDeviceToHost));
        long long t0 = h[0];
        for (int i = 0; i < 20; ++i)
            fprintf(stderr, "unit %d slot %d: mma wait %lld..%lld issued %lld | epi wait %lld..%lld done %lld\n", i / 2, i % 2,
                    h[i * 8] - t0, h[i * 8 + 1] - t0, h[i * 8 + 2] - t0, h[i * 8 + 3] - t0, h[i * 8 + 4] - t0, h[i * 8 + 5] - t0);
        cudaFree(a.prof);
    }
    return NERF_OK;
}

extern "C" int nerf_add_density_noise(float*, int, int64_t, float, uint64_t, int64_t, void*);

static inline int64_t a256(int64_t x) { return (x + 255) & ~(int64_t)255; }

int launch_state_bias(const void* packed, int Dx, int Dd, const float* states, int64_t n_states, float* pes, float* rb, cudaStream_t st);   // tc_train.cu

extern "C" int64_t nerf_g1_render_workspace_bytes(int64_t B, int S, int Dx, int64_t n_states) {
    if (B < 0 || S <= 0 || Dx < 3 || Dx > 16 || n_states < 0) return NERF_ERR_BAD_ARG;
    int64_t P = B * S;
    int64_t n = a256(P * 4) * 2 + a256(P * 12) + 2 * a256(B * 12);   // t, sigma, rgb, rays_o, rays_d
    if (Dx > 3) n += a256(n_states * (Dx - 3) * 2 * NERF_PE_X_OCTAVES * 4) + a256(n_states * 2 * 256 * 4);   // PE(states), bias rows of units 0 and 4
    return n;
}

int launch_fused_x3(const void* packed, FusedArgs& a, cudaStream_t st, bool train = false);      // fused_tc_x3.cu

static int render_fused_impl(bool x3, const void* packed, int Dx, int Dd,
                                    const float* rays_o, const float* rays_d,
                                    const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                                    float focal, int64_t ray_begin, int64_t B, int S,
                                    float near_plane, float far_plane, const float* t_base, const float* t_in,
                                    const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                                    const float* noise, float noise_std,
                                    float normalize_position, int normalize_direction,
                                    const float* states_x, int64_t n_states, int64_t rays_per_state,
                                    float* out_pixels, float* out_depth, float* out_acc, float* out_weights,
                                    float* out_t, float* out_sigma, float* out_rgb,
                                    void* workspace, int64_t workspace_bytes, void* stream) {
    if (!packed || B < 0 || S < 32 || S > 4096) return (S > 0 && S < 32) ? NERF_ERR_UNSUPPORTED : NERF_ERR_BAD_ARG;
    if (!rays_o != !rays_d || Dx < 3 || Dx > 16) return NERF_ERR_BAD_ARG;
    if (x3 && (Dx != 3 || Dd != 3)) return NERF_ERR_UNSUPPORTED;
    if (B == 0) return NERF_OK;                                   // zero rays: nothing to do (empty tensors carry null pointers)
    if (Dx == 3) n_states = 0;
    else if (!states_x || n_states <= 0 || rays_per_state <= 0 ||
             ((rays_o ? 0 : ray_begin) + B - 1) / rays_per_state >= n_states) return NERF_ERR_BAD_ARG;
    if (!rays_o && (!cam_o || !cam_r || n_cams <= 0 || H <= 0 || W <= 0 || ray_begin < 0 ||
                    ray_begin + B > (int64_t)n_cams * H * W)) return NERF_ERR_BAD_ARG;
    if (!workspace || workspace_bytes < nerf_g1_render_workspace_bytes(B, S, Dx, n_states)) return NERF_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t P = B * S;
    uint8_t* w = (uint8_t*)workspace;
    float* ws_t = (float*)w; w += a256(P * 4);
    float* ws_sigma = (float*)w; w += a256(P * 4);
    float* ws_rgb = (float*)w; w += a256(P * 12);
    float* ws_o = (float*)w; w += a256(B * 12);
    float* ws_d = (float*)w; w += a256(B * 12);
    FusedArgs a{};
    if (Dx > 3) {                                                 // states -> PE -> bias rows of units 0 and 4 (tc_layout.cuh)
        float* pes = (float*)w; w += a256(n_states * (Dx - 3) * 2 * NERF_PE_X_OCTAVES * 4);
        float* rb = (float*)w;
        int rc = launch_state_bias(packed, Dx, Dd, states_x, n_states, pes, rb, st); if (rc) return rc;
        a.ray_bias = rb; a.rb_rows = n_states; a.rb_ray0 = rays_o ? 0 : ray_begin; a.rb_div = rays_per_state;
    }
    a.rays_o = rays_o; a.rays_d = rays_d; a.cam_o = cam_o; a.cam_r = cam_r; a.H = H; a.W = W; a.focal = focal;
    a.ray_begin = ray_begin; a.ws_rays_o = ws_o; a.ws_rays_d = ws_d;
    a.B = B; a.S = S; a.P = P; a.n_tiles = (int)((P + 127) / 128);
    a.t_in = t_in; a.t_base = t_base; a.u = u; a.jitter = (jitter || u) ? 1 : 0; a.seed = seed; a.ray_id_base = ray_id_base;
    a.near_plane = near_plane; a.far_plane = far_plane; a.norm = normalize_position; a.normalize_dir = normalize_direction;
    a.out_t = out_t ? out_t : ws_t; a.out_sigma = out_sigma ? out_sigma : ws_sigma; a.out_rgb = out_rgb ? out_rgb : ws_rgb;
    a.dbg = nullptr; a.dbg_layer = -1;
    int rc = x3 ? launch_fused_x3(packed, a, st) : launch_fused(packed, Dx, Dd, a, st); if (rc) return rc;
    if (noise) return NERF_ERR_UNSUPPORTED;   // host-supplied noise is added by the binding on out_sigma before compositing
    if (noise_std != 0.0f) { rc = nerf_add_density_noise(a.out_sigma, S, B, noise_std, seed, ray_id_base, stream); if (rc) return rc; }
    if (out_pixels || out_depth || out_acc || out_weights) {
        const float* ro = rays_o ? rays_o : ws_o; const float* rd = rays_o ? rays_d : ws_d;
        rc = nerf_composite_fwd(a.out_sigma, a.out_rgb, nullptr, ro, rd, a.out_t, B, S, out_pixels, out_depth, out_acc,
                                out_weights, stream);
        if (rc) return rc;
    }
    return NERF_OK;
}

extern "C" int nerf_g1_render_fused(const void* packed, int Dx, int Dd, const float* rays_o, const float* rays_d,
                                    const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                                    float focal, int64_t ray_begin, int64_t B, int S,
                                    float near_plane, float far_plane, const float* t_base, const float* t_in,
                                    const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                                    const float* noise, float noise_std, float normalize_position, int normalize_direction,
                                    const float* states_x, int64_t n_states, int64_t rays_per_state,
                                    float* out_pixels, float* out_depth, float* out_acc, float* out_weights,
                                    float* out_t, float* out_sigma, float* out_rgb,
                                    void* workspace, int64_t workspace_bytes, void* stream) {
    return render_fused_impl(false, packed, Dx, Dd, rays_o, rays_d, cam_o, cam_r, n_cams, H, W, focal, ray_begin, B, S, near_plane, far_plane,
                             t_base, t_in, u, jitter, seed, ray_id_base, noise, noise_std, normalize_position, normalize_direction,
                             states_x, n_states, rays_per_state, out_pixels, out_depth, out_acc, out_weights, out_t, out_sigma, out_rgb,
                             workspace, workspace_bytes, stream);
}

// The same pass with fp32-grade operands (three bf16 products per term, fused_tc_x3.cu); `packed` from nerf_g1_pack_weights_x3.
extern "C" int nerf_g1_render_fused_x3(const void* packed, const float* rays_o, const float* rays_d,
                                       const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                                       float focal, int64_t ray_begin, int64_t B, int S,
                                       float near_plane, float far_plane, const float* t_base, const float* t_in,
                                       const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                                       float noise_std, float normalize_position, int normalize_direction,
                                       float* out_pixels, float* out_depth, float* out_acc, float* out_weights,
                                       float* out_t, float* out_sigma, float* out_rgb,
                                       void* workspace, int64_t workspace_bytes, void* stream) {
    return render_fused_impl(true, packed, 3, 3, rays_o, rays_d, cam_o, cam_r, n_cams, H, W, focal, ray_begin, B, S, near_plane, far_plane,
                             t_base, t_in, u, jitter, seed, ray_id_base, nullptr, noise_std, normalize_position, normalize_direction,
                             nullptr, 0, 1, out_pixels, out_depth, out_acc, out_weights, out_t, out_sigma, out_rgb,
                             workspace, workspace_bytes, stream);
}

// test/debug entry: the MLP alone on explicit rays + t, optionally dumping the normalised
// output of one unit (fp32, before bf16 rounding) -> per-layer parity against the oracle.
extern "C" int nerf_g1_mlp_tc_debug(const void* packed, const float* rays_o, const float* rays_d, const float* t,
                                    int64_t B, int S, float normalize_position, int normalize_direction,
                                    float* out_sigma, float* out_rgb, float* scratch_t, float* dbg, int dbg_layer,
                                    void* stream) {
    if (!packed || !rays_o || !rays_d || !t || !out_sigma || !out_rgb || !scratch_t || B <= 0 || S < 32) return NERF_ERR_BAD_ARG;
    FusedArgs a{};
    a.rays_o = rays_o; a.rays_d = rays_d; a.B = B; a.S = S; a.P = B * S; a.n_tiles = (int)((a.P + 127) / 128);
    a.t_in = t; a.norm = normalize_position; a.normalize_dir = normalize_direction;
    a.out_t = scratch_t; a.out_sigma = out_sigma; a.out_rgb = out_rgb; a.dbg = dbg; a.dbg_layer = dbg_layer;
    return launch_fused(packed, 3, 3, a, (cudaStream_t)stream);
}

// training forward: the fused kernel with TRAIN = true, storing what the backward needs into `ws`
int launch_fused_train(const void* packed, int Dx, int Dd, const float* rays_o, const float* rays_d, int64_t B, int S,
                       float near_plane, float far_plane, const float* t_base, const float* t_in, const float* u, int jitter,
                       uint64_t seed, int64_t ray_id_base, float norm, int normalize_dir,
                       float* out_t, float* out_sigma, float* out_rgb, uint8_t* ws, const TrainLayout& T, cudaStream_t st) {
    FusedArgs a{};
    a.ray_bias = T.EXs ? reinterpret_cast<const float*>(ws + T.off_rb) : nullptr;
    a.rb_rows = B; a.rb_ray0 = 0; a.rb_div = 1;
    a.rays_o = rays_o; a.rays_d = rays_d; a.B = B; a.S = S; a.P = T.P; a.n_tiles = T.n_tiles;
    a.t_in = t_in; a.t_base = t_base; a.u = u; a.jitter = (jitter || u) ? 1 : 0; a.seed = seed; a.ray_id_base = ray_id_base;
    a.near_plane = near_plane; a.far_plane = far_plane; a.norm = norm; a.normalize_dir = normalize_dir;
    a.out_t = out_t; a.out_sigma = out_sigma; a.out_rgb = out_rgb; a.dbg = nullptr; a.dbg_layer = -1;
    a.tr_act = ws + T.off_act; a.tr_pex = ws + T.off_pex; a.tr_ped = ws + T.off_ped;
    a.tr_stats = reinterpret_cast<float2*>(ws + T.off_stats); a.P_pad = T.P_pad; a.tr_tiles = T.tr_tiles;
    return launch_fused(packed, Dx, Dd, a, st, true);
}

// ------------------------------------------------------------------ no-allocation variant ----
// nerf_g1_render_fused for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): the scratch the fused render needs (t, sigma, rgb per point; the generated rays per ray) lives in a
// file-scope device array instead of a caller workspace.  No counterpart in the reference.
// Every GPU that loads the library holds the array (2.6 MB) until exit.
//   * limit: B * S <= NERF_STATIC_MAX_POINTS (131072 points, e.g. 2048 rays x 64 samples), unconditional model (Dx = Dd = 3);
//   * the entry point calls no cudaMalloc / cudaFree / allocating routine: packing, launch and composite are the existing
//     stream-ordered kernels; cudaGetSymbolAddress (once per device) and cudaFuncSetAttribute allocate nothing;
//   * every call shares the array and finds the previous call's contents in it.  Nothing has to start at zero: the MLP kernel
//     writes t / sigma / rgb of every valid point (and the rays of every ray in camera mode) before the composite reads them,
//     and neither kernel reads beyond B * S points;
//   * calls that overlap on the GPU (two streams, or two graphs replayed concurrently) RACE on the array: serialise them.
#define NERF_STATIC_MAX_POINTS 131072
#define NERF_STATIC_MAX_RAYS (NERF_STATIC_MAX_POINTS / 32)          // S >= 32
#define NERF_STATIC_SCRATCH_BYTES (NERF_STATIC_MAX_POINTS * 20 + 2 * ((NERF_STATIC_MAX_RAYS * 12 + 255) & ~255))
static __device__ __align__(256) unsigned char g_render_scratch[NERF_STATIC_SCRATCH_BYTES];

// In host code `g_render_scratch` is a host-side stand-in, not a device address: the device address is asked for once per
// device, with that device current, and cached.
static int render_scratch_address(void** out) {
    static void* addr[NERF_MAX_DEVICES] = {};
    const int dev = nerf_current_device();
    if (dev < 0 || dev >= NERF_MAX_DEVICES) return nerf_set_error_text(NERF_ERR_UNSUPPORTED, "nerf_g1_render_fused_static: device index beyond NERF_MAX_DEVICES");
    if (!addr[dev]) {
        void* p = nullptr;
        NERF_CUDA(cudaGetSymbolAddress(&p, g_render_scratch));
        if (!p) return nerf_set_error_text(NERF_ERR_WORKSPACE, "nerf_g1_render_fused_static: cudaGetSymbolAddress returned a null address");
        addr[dev] = p;
    }
    *out = addr[dev];
    return NERF_OK;
}

extern "C" int64_t nerf_g1_render_static_max_points(void) { return NERF_STATIC_MAX_POINTS; }

extern "C" int nerf_g1_render_fused_static(const void* packed, const float* rays_o, const float* rays_d,
                                           const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                                           float focal, int64_t ray_begin, int64_t B, int S,
                                           float near_plane, float far_plane, const float* t_base,
                                           const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                                           float noise_std, float normalize_position, int normalize_direction,
                                           float* out_pixels, float* out_depth, float* out_acc, void* stream) {
    if (B < 0 || S <= 0) return NERF_ERR_BAD_ARG;
    if (B * (int64_t)S > NERF_STATIC_MAX_POINTS || nerf_g1_render_workspace_bytes(B, S, 3, 0) > (int64_t)NERF_STATIC_SCRATCH_BYTES)
        return nerf_set_error_text(NERF_ERR_WORKSPACE, "nerf_g1_render_fused_static: B * S exceeds the static scratch (131072 points): "
                                                       "split the call or use nerf_g1_render_fused with a caller workspace");
    void* scratch = nullptr;
    int rc = render_scratch_address(&scratch); if (rc) return rc;
    return render_fused_impl(false, packed, 3, 3, rays_o, rays_d, cam_o, cam_r, n_cams, H, W, focal, ray_begin, B, S, near_plane, far_plane,
                             t_base, nullptr, u, jitter, seed, ray_id_base, nullptr, noise_std, normalize_position, normalize_direction,
                             nullptr, 0, 1, out_pixels, out_depth, out_acc, nullptr, nullptr, nullptr, nullptr,
                             scratch, NERF_STATIC_SCRATCH_BYTES, stream);
}
