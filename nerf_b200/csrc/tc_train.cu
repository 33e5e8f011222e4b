// bf16 training step on tcgen05: host entry points around the three tensor-core kernels
//   g1_fused_tc_kernel<CG, true>  forward, stores slab images of every unit's output   (fused_tc.cu)
//   g1_bwd_tc_kernel<CG>          data-gradient chain + LayerNorm/ReLU backward         (tc_bwd_kernel.cuh)
//   g1_wgrad_tc_kernel            dW' = dZ^T X, db' = dZ^T 1, head gradients            (tc_wgrad.cu)
// plus the transposed weight packing for the backward MMAs and the "unfold" kernels that turn
// gradients w.r.t. the folded parameters (W' = W diag(gamma_prev), b' = b + W beta_prev) into the
// gradients of the 44 state_dict tensors (reference: autograd of train_conditional_nerf.py:120-135).
#include "tc_bwd_kernel.cuh"
#include <stdio.h>
#include <stdlib.h>

struct WgradArgs;
int launch_wgrad_tc_units(const uint8_t* ws, const TrainLayout& T, cudaStream_t st);
int launch_fused_train(const void* packed, int Dx, int Dd, const float* rays_o, const float* rays_d, int64_t B, int S,
                       float near_plane, float far_plane, const float* t_base, const float* t_in, const float* u, int jitter,
                       uint64_t seed, int64_t ray_id_base, float norm, int normalize_dir,
                       float* out_t, float* out_sigma, float* out_rgb, uint8_t* ws, const TrainLayout& T, cudaStream_t st);

// ---- transposed weight stages for g_{l-1} = dz_l W'_l: stage (l-1)*8 + ks*2 + nh holds
//      [128 rows r = input feature nh*128 + r] x [64 cols c = output feature ks*64 + c], K-major SW128
__global__ void pack_bwd_stage_kernel(const float* __restrict__ params, PackLayout L, uint8_t* __restrict__ packed) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 72ll * 128 * 8) return;
    const int c16 = (int)(i & 7), r = (int)((i >> 3) & 127), stage = (int)(i >> 10);
    const int l = 1 + stage / 8, rel = stage % 8, ks = rel >> 1, nh = rel & 1;
    const int kin = nh * 128 + r;
    const float* W = params + L.off_W[l];
    const float g = params[L.off_gamma[l - 1] + kin];
    const int Kt = L.Kt[l];
    uint32_t out[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int n0 = ks * 64 + c16 * 8 + q * 2;
        out[q] = tc::pack_bf16x2(W[(size_t)n0 * Kt + kin] * g, W[(size_t)(n0 + 1) * Kt + kin] * g);
    }
    *reinterpret_cast<uint4*>(packed + (size_t)stage * TC_STAGE_BYTES + tc::sw128_offset(r, c16)) = make_uint4(out[0], out[1], out[2], out[3]);
}

extern "C" int64_t nerf_g1_packed_bwd_bytes(int Dx, int Dd) {
    PackLayout L;
    if (make_pack_layout(Dx, Dd, L) != NERF_OK) return NERF_ERR_BAD_ARG;
    return 72ll * TC_STAGE_BYTES;
}

extern "C" int nerf_g1_pack_weights_bwd(const float* params, int Dx, int Dd, void* packed_bwd, void* stream) {
    if (!params || !packed_bwd) return NERF_ERR_BAD_ARG;
    PackLayout L; int rc = make_pack_layout(Dx, Dd, L); if (rc) return rc;
    const int64_t n = 72ll * 128 * 8;
    pack_bwd_stage_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(params, L, (uint8_t*)packed_bwd);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int64_t nerf_g1_train_workspace_bytes(int64_t B, int S, int Dx) {
    if (B <= 0 || S <= 0 || Dx < 3 || Dx > 16) return NERF_ERR_BAD_ARG;
    TrainLayout T; make_train_layout(B, S, T, Dx);
    return T.total;
}

// ---- conditional inputs: the states' share of units 0 and 4 as a per-ray bias (tc_layout.cuh) ----
// pes[ray][c*20 + j] = sin(pi 2^j s_c), [c*20 + 10 + j] = cos(pi 2^j s_c): the layout of nerf_encode_g1 for c >= 3
__global__ void pe_states_kernel(const float* __restrict__ states, int64_t B, int Sx, float* __restrict__ pes) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * Sx * NERF_PE_X_OCTAVES) return;
    const int j = (int)(i % NERF_PE_X_OCTAVES); const int64_t rc = i / NERF_PE_X_OCTAVES;
    const int c = (int)(rc % Sx); const int64_t ray = rc / Sx;
    float s, co; sincospif(ldexpf(states[ray * Sx + c], j), &s, &co);
    float* o = pes + ray * (2 * NERF_PE_X_OCTAVES * Sx) + c * 2 * NERF_PE_X_OCTAVES;
    o[j] = s; o[NERF_PE_X_OCTAVES + j] = co;
}
// rb[u][ray][n] = sum_k sw[u][k][n] pes[ray][k]   (block = RB_RAYS rays, thread = n: a weight is loaded once for the block's
// rays -- a block per ray re-read the 80 KB weight block 8192 times, 34 us per step; the sum over k keeps its order)
#define RB_RAYS 16
__global__ void __launch_bounds__(256) ray_bias_kernel(const float* __restrict__ sw, const float* __restrict__ pes, int64_t B, int EXs,
                                                      float* __restrict__ rb) {
    extern __shared__ float pe_s[];                                  // [RB_RAYS][EXs]
    const int64_t r0 = (int64_t)blockIdx.x * RB_RAYS; const int u = blockIdx.y, n = threadIdx.x;
    const int nr = (int)(B - r0 < RB_RAYS ? B - r0 : RB_RAYS);
    for (int i = threadIdx.x; i < RB_RAYS * EXs; i += 256) pe_s[i] = (i / EXs) < nr ? pes[r0 * EXs + i] : 0.f;
    __syncthreads();
    const float* w = sw + (size_t)u * EXs * 256 + n;
    float acc[RB_RAYS];
#pragma unroll
    for (int r = 0; r < RB_RAYS; ++r) acc[r] = 0.f;
#pragma unroll 4
    for (int k = 0; k < EXs; ++k) {
        const float wk = w[(size_t)k * 256];
#pragma unroll
        for (int r = 0; r < RB_RAYS; ++r) acc[r] = fmaf(wk, pe_s[r * EXs + k], acc[r]);
    }
#pragma unroll
    for (int r = 0; r < RB_RAYS; ++r) if (r < nr) rb[((size_t)u * B + r0 + r) * 256 + n] = acc[r];
}
// rsum[u][ray][n] = sum over the ray's S points of dz_l[p][n], l = 0 (u = 0) or 4 (u = 1), read from the slab image.
// Block = ray; thread = (point group pg of 8, 16-byte chunk c of the row's 32): 128-byte segments per warp-load.
__global__ void __launch_bounds__(256) dz_ray_sum_kernel(const uint8_t* __restrict__ dz, int tr_tiles, int64_t B, int S, float* __restrict__ rsum) {
    __shared__ float part[8][256];
    const int64_t ray = blockIdx.x; const int u = blockIdx.y;
    const int pg = threadIdx.x >> 5, c = threadIdx.x & 31;
    const uint8_t* img = dz + (size_t)(u ? 4 : 0) * tr_tiles * (4 * TC_SLAB_BYTES);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int k = pg; k < S; k += 8) {
        const int64_t p = ray * S + k; const int64_t tile = p >> 7; const uint32_t row = (uint32_t)(p & 127);
        const uint4 v = *reinterpret_cast<const uint4*>(img + (size_t)tile * (4 * TC_SLAB_BYTES) + (size_t)(c >> 3) * TC_SLAB_BYTES + tc::sw128_offset(row, c & 7));
        acc[0] += bf_lo(v.x); acc[1] += bf_hi(v.x); acc[2] += bf_lo(v.y); acc[3] += bf_hi(v.y);
        acc[4] += bf_lo(v.z); acc[5] += bf_hi(v.z); acc[6] += bf_lo(v.w); acc[7] += bf_hi(v.w);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) part[pg][c * 8 + j] = acc[j];
    __syncthreads();
    const int n = threadIdx.x;
    float t = 0.f;
#pragma unroll
    for (int g = 0; g < 8; ++g) t += part[g][n];
    rsum[((size_t)u * B + ray) * 256 + n] = t;
}
// gs[u][n][k] += sum_ray rsum[u][ray][n] pes[ray][k]: block = (chunk of SG_RAYS rays, u, chunk of KC columns), thread = n holding
// its KC partial sums in registers (one coalesced load + KC FMAs per ray, eight loads in flight), EXs <= 20 * 13.  One atomic
// per output element and block: with 64 rays a block the 2.6 M contended atomics were most of the kernel's 48 us.
#define SG_RAYS 128
template <int KC>
__global__ void __launch_bounds__(256) state_grad_kernel(const float* __restrict__ rsum, const float* __restrict__ pes, int64_t B, int EXs,
                                                        float* __restrict__ gs) {
    __shared__ float pe_s[64][KC];
    const int u = blockIdx.y, n = threadIdx.x, k0 = blockIdx.z * KC;
    float acc[KC];
#pragma unroll
    for (int k = 0; k < KC; ++k) acc[k] = 0.f;
    const int64_t b0 = (int64_t)blockIdx.x * SG_RAYS, b1 = b0 + SG_RAYS < B ? b0 + SG_RAYS : B;
    for (int64_t r0 = b0; r0 < b1; r0 += 64) {
        const int nr = (int)(b1 - r0 < 64 ? b1 - r0 : 64);
        __syncthreads();
        for (int i = threadIdx.x; i < 64 * KC; i += 256) { const int r = i / KC, k = i % KC; pe_s[r][k] = (r < nr && k0 + k < EXs) ? pes[(r0 + r) * EXs + k0 + k] : 0.f; }
        __syncthreads();
        const float* src = rsum + ((size_t)u * B + r0) * 256 + n;
        for (int r = 0; r < 64; r += 8) {
            float v[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = r + q < nr ? src[(size_t)(r + q) * 256] : 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q)
#pragma unroll
                for (int k = 0; k < KC; ++k) acc[k] = fmaf(v[q], pe_s[r + q][k], acc[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < KC; ++k) if (k0 + k < EXs) atomicAdd(gs + ((size_t)u * 256 + n) * EXs + k0 + k, acc[k]);
}
// ---- unfold: gradients of folded parameters -> gradients of the state_dict tensors ----
// gw [10][256][336]: G_l = dL/dW'_l (cols 0..255 hidden, 256..319 encoding), col 320 = c_l = dL/db'_l
// gh [2][256][16]:   gh[0][k][0] = sum_p d_sigma n_7[k];  gh[1][k][1+c] = sum_p d_rgb_c n_9[k]
// hsum [4]:          sum_p (d_sigma, d_rgb)
//   dW_l[i,k] = G_l[i,k] gamma_{l-1}[k] + c_l[i] beta_{l-1}[k]  (hidden k);  = G_l[i, 256+k2]  (encoding)
//   db_l = c_l;  heads: dw[k] = gamma[k] H[k] + beta[k] s,  db = s
//   conditional inputs: the state columns of units 0 and 4 come from gs (state_grad_kernel)
__device__ __forceinline__ void unfold_weights_body(const float* __restrict__ params, const PackLayout& L, const float* __restrict__ gw,
                                                    const float* __restrict__ gh, const float* __restrict__ hsum, const float* __restrict__ gs,
                                                    float* __restrict__ grads) {
    const int l = blockIdx.y;
    if (l < 10) {
        const int Kt = L.Kt[l], hid = (l == 0) ? 0 : 256;
        const float* G = gw + (size_t)l * 256 * TC_GW_LD;
        const uint32_t row_len = (uint32_t)Kt + 1u, n_el = 256u * row_len;       // 32-bit index math: the kernel is a few elements per thread
        for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n_el; e += gridDim.x * blockDim.x) {
            const int i = (int)(e / row_len), k = (int)(e - (uint32_t)i * row_len);
            const float c = G[(size_t)i * TC_GW_LD + 320];
            if (k == Kt) { grads[L.off_b[l] + i] = c; continue; }
            float v;
            if (k < hid) v = fmaf(G[(size_t)i * TC_GW_LD + k], params[L.off_gamma[l - 1] + k], c * params[L.off_beta[l - 1] + k]);
            else if (l != 8 && k - hid >= 2 * NERF_PE_X_OCTAVES * 3) v = gs[((size_t)(l ? 1 : 0) * 256 + i) * L.EXs + (k - hid - 2 * NERF_PE_X_OCTAVES * 3)];
            else v = G[(size_t)i * TC_GW_LD + 256 + (k - hid)];
            grads[L.off_W[l] + (size_t)i * Kt + k] = v;
        }
    } else {
        for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < 4 * 256 + 4; e += gridDim.x * blockDim.x) {
            if (e >= 4 * 256) {
                const int c = e - 4 * 256;
                if (c == 0) grads[L.off_head_b[0]] = hsum[0]; else grads[L.off_head_b[1] + c - 1] = hsum[c];
                continue;
            }
            const int c = e >> 8, k = e & 255, unit = c == 0 ? 7 : 9;
            const float H = gh[(size_t)(c == 0 ? 0 : 1) * 256 * 16 + k * 16 + c];
            const float v = fmaf(params[L.off_gamma[unit] + k], H, params[L.off_beta[unit] + k] * hsum[c]);
            if (c == 0) grads[L.off_head_w[0] + k] = v; else grads[L.off_head_w[1] + (size_t)(c - 1) * 256 + k] = v;
        }
    }
}
//   dgamma_j[k] = sum_i G_{j+1}[i,k] W_{j+1}[i,k]  (+ head: w[k] H[k]);  dbeta_j[k] = sum_i c_{j+1}[i] W_{j+1}[i,k]  (+ head: w[k] s)
__device__ __forceinline__ void unfold_affine_body(const float* __restrict__ params, const PackLayout& L, const float* __restrict__ gw,
                                                   const float* __restrict__ gh, const float* __restrict__ hsum, float* __restrict__ grads) {
    const int j = blockIdx.x, k = threadIdx.x & 255, part = threadIdx.x >> 8;      // 1024 threads: 4 row groups x 256 columns
    __shared__ float sg[4][256], sb[4][256];
    float dg = 0.f, db = 0.f;
    if (j < 9) {
        const int l = j + 1, Kt = L.Kt[l];
        const float* G = gw + (size_t)l * 256 * TC_GW_LD;
        const float* W = params + L.off_W[l];
        float dgs[8], dbs[8];                                       // 8 rows in flight: the loop is load-latency bound
#pragma unroll
        for (int q = 0; q < 8; ++q) dgs[q] = dbs[q] = 0.f;
        for (int i = part * 64; i < part * 64 + 64; i += 8) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float w = W[(size_t)(i + q) * Kt + k];
                dgs[q] = fmaf(G[(size_t)(i + q) * TC_GW_LD + k], w, dgs[q]);
                dbs[q] = fmaf(G[(size_t)(i + q) * TC_GW_LD + 320], w, dbs[q]);
            }
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) { dg += dgs[q]; db += dbs[q]; }
    }
    sg[part][k] = dg; sb[part][k] = db;
    __syncthreads();
    if (part == 0) {
        dg = sg[0][k] + sg[1][k] + sg[2][k] + sg[3][k]; db = sb[0][k] + sb[1][k] + sb[2][k] + sb[3][k];
        if (j == 7) {
            const float w = params[L.off_head_w[0] + k];
            dg = fmaf(w, gh[k * 16], dg); db = fmaf(w, hsum[0], db);
        } else if (j == 9) {
            for (int c = 0; c < 3; ++c) {
                const float w = params[L.off_head_w[1] + (size_t)c * 256 + k];
                dg = fmaf(w, gh[256 * 16 + k * 16 + 1 + c], dg); db = fmaf(w, hsum[1 + c], db);
            }
        }
        grads[L.off_gamma[j] + k] = dg; grads[L.off_beta[j] + k] = db;
    }
}

// one launch: rows 0..10 of the grid unfold the weights (grid-stride), row 11 the LayerNorm affines (block = unit)
__global__ void __launch_bounds__(1024) unfold_kernel(const float* __restrict__ params, PackLayout L, const float* __restrict__ gw,
                                                      const float* __restrict__ gh, const float* __restrict__ hsum,
                                                      const float* __restrict__ gs, float* __restrict__ grads) {
    if (blockIdx.y < 11) unfold_weights_body(params, L, gw, gh, hsum, gs, grads);
    else if (blockIdx.x < 10) unfold_affine_body(params, L, gw, gh, hsum, grads);
}

template <int CG>
static int launch_bwd(BwdArgs& a, cudaStream_t st) {
    a.prof = nullptr;
    static const bool want_prof = nerf_env_set("NERF_TC_PROF");     // debug timeline: allocates and synchronises, never on by default
    if (want_prof) { NERF_CUDA(cudaMalloc(&a.prof, 192 * sizeof(long long))); NERF_CUDA(cudaMemset(a.prof, 0, 192 * sizeof(long long))); }
    NERF_CUDA(nerf_smem_opt_in<g1_bwd_tc_kernel<CG>>(TB_SMEM_BYTES));
    const int sms = nerf_num_sms();
    if (CG == 2) {
        int pairs = (a.n_tiles + 1) / 2, clusters = pairs < sms / 2 ? pairs : sms / 2;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(TB_THREADS); cfg.dynamicSmemBytes = TB_SMEM_BYTES; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension; attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        NERF_CUDA(cudaLaunchKernelEx(&cfg, g1_bwd_tc_kernel<CG>, a));
    } else {
        int grid = a.n_tiles < sms ? a.n_tiles : sms;
        g1_bwd_tc_kernel<CG><<<grid, TB_THREADS, TB_SMEM_BYTES, st>>>(a);
    }
    NERF_CUDA_CHECK_LAUNCH();
    if (a.prof) {                                                 // debug: stamps relative to the first MMA wait of unit 9
        long long h[192];
        NERF_CUDA(cudaStreamSynchronize(st));
        NERF_CUDA(cudaMemcpy(h, a.prof, sizeof(h), cudaMemcpyDeviceToHost));
        const long long t0 = h[9 * 2 * 8 + 3];
        for (int x = 0; x < 2; ++x)
            fprintf(stderr, "bwd unit 5 slot %d: n_full seen at %lld %lld %lld %lld, pass 1 done %lld\n", x, h[160 + x * 8] - t0, h[161 + x * 8] - t0,
                    h[162 + x * 8] - t0, h[163 + x * 8] - t0, h[164 + x * 8] - t0);
        for (int x = 0; x < 2; ++x)
            fprintf(stderr, "bwd unit 6->5 slot %d io thread: refill loop entered %lld, slab 0 free %lld, dz store read out %lld, slab 3 free %lld\n", x,
                    h[176 + x * 4] - t0, h[177 + x * 4] - t0, h[178 + x * 4] - t0, h[179 + x * 4] - t0);
        for (int l = 9; l >= 0; --l)
            for (int x = 0; x < 2; ++x) {
                const long long* r = h + (l * 2 + x) * 8;
                fprintf(stderr, "bwd unit %d slot %d: mma wait %lld..%lld issued %lld | epi wait %lld..%lld math done %lld synced %lld\n", l, x,
                        r[0] - t0, r[1] - t0, r[2] - t0, r[3] - t0, r[4] - t0, r[5] - t0, r[6] - t0);
            }
        cudaFree(a.prof);
    }
    return NERF_OK;
}

// pes [n_states][EXs] <- PE(states [n_states][Dx-3]); rb [2][n_states][256] <- state weight columns of units 0, 4 x pes
int launch_state_bias(const void* packed, int Dx, int Dd, const float* states, int64_t n_states, float* pes, float* rb, cudaStream_t st) {
    PackLayout L; int rc = make_pack_layout(Dx, Dd, L); if (rc) return rc;
    const int Sx = Dx - 3;
    const int64_t n = n_states * Sx * NERF_PE_X_OCTAVES;
    pe_states_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(states, n_states, Sx, pes);
    NERF_CUDA_CHECK_LAUNCH();
    ray_bias_kernel<<<dim3((unsigned)((n_states + RB_RAYS - 1) / RB_RAYS), 2), 256, (size_t)RB_RAYS * L.EXs * sizeof(float), st>>>(
        reinterpret_cast<const float*>((const uint8_t*)packed + tc_packed_state_w_offset(L)), pes, n_states, L.EXs, rb);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int nerf_g1_train_fwd_tc(const void* packed, int Dx, int Dd, const float* rays_o, const float* rays_d,
                                    int64_t B, int S, float near_plane, float far_plane, const float* t_base,
                                    const float* t_in, const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                                    float normalize_position, int normalize_direction, const float* states_x,
                                    float* out_t, float* out_sigma, float* out_rgb,
                                    void* workspace, int64_t workspace_bytes, void* stream) {
    if (!packed || !rays_o || !rays_d || !out_t || !out_sigma || !out_rgb || B <= 0 || S < 32) return NERF_ERR_BAD_ARG;
    if (Dx < 3 || Dx > 16 || (Dx > 3 && !states_x)) return NERF_ERR_BAD_ARG;
    if (Dd != 3) return NERF_ERR_UNSUPPORTED;
    PackLayout L; int rc = make_pack_layout(Dx, Dd, L); if (rc) return rc;
    TrainLayout T; make_train_layout(B, S, T, Dx);
    if (!workspace || workspace_bytes < T.total) return NERF_ERR_WORKSPACE;
    if (T.EXs) {                                                  // states -> PE -> per-ray bias of units 0 and 4
        uint8_t* ws = (uint8_t*)workspace;
        rc = launch_state_bias(packed, Dx, Dd, states_x, B, reinterpret_cast<float*>(ws + T.off_pes),
                               reinterpret_cast<float*>(ws + T.off_rb), (cudaStream_t)stream);
        if (rc) return rc;
    }
    return launch_fused_train(packed, Dx, Dd, rays_o, rays_d, B, S, near_plane, far_plane, t_base, t_in, u, jitter, seed,
                              ray_id_base, normalize_position, normalize_direction, out_t, out_sigma, out_rgb,
                              (uint8_t*)workspace, T, (cudaStream_t)stream);
}

// stage: 0 = all, 1 = stop after the data-gradient kernel, 2 = stop after the weight-gradient GEMM (tests),
//        3 = the weight-gradient GEMM alone on whatever the workspace holds (timing)
extern "C" int nerf_g1_train_bwd_tc(const void* packed, const void* packed_bwd, const float* params, int Dx, int Dd,
                                    const float* d_sigma, const float* d_rgb, int64_t B, int S, float* grads,
                                    void* workspace, int64_t workspace_bytes, int stage, void* stream) {
    if (!packed || !packed_bwd || !params || !d_sigma || !d_rgb || !grads || B <= 0 || S < 32) return NERF_ERR_BAD_ARG;
    if (Dd != 3) return NERF_ERR_UNSUPPORTED;
    PackLayout L; int rc = make_pack_layout(Dx, Dd, L); if (rc) return rc;
    TrainLayout T; make_train_layout(B, S, T, Dx);
    if (!workspace || workspace_bytes < T.total) return NERF_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* ws = (uint8_t*)workspace;
    NERF_CUDA(cudaMemsetAsync(ws + T.zero_begin, 0, (size_t)T.zero_bytes, st));
    BwdArgs a{};
    a.stages = (const uint8_t*)packed_bwd;
    a.vec = reinterpret_cast<const float*>((const uint8_t*)packed + tc_packed_vec_offset(L));
    a.act = ws + T.off_act; a.stats = reinterpret_cast<const float2*>(ws + T.off_stats);
    a.d_sigma = d_sigma; a.d_rgb = d_rgb; a.dz = ws + T.off_dz; a.dh = ws + T.off_dh;
    a.hsum = reinterpret_cast<float*>(ws + T.off_hsum);
    a.P = T.P; a.P_pad = T.P_pad; a.n_tiles = T.n_tiles; a.tr_tiles = T.tr_tiles;
    static const uint32_t wait_hint = nerf_env_u32("NERF_TC_WAIT_HINT", 0x989680u);
    static const int pf = (int)nerf_env_u32("NERF_BWD_PF", 11);
    static const int cg = nerf_env_u32("NERF_TC_CG", 2) == 1 ? 1 : 2;
    a.wait_hint = wait_hint; a.pf_dist = pf % 10; a.pf_next = pf >= 10;
    if (stage != 3) { rc = (cg == 2) ? launch_bwd<2>(a, st) : launch_bwd<1>(a, st); if (rc) return rc; }
    if (stage == 1) return NERF_OK;
    rc = launch_wgrad_tc_units(ws, T, st); if (rc) return rc;
    if (stage == 2 || stage == 3) return NERF_OK;
    const float* gw = reinterpret_cast<const float*>(ws + T.off_gw);
    const float* gh = reinterpret_cast<const float*>(ws + T.off_gh);
    const float* hsum = reinterpret_cast<const float*>(ws + T.off_hsum);
    const float* gs = reinterpret_cast<const float*>(ws + T.off_gs);
    if (T.EXs) {                                                  // weight gradients of the state columns (forward must have run on this workspace)
        float* rsum = reinterpret_cast<float*>(ws + T.off_rsum);
        dz_ray_sum_kernel<<<dim3((unsigned)B, 2), 256, 0, st>>>(ws + T.off_dz, T.tr_tiles, B, S, rsum);
        NERF_CUDA_CHECK_LAUNCH();
        state_grad_kernel<40><<<dim3((unsigned)((B + SG_RAYS - 1) / SG_RAYS), 2, (T.EXs + 39) / 40), 256, 0, st>>>(
            rsum, reinterpret_cast<const float*>(ws + T.off_pes), B, T.EXs, reinterpret_cast<float*>(ws + T.off_gs));
        NERF_CUDA_CHECK_LAUNCH();
    }
    unfold_kernel<<<dim3(20, 12), 1024, 0, st>>>(params, L, gw, gh, hsum, gs, grads);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

// ======================= fp32-grade ("bf16x3") training step: host entry points =======================
// Forward: g1_fused_x3_kernel<true> (fused_tc_x3.cu) stores hi / lo slab images, statistics and the ReLU mask bit;
// backward: g1_bwd_x3_kernel (tc_bwd_x3.cu), three passes of the weight-gradient GEMM, the same unfold.
#include "tc_bwd_x3.cuh"
#include "tc_fused_kernel.cuh"
int launch_fused_x3(const void* packed, FusedArgs& a, cudaStream_t st, bool train);              // fused_tc_x3.cu
int launch_wgrad_tc_units_x3(const uint8_t* ws, const X3TrainLayout& X, cudaStream_t st);      // tc_wgrad.cu

// transposed stages as pack_bwd_stage_kernel, split: part 0 = bf16(w'), part 1 = bf16(w' - part 0)
__global__ void pack_bwd_stage_x3_kernel(const float* __restrict__ params, PackLayout L, uint8_t* __restrict__ packed) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 72ll * 128 * 8) return;
    const int c16 = (int)(i & 7), r = (int)((i >> 3) & 127), stage = (int)(i >> 10);
    const int l = 1 + stage / 8, rel = stage % 8, ks = rel >> 1, nh = rel & 1;
    const int kin = nh * 128 + r;
    const float* W = params + L.off_W[l];
    const float g = params[L.off_gamma[l - 1] + kin];
    const int Kt = L.Kt[l];
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int n0 = ks * 64 + c16 * 8 + q * 2;
        const float w0 = W[(size_t)n0 * Kt + kin] * g, w1 = W[(size_t)(n0 + 1) * Kt + kin] * g;
        hi[q] = tc::pack_bf16x2(w0, w1);
        lo[q] = tc::pack_bf16x2(w0 - bf_lo(hi[q]), w1 - bf_hi(hi[q]));
    }
    const size_t off = (size_t)stage * TC_STAGE_BYTES + tc::sw128_offset(r, c16);
    *reinterpret_cast<uint4*>(packed + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(packed + 72ll * TC_STAGE_BYTES + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
}

extern "C" int64_t nerf_g1_packed_bwd_x3_bytes(void) { return 2 * 72ll * TC_STAGE_BYTES; }

extern "C" int nerf_g1_pack_weights_bwd_x3(const float* params, void* packed_bwd, void* stream) {
    if (!params || !packed_bwd) return NERF_ERR_BAD_ARG;
    PackLayout L; int rc = make_pack_layout(3, 3, L); if (rc) return rc;
    const int64_t n = 72ll * 128 * 8;
    pack_bwd_stage_x3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(params, L, (uint8_t*)packed_bwd);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}

extern "C" int64_t nerf_g1_train_x3_workspace_bytes(int64_t B, int S) {
    if (B <= 0 || S <= 0) return NERF_ERR_BAD_ARG;
    X3TrainLayout X; make_x3_train_layout(B, S, X);
    return X.total;
}

extern "C" int nerf_g1_train_fwd_x3(const void* packed_x3, const float* rays_o, const float* rays_d, int64_t B, int S,
                                    float near_plane, float far_plane, const float* t_base, const float* t_in, const float* u,
                                    int jitter, uint64_t seed, int64_t ray_id_base, float normalize_position, int normalize_direction,
                                    float* out_t, float* out_sigma, float* out_rgb,
                                    void* workspace, int64_t workspace_bytes, void* stream) {
    if (!packed_x3 || !rays_o || !rays_d || !out_t || !out_sigma || !out_rgb || B <= 0 || S < 32) return NERF_ERR_BAD_ARG;
    X3TrainLayout X; make_x3_train_layout(B, S, X);
    if (!workspace || workspace_bytes < X.total) return NERF_ERR_WORKSPACE;
    uint8_t* ws = (uint8_t*)workspace;
    const TrainLayout& T = X.T;
    FusedArgs a{};
    a.rays_o = rays_o; a.rays_d = rays_d; a.B = B; a.S = S; a.P = T.P; a.n_tiles = T.n_tiles;
    a.t_in = t_in; a.t_base = t_base; a.u = u; a.jitter = (jitter || u) ? 1 : 0; a.seed = seed; a.ray_id_base = ray_id_base;
    a.near_plane = near_plane; a.far_plane = far_plane; a.norm = normalize_position; a.normalize_dir = normalize_direction;
    a.out_t = out_t; a.out_sigma = out_sigma; a.out_rgb = out_rgb; a.dbg = nullptr; a.dbg_layer = -1;
    a.tr_act = ws + X.off_act3; a.tr_pex = ws + X.off_pex3; a.tr_ped = ws + X.off_ped3;     // interleaved (hi, lo) images
    a.tr_mask = reinterpret_cast<uint32_t*>(ws + X.off_mask);
    a.tr_stats = reinterpret_cast<float2*>(ws + T.off_stats); a.P_pad = T.P_pad; a.tr_tiles = T.tr_tiles;
    return launch_fused_x3(packed_x3, a, (cudaStream_t)stream, true);
}

// stage: 0 = all, 1 = stop after the data-gradient kernel, 2 = stop after the weight-gradient GEMMs (tests / timing)
extern "C" int nerf_g1_train_bwd_x3(const void* packed_x3, const void* packed_bwd_x3, const float* params,
                                    const float* d_sigma, const float* d_rgb, int64_t B, int S, float* grads,
                                    void* workspace, int64_t workspace_bytes, int stage, void* stream) {
    if (!packed_x3 || !packed_bwd_x3 || !params || !d_sigma || !d_rgb || !grads || B <= 0 || S < 32) return NERF_ERR_BAD_ARG;
    PackLayout L; int rc = make_pack_layout(3, 3, L); if (rc) return rc;
    X3Layout XL; make_x3_layout(XL);
    X3TrainLayout X; make_x3_train_layout(B, S, X);
    if (!workspace || workspace_bytes < X.total) return NERF_ERR_WORKSPACE;
    const TrainLayout& T = X.T;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* ws = (uint8_t*)workspace;
    NERF_CUDA(cudaMemsetAsync(ws + T.zero_begin, 0, (size_t)T.zero_bytes, st));
    BwdX3Args a{};
    a.stages_hi = (const uint8_t*)packed_bwd_x3; a.stages_lo = a.stages_hi + 72ll * TC_STAGE_BYTES;
    a.vec = reinterpret_cast<const float*>((const uint8_t*)packed_x3 + XL.weight_bytes);
    a.act = ws + X.off_act3; a.mask = reinterpret_cast<const uint32_t*>(ws + X.off_mask); a.stats = reinterpret_cast<const float2*>(ws + T.off_stats);
    a.d_sigma = d_sigma; a.d_rgb = d_rgb; a.dz = ws + X.off_dz3;
    a.dh = ws + X.off_dh3; a.hsum = reinterpret_cast<float*>(ws + T.off_hsum);
    a.P = T.P; a.P_pad = T.P_pad; a.n_tiles = T.n_tiles; a.tr_tiles = T.tr_tiles;
    rc = launch_bwd_x3(a, st); if (rc) return rc;
    if (stage == 1) return NERF_OK;
    rc = launch_wgrad_tc_units_x3(ws, X, st); if (rc) return rc;
    if (stage == 2) return NERF_OK;
    unfold_kernel<<<dim3(20, 12), 1024, 0, st>>>(params, L, reinterpret_cast<const float*>(ws + T.off_gw), reinterpret_cast<const float*>(ws + T.off_gh),
                                                 reinterpret_cast<const float*>(ws + T.off_hsum), reinterpret_cast<const float*>(ws + T.off_gs), grads);
    NERF_CUDA_CHECK_LAUNCH();
    return NERF_OK;
}
