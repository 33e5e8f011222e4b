// Host orchestration of the fp32 G1 MLP (forward + backward), SURVEY A.3:
//   block_0: 4 units on ex; block_1: 4 units on cat[h, ex]; density head;
//   block_2: 2 units on cat[h, ed]; colour head.   unit = LayerNorm(ReLU(Linear)).
// Workspace (floats): n[10][P*256] | mu[10][P] | rstd[10][P] | G[P*256] | dz[P*256]
// (without save_for_backward: two ping-pong n buffers only).
#include "mlp_fp32_gemm.cuh"
#include "mlp_fp32_rows.cuh"

struct G1Layout {
    int Dx, Dd, EX, ED;
    int64_t off[44];
    int64_t total;
    // per unit l (0..9): indices into off[] of W, b, gamma, beta
    int iW(int l) const { return 4 + 4 * l; }
    int ib(int l) const { return 5 + 4 * l; }
    int ig(int l) const { return 6 + 4 * l; }
    int ibt(int l) const { return 7 + 4 * l; }
    int Kt(int l) const { return l == 0 ? EX : (l == 4 ? HID + EX : (l == 8 ? HID + ED : HID)); }
};
static int make_layout(int Dx, int Dd, G1Layout& L) {
    if (Dx < 3 || Dd < 3 || Dx > 64 || Dd > 64) return NERF_ERR_BAD_ARG;
    L.Dx = Dx; L.Dd = Dd; L.EX = 2 * NERF_PE_X_OCTAVES * Dx; L.ED = 2 * NERF_PE_D_OCTAVES * Dd;
    L.total = nerf_g1_param_layout(Dx, Dd, L.off);
    return NERF_OK;
}
static inline int64_t align64(int64_t x) { return (x + 63) & ~(int64_t)63; }

struct Fp32Ws { float* n[10]; float* mu[10]; float* rstd[10]; float* G; float* dz; };
static int64_t carve(int64_t P, int save, void* base, Fp32Ws* ws) {
    int64_t act = align64(P * HID), vec = align64(P), o = 0;
    float* b = (float*)base;
    if (save) {
        for (int l = 0; l < 10; ++l) { if (ws) ws->n[l] = b + o; o += act; }
        for (int l = 0; l < 10; ++l) { if (ws) ws->mu[l] = b + o; o += vec; }
        for (int l = 0; l < 10; ++l) { if (ws) ws->rstd[l] = b + o; o += vec; }
        if (ws) ws->G = b + o; o += act;
        if (ws) ws->dz = b + o; o += act;
    } else {
        // n7 must survive block_2 (density head reads it first, so 2 buffers suffice)
        for (int l = 0; l < 10; ++l) { if (ws) { ws->n[l] = b + (l & 1) * act; ws->mu[l] = nullptr; ws->rstd[l] = nullptr; } }
        o = 2 * act;
        if (ws) { ws->G = nullptr; ws->dz = nullptr; }
    }
    return o * (int64_t)sizeof(float);
}

extern "C" int64_t nerf_g1_mlp_workspace_bytes(int64_t P, int Dx, int Dd, int save_for_backward) {
    if (P < 0 || Dx < 3 || Dd < 3) return NERF_ERR_BAD_ARG;
    return carve(P, save_for_backward, nullptr, nullptr);
}

static ASource unit_source(const G1Layout& L, const Fp32Ws& ws, const float* params, const float* ex,
                           const float* ed, int l, int S) {
    ASource s{};
    if (l > 0) { s.n_prev = ws.n[l - 1]; s.gamma = params + L.off[L.ig(l - 1)]; s.beta = params + L.off[L.ibt(l - 1)]; s.K1 = HID; }
    s.row_div2 = 1;
    if (l == 0 || l == 4) { s.seg2 = ex; s.K2 = L.EX; }
    if (l == 8) { s.seg2 = ed; s.K2 = L.ED; s.row_div2 = S; }
    return s;
}

static unsigned row_grid(int64_t P) {
    int64_t blocks = (P + 7) / 8, cap = (int64_t)nerf_num_sms() * 4;
    return (unsigned)(blocks < cap ? blocks : cap);
}

int nerf_g1_mlp_fwd_fp32(const float* params, const float* ex, const float* ed, int64_t P, int S,
                         int Dx, int Dd, int save, float* sigma, float* rgb, void* workspace,
                         int64_t workspace_bytes, cudaStream_t st) {
    G1Layout L; int rc = make_layout(Dx, Dd, L); if (rc) return rc;
    Fp32Ws ws; if (carve(P, save, workspace, &ws) > workspace_bytes) return NERF_ERR_WORKSPACE;
    const unsigned tiles = (unsigned)((P + BM - 1) / BM);
    for (int l = 0; l < 10; ++l) {
        ASource src = unit_source(L, ws, params, ex, ed, l, S);
        mlp_tile_kernel<MODE_FWD><<<tiles, 256, 0, st>>>(src, params + L.off[L.iW(l)], L.Kt(l),
            params + L.off[L.ib(l)], ws.n[l], ws.mu[l], ws.rstd[l], P);
        NERF_CUDA_CHECK_LAUNCH();
        if (l == 7)
            head_fwd_kernel<1><<<row_grid(P), 256, 0, st>>>(ws.n[7], params + L.off[L.ig(7)], params + L.off[L.ibt(7)],
                params + L.off[0], params + L.off[1], sigma, P);
        if (l == 9)
            head_fwd_kernel<3><<<row_grid(P), 256, 0, st>>>(ws.n[9], params + L.off[L.ig(9)], params + L.off[L.ibt(9)],
                params + L.off[2], params + L.off[3], rgb, P);
        NERF_CUDA_CHECK_LAUNCH();
    }
    return NERF_OK;
}

int nerf_g1_mlp_bwd_fp32(const float* params, const float* ex, const float* ed, int64_t P, int S,
                         int Dx, int Dd, const float* d_sigma, const float* d_rgb, float* grads,
                         void* workspace, int64_t workspace_bytes, cudaStream_t st) {
    G1Layout L; int rc = make_layout(Dx, Dd, L); if (rc) return rc;
    Fp32Ws ws; if (carve(P, 1, workspace, &ws) > workspace_bytes) return NERF_ERR_WORKSPACE;
    const unsigned tiles = (unsigned)((P + BM - 1) / BM);
    // split the row reduction of wgrad so that the grid is a few waves of 148 SMs
    int64_t rows_per_split = 4096;
    const int64_t splits = (P + rows_per_split - 1) / rows_per_split;
    bool haveG = false;
    for (int l = 9; l >= 0; --l) {
        RowBwdArgs a{};
        a.G = haveG ? ws.G : nullptr;
        a.n = ws.n[l]; a.mu = ws.mu[l]; a.rstd = ws.rstd[l];
        a.gamma = params + L.off[L.ig(l)]; a.beta = params + L.off[L.ibt(l)];
        a.dz = ws.dz; a.d_gamma = grads + L.off[L.ig(l)]; a.d_beta = grads + L.off[L.ibt(l)]; a.d_bias = grads + L.off[L.ib(l)];
        a.P = P;
        if (l == 9) {
            a.dh = d_rgb; a.Wh = params + L.off[2]; a.d_Wh = grads + L.off[2]; a.d_bh = grads + L.off[3];
            row_bwd_kernel<3><<<row_grid(P), 256, 0, st>>>(a);
        } else if (l == 7) {
            a.dh = d_sigma; a.Wh = params + L.off[0]; a.d_Wh = grads + L.off[0]; a.d_bh = grads + L.off[1];
            row_bwd_kernel<1><<<row_grid(P), 256, 0, st>>>(a);
        } else {
            row_bwd_kernel<0><<<row_grid(P), 256, 0, st>>>(a);
        }
        NERF_CUDA_CHECK_LAUNCH();
        ASource src = unit_source(L, ws, params, ex, ed, l, S);
        const int Kt = L.Kt(l);
        dim3 wg((unsigned)((Kt + 63) / 64), HID / 64, (unsigned)splits);
        mlp_wgrad_kernel<<<wg, 256, 0, st>>>(ws.dz, src, Kt, grads + L.off[L.iW(l)], P, rows_per_split);
        NERF_CUDA_CHECK_LAUNCH();
        if (l > 0) {
            ASource d{}; d.n_prev = ws.dz; d.K1 = HID;
            mlp_tile_kernel<MODE_DGRAD><<<tiles, 256, 0, st>>>(d, params + L.off[L.iW(l)], Kt, nullptr, ws.G, nullptr, nullptr, P);
            NERF_CUDA_CHECK_LAUNCH();
            haveG = true;
        }
    }
    return NERF_OK;
}
