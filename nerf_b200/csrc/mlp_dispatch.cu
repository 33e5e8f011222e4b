// C-ABI entry points of the coordinate MLP on EXPLICIT encodings: fp32 FFMA tiles (the <= 1e-3 parity mode).  The bf16
// tcgen05 contraction fuses the encoder and has its own entry points (nerf_g1_render_fused, nerf_g1_train_fwd_tc / _bwd_tc).
#include "common.cuh"

int nerf_g1_mlp_fwd_fp32(const float*, const float*, const float*, int64_t, int, int, int, int, float*, float*,
                         void*, int64_t, cudaStream_t);
int nerf_g1_mlp_bwd_fp32(const float*, const float*, const float*, int64_t, int, int, int, const float*,
                         const float*, float*, void*, int64_t, cudaStream_t);

extern "C" int nerf_g1_mlp_fwd(const float* params, const float* ex, const float* ed, int64_t P, int S,
                               int Dx, int Dd, int save_for_backward,
                               float* sigma, float* rgb, void* workspace, int64_t workspace_bytes,
                               void* stream) {
    if (P == 0 && params && S > 0) return NERF_OK;
    if (!params || !ex || !ed || !sigma || !rgb || P < 0 || S <= 0 || (P % S) != 0) return NERF_ERR_BAD_ARG;
    if (P == 0) return NERF_OK;
    if (!workspace) return NERF_ERR_WORKSPACE;
    return nerf_g1_mlp_fwd_fp32(params, ex, ed, P, S, Dx, Dd, save_for_backward, sigma, rgb, workspace,
                                workspace_bytes, (cudaStream_t)stream);
}

extern "C" int nerf_g1_mlp_bwd(const float* params, const float* ex, const float* ed, int64_t P, int S,
                               int Dx, int Dd, const float* d_sigma, const float* d_rgb,
                               float* grads, void* workspace, int64_t workspace_bytes, void* stream) {
    if (P == 0 && params && grads && S > 0) return NERF_OK;
    if (!params || !ex || !ed || !d_sigma || !d_rgb || !grads || P < 0 || S <= 0 || (P % S) != 0) return NERF_ERR_BAD_ARG;
    if (P == 0) return NERF_OK;
    if (!workspace) return NERF_ERR_WORKSPACE;
    return nerf_g1_mlp_bwd_fp32(params, ex, ed, P, S, Dx, Dd, d_sigma, d_rgb, grads, workspace,
                                workspace_bytes, (cudaStream_t)stream);
}
