/*
 * nerf_b200 -- C ABI of the B200-native NeRF hot path (libnerf_b200.so).
 *
 * The reference (brandontrabucco/nerf) has no FFI: its boundary is the Python
 * method surface of nerf.model.NeRF.  Each entry point below is what a binding
 * for one of those methods calls; the comment above it cites the reference
 * code it replaces (paths relative to the reference repository root).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - tensors are dense, row-major, fp32 unless stated otherwise;
 *   - the library never allocates: callers pass outputs and workspaces
 *     (sizes from the *_workspace_bytes queries);
 *   - `stream` is a cudaStream_t passed as void*; calls are stream-ordered
 *     and re-entrant;
 *   - return value: 0 = ok, NERF_ERR_* < 0 = argument error, > 0 = cudaError_t.
 *   - there is no CPU fallback anywhere.
 */
#ifndef NERF_B200_H_
#define NERF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NERF_OK 0
#define NERF_ERR_BAD_ARG (-1)
#define NERF_ERR_UNSUPPORTED (-2)
#define NERF_ERR_WORKSPACE (-3)

#define NERF_HIDDEN 256
#define NERF_PE_X_OCTAVES 10
#define NERF_PE_D_OCTAVES 6
#define NERF_G1_LAYERS 10

/* The MLP contraction comes in two forms, each with its own entry points:
 *   fp32 FFMA tiles on explicit encodings (the <= 1e-3 parity mode): nerf_g1_mlp_fwd / nerf_g1_mlp_bwd;
 *   bf16 tcgen05 kind::f16 (fp32 accumulate in TMEM) fused with the encoder: nerf_g1_render_fused,
 *   nerf_g1_train_fwd_tc / nerf_g1_train_bwd_tc. */

int nerf_abi_version(void);
/* Human readable name of the last CUDA error seen by this library (thread local). */
const char* nerf_last_error(void);

/* ------------------------------------------------------------------ rays --- */

/* NeRF.generate_rays + rays_to_world_coordinates + the flatten in
 * render_image (nerf/model.py:242-278, 336-367, 727-753) for the flattened ray
 * ids [ray_begin, ray_end): id = (cam*H + row)*W + col.  cam_o [n_cams,3],
 * cam_r [n_cams,3,3]; outputs [ray_end-ray_begin, 3].  Either output may be NULL. */
int nerf_generate_rays(const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                       float focal, int64_t ray_begin, int64_t ray_end,
                       float* rays_o, float* rays_d, void* stream);

/* PixelRayDataset.__getitem__ + default collate for a whole batch
 * (nerf/dataset.py:283-316): idx -> (image, row, col), gather pixel, pose,
 * states; rotate the camera ray to world.  images [N,H,W,C] fp32, poses
 * [N,3|4,4] (pose_rows = 3 or 4), states_x [N,Sx] or NULL.  idx int64 [B].
 * Outputs: rays_o/rays_d [B,3], pixels [B,C], states_out [B,Sx] (or NULL),
 * bhw int64 [B,3] = (image_bi, image_hi, image_wi) (or NULL). */
int nerf_gather_pixel_rays(const float* images, const float* poses, int pose_rows,
                           const float* states_x, int Sx, int N, int H, int W, int C,
                           float focal, const int64_t* idx, int64_t B,
                           float* rays_o, float* rays_d, float* pixels, float* states_out,
                           int64_t* bhw, void* stream);

/* G1 sampler: linspace(near, far, S) with the stratified jitter of
 * NeRF.sample_along_rays (nerf/model.py:422-433).  u [B,S] uniforms in [0,1)
 * or NULL; if u is NULL and jitter != 0 the uniforms come from Philox4x32-7
 * keyed by (seed, ray_id_base + ray, sample).  t_base [S] (device) optionally
 * supplies the un-jittered ladder (the host binding passes torch.linspace's own
 * output so the bins are bit-identical to the reference's); NULL = computed in
 * the kernel.  t_out [B,S]. */
int nerf_sample_linear(float near_plane, float far_plane, const float* t_base, int S,
                       int64_t B, const float* u, int jitter, uint64_t seed,
                       int64_t ray_id_base, float* t_out, void* stream);

/* HEAD sampler: 2^linspace(-9.43633744014, 0, S) * box_diag with the same
 * jitter (nerf/model.py:412-435). */
int nerf_sample_geometric(float box_diag, const float* t_base, int S, int64_t B,
                          const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                          float* t_out, void* stream);

/* density noise of the training pass (nerf/model.py:652-654): sigma [B,S] +=
 * N(0,1)*std in place, Philox4x32-7 keyed by (seed, ray_id_base + ray, sample). */
int nerf_add_density_noise(float* sigma, int S, int64_t B, float std, uint64_t seed,
                           int64_t ray_id_base, void* stream);

/* Hierarchical inverse-CDF sampler (promised at nerf/model.py:191-193, never
 * implemented there; NeRF paper sec. 5.2): per ray, n_fine draws from the pdf of
 * the interior coarse weights over the midpoint bins, merged and sorted with the
 * coarse t.  t [B,S], weights [B,S] -> t_out [B,S+n_fine] ascending.
 * u_fine [B,n_fine] or NULL (= linspace(0,1,n_fine)). */
int nerf_sample_pdf_merge(const float* t, const float* weights, int S, int n_fine,
                          const float* u_fine, int64_t B, float* t_out, void* stream);

/* G1 encoder (layout of NeRF.positional_encoding, nerf/model.py:235-240; SURVEY
 * A.3): ex [B,S,20*(3+Sx)] from cat[(o+d t)/normalize, states_x], ed
 * [B,12*(3+Sd)] from cat[d/|d|, states_d].  points_out [B,S,3] optional. */
int nerf_encode_g1(const float* rays_o, const float* rays_d, const float* t,
                   const float* states_x, int Sx, const float* states_d, int Sd,
                   int64_t B, int S, float normalize_position, int normalize_direction,
                   float* ex, float* ed, float* points_out, void* stream);

/* ------------------------------------------------------------ compositing --- */

/* NeRF.alpha_compositing_coefficients + the colour composite
 * (nerf/model.py:462-469, 660) plus depth = sum w t and acc = sum w.
 * sigma [B,S] (pre-relu, noise already added), rgb [B,S,3] (pre-sigmoid).
 * Sample positions: either points [B,S,3] (exactly the reference's argument) or,
 * when points == NULL, rays_o/rays_d [B,3] + t [B,S] (p = o + d t is rebuilt in
 * registers with the reference's rounding).  t may be NULL only if depth is.
 * Outputs (each optional): pixels [B,3], depth [B], acc [B], weights [B,S]. */
int nerf_composite_fwd(const float* sigma, const float* rgb, const float* points,
                       const float* rays_o, const float* rays_d, const float* t,
                       int64_t B, int S, float* pixels, float* depth, float* acc,
                       float* weights, void* stream);

/* Backward of the above w.r.t. sigma and rgb.  g_* are the incoming gradients
 * (g_depth/g_acc/g_weights may be NULL).  saved_pixels [B,3] / saved_depth [B] /
 * saved_acc [B] are the forward call's outputs (each optional): with them
 * sum_j w_j s_j is known up front and the suffix sums need no extra pass over
 * the ray.  d_sigma [B,S], d_rgb [B,S,3]. */
int nerf_composite_bwd(const float* sigma, const float* rgb, const float* points,
                       const float* rays_o, const float* rays_d, const float* t,
                       int64_t B, int S, const float* g_pixels, const float* g_depth,
                       const float* g_acc, const float* g_weights,
                       const float* saved_pixels, const float* saved_depth, const float* saved_acc,
                       float* d_sigma, float* d_rgb, void* stream);

/* ------------------------------------------------------------------- MLP --- */

/* Layout of the flat G1 parameter vector = state_dict order of
 * examples/nerf.pth, generalised to Dx = 3+Sx position inputs and
 * Dd = 3+Sd direction inputs.  offsets_host receives 44 element offsets in
 * state_dict order; returns the total element count (638468 for Dx=Dd=3). */
int64_t nerf_g1_param_layout(int Dx, int Dd, int64_t* offsets_host);

/* Workspace (bytes) needed by nerf_g1_mlp_fwd for P points. */
int64_t nerf_g1_mlp_workspace_bytes(int64_t P, int Dx, int Dd, int save_for_backward);

/* The coordinate MLP of SURVEY A.3 (block_0 -> cat -> block_1 -> density;
 * cat dir -> block_2 -> color): ex [P,20 Dx], ed [P/S,12 Dd] (one row per ray)
 * -> sigma [P], rgb [P,3].  params = flat vector above.  With
 * save_for_backward != 0 the workspace keeps what nerf_g1_mlp_bwd needs. */
int nerf_g1_mlp_fwd(const float* params, const float* ex, const float* ed, int64_t P, int S,
                    int Dx, int Dd, int save_for_backward,
                    float* sigma, float* rgb, void* workspace, int64_t workspace_bytes,
                    void* stream);

/* Backward: d_sigma [P], d_rgb [P,3] -> ACCUMULATES into grads (flat, same layout
 * as params).  Uses the workspace written by the matching forward call. */
int nerf_g1_mlp_bwd(const float* params, const float* ex, const float* ed, int64_t P, int S,
                    int Dx, int Dd, const float* d_sigma, const float* d_rgb,
                    float* grads, void* workspace, int64_t workspace_bytes, void* stream);

/* ----------------------------------------------------------------- fused --- */

/* bytes of the bf16 weight image built by nerf_g1_pack_weights */
int64_t nerf_g1_packed_bytes(int Dx, int Dd);
/* fp32 flat params -> bf16 UMMA-ready (128B-swizzled, K-major) weight slabs with
 * the previous LayerNorm's affine folded in, plus fp32 bias/head vectors. */
int nerf_g1_pack_weights(const float* params, int Dx, int Dd, void* packed, void* stream);

/* Workspace of nerf_g1_render_fused: t, sigma, rgb per point and rays per ray; with conditional
 * inputs (Dx > 3) also the encoded states and their bias rows (2 KB per state row). */
int64_t nerf_g1_render_workspace_bytes(int64_t B, int S, int Dx, int64_t n_states);

/* Fused render (tcgen05 path): ray-gen (or explicit rays) -> stratified sample -> encode ->
 * MLP, one persistent kernel whose activations never leave the SM, then the compositing
 * kernel; replaces the body of NeRF.render_rays / one chunk of NeRF.render_image
 * (nerf/model.py:596-668, 727-761).
 * Rays: rays_o/rays_d [B,3], or both NULL -> camera rays of the flattened ids
 * [ray_begin, ray_begin+B) from cam_o [n_cams,3], cam_r [n_cams,3,3], H, W, focal.
 * Samples: t_in [B,S] if given, else the ladder t_base [S] (NULL: linspace(near,far,S)
 * computed in-kernel) with jitter from u [B,S] or, when u == NULL and jitter != 0, from
 * Philox keyed by (seed, ray_id_base + ray, sample).
 * Density noise: noise_std != 0 adds Philox N(0,1)*std to sigma before compositing.
 * Conditional inputs (G2, train_conditional_nerf.py:120-153; Dx = 3 + Sx): states_x [n_states,Sx];
 * ray r of this call (camera mode: flattened id ray_begin + r) reads row r / rays_per_state --
 * rays_per_state = 1 with one row per ray (render_rays), H*W with one row per camera
 * (render_image).  The states' share of the two layers that see the position encoding is applied
 * as a per-row bias computed in fp32 (csrc/tc_layout.cuh).  Dx == 3: states_x is ignored.
 * Outputs (each optional): out_pixels [B,3], out_depth/out_acc [B], out_weights [B,S],
 * out_t [B,S], out_sigma [B,S] (after noise), out_rgb [B,S,3].  Needs S >= 32. */
int nerf_g1_render_fused(const void* packed, int Dx, int Dd,
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
                         void* workspace, int64_t workspace_bytes, void* stream);

/* fp32-grade form of nerf_g1_render_fused for the reference's fp32 arithmetic (nn.Linear = fp32 addmm,
 * nerf/model.py:525-542) on the tensor cores: every GEMM operand is split into two bf16 parts and each product is
 * formed as hi*hi + lo*hi + hi*lo with fp32 accumulation in TMEM (~2^-17 relative; csrc/fused_tc_x3.cu).
 * nerf_g1_pack_weights_x3: flat fp32 parameters (Dx = Dd = 3) -> the hi/lo weight image of nerf_g1_packed_x3_bytes().
 * nerf_g1_render_fused_x3: arguments as nerf_g1_render_fused without the conditional inputs and the host noise;
 * workspace = nerf_g1_render_workspace_bytes(B, S, 3, 0). */
int64_t nerf_g1_packed_x3_bytes(void);
int nerf_g1_pack_weights_x3(const float* params, void* packed, void* stream);
int nerf_g1_render_fused_x3(const void* packed, const float* rays_o, const float* rays_d,
                            const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                            float focal, int64_t ray_begin, int64_t B, int S,
                            float near_plane, float far_plane, const float* t_base, const float* t_in,
                            const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                            float noise_std, float normalize_position, int normalize_direction,
                            float* out_pixels, float* out_depth, float* out_acc, float* out_weights,
                            float* out_t, float* out_sigma, float* out_rgb,
                            void* workspace, int64_t workspace_bytes, void* stream);

/* nerf_g1_render_fused for callers that cannot allocate a workspace (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  This entry point has no counterpart in the reference.
 * The scratch of the fused render (t, sigma, rgb per point, generated rays) lives in a file-scope
 * device array of the library (2.6 MB, held by every GPU that loads the library until exit): the call allocates nothing
 * and calls no cudaMalloc / cudaFree.  Limit: B * S <= nerf_g1_render_static_max_points() = 131072 (e.g. 2048 rays x 64
 * samples), S >= 32, unconditional model; beyond it the call returns NERF_ERR_WORKSPACE and nerf_last_error() says why.
 * Arguments as nerf_g1_render_fused (packed from nerf_g1_pack_weights(params, 3, 3, ...)); outputs (each optional):
 * out_pixels [B,3], out_depth [B], out_acc [B].  Every call shares the array (it keeps the last call's contents; nothing
 * in it needs clearing: each call writes what it reads): calls that overlap on the GPU race -- serialise them. */
int64_t nerf_g1_render_static_max_points(void);
int nerf_g1_render_fused_static(const void* packed, const float* rays_o, const float* rays_d,
                                const float* cam_o, const float* cam_r, int n_cams, int H, int W,
                                float focal, int64_t ray_begin, int64_t B, int S,
                                float near_plane, float far_plane, const float* t_base,
                                const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                                float noise_std, float normalize_position, int normalize_direction,
                                float* out_pixels, float* out_depth, float* out_acc, void* stream);

/* Test hook: the tcgen05 MLP alone on explicit rays and sample positions t [B,S]; when dbg
 * != NULL the normalised output of unit dbg_layer (fp32, [B*S,256]) is written there.
 * scratch_t [B,S] is overwritten. */
int nerf_g1_mlp_tc_debug(const void* packed, const float* rays_o, const float* rays_d, const float* t,
                         int64_t B, int S, float normalize_position, int normalize_direction,
                         float* out_sigma, float* out_rgb, float* scratch_t, float* dbg, int dbg_layer,
                         void* stream);

/* ------------------------------------------------- bf16 training (tcgen05) --- */

/* Slab image: the HBM format the bf16 training kernels exchange activations and gradients in.
 * A [P x 64*ns] bf16 matrix is stored per 128-point tile as ns consecutive 16 KB slabs
 * [128 rows x 128 B]; inside a slab the 16-byte chunk c of row r sits at r*128 + ((c ^ (r&7))<<4)
 * (the 128-byte-swizzled UMMA operand image, usable K-major and MN-major without a transpose). */

/* Test hook for the weight-gradient GEMM (autograd of nn.Linear.weight / .bias, a12), P = 128*n_tiles:
 *   out[256][ld] += dz^T [256 x P] * [xmain (cols 0..255) | xenc (cols 256..319) | 1 (cols 320..335, if ones)]
 * dz, xmain: slab images with ns=4; xenc: ns=1; each may be NULL.  xsmall (excludes xmain): a
 * [P x 16] bf16 operand stored per tile as [16 point groups][2 column halves][8 points][8 columns]
 * (4 KB), result in cols 0..15. */
int nerf_tc_wgrad_debug(const void* dz, const void* xmain, const void* xenc, const void* xsmall, int ones,
                        int n_tiles, float* out, int ld, void* stream);

/* bf16 training step of the G1 model = autograd of the loop body train_conditional_nerf.py:120-135
 * / example.ipynb cell 8 on tcgen05.  The caller owns one workspace of
 * nerf_g1_train_workspace_bytes(B, S, Dx) bytes shared by the forward and backward calls.
 * Conditional models (train_conditional_nerf.py:97-104, Dx = 3 + Sx): states_x [B, Sx] are per-ray
 * constants, so their share of units 0 and 4 is added as a per-ray bias (one small fp32 GEMM per step)
 * and their weight gradients come from per-ray sums of dz; Dd must be 3 (states_d: fp32 path). */
int64_t nerf_g1_train_workspace_bytes(int64_t B, int S, int Dx);
/* transposed bf16 weight stages (gamma folded) for the data-gradient MMAs */
int64_t nerf_g1_packed_bwd_bytes(int Dx, int Dd);
int nerf_g1_pack_weights_bwd(const float* params, int Dx, int Dd, void* packed_bwd, void* stream);
/* forward: sample (as nerf_g1_render_fused) -> encode -> MLP; writes t, sigma, rgb [B,S(,3)] and
 * keeps every unit's normalised output (bf16 slab images) + LayerNorm statistics in the workspace */
int nerf_g1_train_fwd_tc(const void* packed, int Dx, int Dd, const float* rays_o, const float* rays_d,
                         int64_t B, int S, float near_plane, float far_plane, const float* t_base,
                         const float* t_in, const float* u, int jitter, uint64_t seed, int64_t ray_id_base,
                         float normalize_position, int normalize_direction, const float* states_x,
                         float* out_t, float* out_sigma, float* out_rgb,
                         void* workspace, int64_t workspace_bytes, void* stream);
/* backward: d_sigma [B,S], d_rgb [B,S,3] (from nerf_composite_bwd) -> grads (flat fp32, state_dict
 * order, OVERWRITTEN).  stage = 0 runs everything; 1 / 2 stop after the data-gradient kernel / the
 * weight-gradient GEMM (test hooks reading the workspace). */
int nerf_g1_train_bwd_tc(const void* packed, const void* packed_bwd, const float* params, int Dx, int Dd,
                         const float* d_sigma, const float* d_rgb, int64_t B, int S, float* grads,
                         void* workspace, int64_t workspace_bytes, int stage, void* stream);

/* fp32-grade training step on the tensor cores ("bf16x3"): the same autograd (train_conditional_nerf.py:120-135, notebook
 * cell 8) with the reference's fp32 arithmetic (nn.Linear = fp32 addmm, nerf/model.py:525-542) carried by tcgen05: every
 * GEMM operand of the forward, the data-gradient chain and the weight-gradient GEMMs is split into two bf16 parts and
 * each product formed as hi*hi + lo*hi + hi*lo with fp32 accumulation (~2^-17 relative); LayerNorm / ReLU forward and
 * backward, the heads and the unfold are fp32.  Unconditional model (Dx = Dd = 3).  packed_x3 from
 * nerf_g1_pack_weights_x3, packed_bwd_x3 (nerf_g1_packed_bwd_x3_bytes()) from nerf_g1_pack_weights_bwd_x3; one workspace of
 * nerf_g1_train_x3_workspace_bytes(B, S) shared by the two calls (21 KB per point).  Arguments as nerf_g1_train_fwd_tc /
 * nerf_g1_train_bwd_tc; grads (flat fp32, state_dict order) are OVERWRITTEN. */
int64_t nerf_g1_packed_bwd_x3_bytes(void);
int nerf_g1_pack_weights_bwd_x3(const float* params, void* packed_bwd_x3, void* stream);
int64_t nerf_g1_train_x3_workspace_bytes(int64_t B, int S);
int nerf_g1_train_fwd_x3(const void* packed_x3, const float* rays_o, const float* rays_d, int64_t B, int S,
                         float near_plane, float far_plane, const float* t_base, const float* t_in, const float* u,
                         int jitter, uint64_t seed, int64_t ray_id_base, float normalize_position, int normalize_direction,
                         float* out_t, float* out_sigma, float* out_rgb,
                         void* workspace, int64_t workspace_bytes, void* stream);
int nerf_g1_train_bwd_x3(const void* packed_x3, const void* packed_bwd_x3, const float* params,
                         const float* d_sigma, const float* d_rgb, int64_t B, int S, float* grads,
                         void* workspace, int64_t workspace_bytes, int stage, void* stream);

/* ------------------------------------------------ HEAD generation (mip-NeRF) --- */

/* NeRF.integrated_pe (nerf/model.py:544-551 -> cast_rays :112-136, conical_frustum_to_gaussian
 * :74-81, lift_gaussian :35-45, integrated_pos_enc :139-163, expected_sin :24-30).
 * t [B,S] fenceposts -> S-1 intervals per ray: mean_out/cov_out [B,S-1,3] (optional),
 * feat_out [B,S-1,6*n_deg] (sin block, then sin(.+pi/2) block; index deg*3+c).
 * base_radius = 1/(sqrt(3)*focal_length of the constructor); degrees min_deg .. min_deg+n_deg-1. */
int nerf_head_encode(const float* rays_o, const float* rays_d, const float* t, int64_t B, int S,
                     float base_radius, int min_deg, int n_deg, float* mean_out, float* cov_out,
                     float* feat_out, void* stream);

/* NeRF.prediction_heads (nerf/model.py:525-542): 5 x (Linear, LayerNorm, ReLU) + Linear n_out.
 * params_host: HOST array of 22 DEVICE pointers in state_dict order (prediction_heads.0.weight,
 * .0.bias, .1.weight, .1.bias, .3.weight, ... .15.weight, .15.bias).  feat [P,n_in], out [P,n_out]. */
int64_t nerf_head_mlp_workspace_bytes(int64_t P);
int nerf_head_mlp_fwd(const float* const* params_host, const float* feat, int64_t P, int n_in, int n_out,
                      float* out, void* workspace, int64_t workspace_bytes, void* stream);

/* bf16 tensor-core (tcgen05) form of NeRF.forward for the HEAD generation (nerf/model.py:553-594):
 * integrated_pe + prediction_heads in one persistent kernel, activations on chip.
 *   nerf_head_pack_weights: the 22 state_dict tensors (HOST array of DEVICE pointers, as for
 *     nerf_head_mlp_fwd) -> operand image of nerf_head_packed_bytes() bytes; n_in = 3*encoding_size.
 *   nerf_head_fwd_tc: rays [B,3], fenceposts t [B,S] -> out [B,S-1,n_out] (density | colour |
 *     segmentation rows, fp32) and mean_out [B,S-1,3] (the Gaussian means nerf_head_composite takes as
 *     points).  log_softmax_seg != 0: the segmentation columns hold log_softmax(logits) (model.py:662),
 *     to be composited with seg_is_log_softmax = 1.  Built for encoding_size = 32 (n_deg = 16), n_out <= 64. */
int64_t nerf_head_packed_bytes(void);
int nerf_head_pack_weights(const float* const* params_host, int n_in, int n_out, void* packed, void* stream);
int nerf_head_fwd_tc(const void* packed, const float* rays_o, const float* rays_d, const float* t, int64_t B, int S,
                     float base_radius, int min_deg, int n_deg, int n_out, int log_softmax_seg, float* out, float* mean_out,
                     void* stream);

/* Colour + segmentation composite of NeRF.render_rays (nerf/model.py:652-663) on the trunk's
 * interleaved rows out [B,S,n_out] = [density | colour(3) | segmentation(n_seg)], n_out = 4+n_seg,
 * n_seg <= 64; points [B,S,3] are the Gaussian means.  noise [B,S] (or NULL) * noise_std is added
 * to the density.  pixels [B,3], seg_out [B,n_seg], weights [B,S] (optional).  seg_is_log_softmax: the
 * segmentation columns already went through log_softmax (nerf_head_fwd_tc can emit them that way). */
int nerf_head_composite(const float* out, int n_out, int n_seg, const float* points, const float* noise,
                        float noise_std, int64_t B, int S, float* pixels, float* seg_out, float* weights,
                        int seg_is_log_softmax, void* stream);

/* Backward of the HEAD generation in fp32: autograd of NeRF.render_rays (nerf/model.py:596-668) w.r.t. the 22
 * prediction_heads tensors.  The reference has no training script for this generation; its nn.Module is
 * differentiable through torch autograd, and these entry points are that backward.
 *   nerf_head_mlp_fwd_train: nerf_head_mlp_fwd that keeps every unit's normalised output and LayerNorm
 *     statistics in the workspace (nerf_head_mlp_train_workspace_bytes(P): 7 x [P,256] + 10 x [P] floats).
 *   nerf_head_composite_bwd: gradients d_pixels [B,3] / d_seg [B,n_seg] (either may be NULL = zero) of
 *     nerf_head_composite's outputs -> d_out [B*S, ldo] rows (density | colour | segmentation logits | zero pad),
 *     ldo a multiple of 4, >= n_out.  seg_fwd [B,n_seg] is the forward's seg_out.  S <= 3072.
 *   nerf_head_mlp_bwd: d_out [P, ldo = 64] -> grads_host: HOST array of 22 DEVICE pointers shaped like
 *     params_host; gradients are ACCUMULATED (zero them first).  workspace = the one nerf_head_mlp_fwd_train filled. */
int64_t nerf_head_mlp_train_workspace_bytes(int64_t P);
int nerf_head_mlp_fwd_train(const float* const* params_host, const float* feat, int64_t P, int n_in, int n_out,
                            float* out, void* workspace, int64_t workspace_bytes, void* stream);
int nerf_head_composite_bwd(const float* out, int n_out, int n_seg, const float* points, const float* noise,
                            float noise_std, int64_t B, int S, const float* seg_fwd, const float* d_pixels,
                            const float* d_seg, float* d_out, int ldo, void* stream);
int nerf_head_mlp_bwd(const float* const* params_host, const float* feat, int64_t P, int n_in, int n_out,
                      const float* d_out, int ldo, float* const* grads_host, void* workspace,
                      int64_t workspace_bytes, void* stream);

/* -------------------------------------------------------------- optimiser --- */

/* torch.optim.Adam(lr, betas, eps) step over a flat buffer
 * (train_conditional_nerf.py:106-107, 135): one launch for all 44 tensors.
 * grad_scale multiplies the gradient first (1/world_size after an all-reduce). */
int nerf_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq,
                   int64_t n, int64_t step, float lr, float beta1, float beta2, float eps,
                   float grad_scale, void* stream);

/* The data-parallel exchange step (one all-reduce of the flat gradient per step, then the same Adam on every
 * rank -- train_conditional_nerf.py:130-135 run on several GPUs) as ONE kernel over NVLink / NVSwitch peer memory:
 * cross-GPU barrier -> all-reduce -> Adam.  The gradient sits in symmetric memory (the same allocation on every rank, peer
 * mapped).  grads_mc: this rank's MULTICAST mapping of that allocation (NVLS: the sum is read with multimem.ld_reduce) or
 * NULL (one-shot all-reduce with P2P loads); grads_peers / flags_peers: DEVICE arrays [world] of every rank's mapping of
 * the gradient ALLOCATION and of the allocation holding a zero-initialised flag block of `world` uint64; grad_offset /
 * flag_offset: element offsets (float / uint64) of the buffer to reduce and of the flag block inside those allocations
 * (callers alternate between two gradient buffers by epoch parity); epoch: 1, 2, 3 ... (same on every rank).
 * n (elements, multiple of 4) and the remaining arguments as nerf_adam_step.  All ranks must make the call. */
int nerf_dp_allreduce_adam(const float* grads_mc, const float* const* grads_peers, uint64_t* const* flags_peers,
                           int64_t grad_offset, int64_t flag_offset, int rank, int world, uint64_t epoch,
                           float* params, float* exp_avg, float* exp_avg_sq, int64_t n, int64_t step,
                           float lr, float beta1, float beta2, float eps, float grad_scale, void* stream);

/* mean((pixels - target)^2) and its gradient w.r.t. pixels
 * (train_conditional_nerf.py:132): loss_out [1] (overwritten), g_pixels [B,3]. */
int nerf_mse_loss(const float* pixels, const float* target, int64_t n, float* loss_out,
                  float* g_pixels, void* stream);

/* The loss side of one training step (train_conditional_nerf.py:120-133 after the MLP) in ONE launch:
 * sigma + N(0,1)*noise_std (the draw of nerf_add_density_noise; sigma itself is not modified) ->
 * composite (nerf/model.py:437-469,660) -> mean((pixels - target)^2) -> its gradient back through the
 * composite.  Same results as nerf_add_density_noise + nerf_composite_fwd + nerf_mse_loss +
 * nerf_composite_bwd.  pixels [B,3] optional; loss_out [1] overwritten; d_sigma [B,S], d_rgb [B,S,3].
 * S <= 256 (NERF_ERR_UNSUPPORTED above: use the four calls). */
int nerf_composite_mse_train(const float* sigma, const float* rgb, const float* rays_o, const float* rays_d,
                             const float* t, const float* target, int64_t B, int S,
                             float noise_std, uint64_t seed, int64_t ray_id_base,
                             float* pixels, float* loss_out, float* d_sigma, float* d_rgb, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H_ */
