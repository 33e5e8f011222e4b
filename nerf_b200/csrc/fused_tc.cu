// tcgen05 fused path (placeholder until the kernel lands)
#include "common.cuh"
extern "C" int64_t nerf_g1_packed_bytes(int Dx, int Dd) { (void)Dx; (void)Dd; return 0; }
extern "C" int nerf_g1_pack_weights(const float*, int, int, void*, void*) { return NERF_ERR_UNSUPPORTED; }
extern "C" int nerf_g1_render_fused(const void*, int, int, const float*, const float*, const float*, const float*, int, int, int,
                         float, int64_t, int64_t, int, float, float, const float*, const float*, const float*, const float*,
                         const float*, float, int, float*, float*, float*, float*, float*, float*, void*) { return NERF_ERR_UNSUPPORTED; }
