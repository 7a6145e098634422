// Entry points declared in include/cardioflow.h whose kernels are not written yet (they are replaced one by
// one by real implementations; each returns CF_ERR_UNSUPPORTED so a caller fails loudly).
#include "common.cuh"
using namespace cf;
extern "C" {
size_t cf_fmap_pyramid_level_offset(int, int, int, int, int) { return 0; }
size_t cf_fmap_pyramid_bytes(int, int, int, int, int) { return 0; }
int cf_fmap_pyramid_build(const float*, int, int, int, int, int, float*, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
int cf_corr_lookup_fused(const float*, const float*, const float*, int, int, int, int, int, int, float*, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
int cf_warp2d_bwd(const float*, const float*, const float*, int, int, int, int, float*, float*, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
int cf_corr_lookup_bwd(const float*, const float*, int, int, int, int, int, float*, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
int cf_corr_pool_bwd(float*, int, int, int, int, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
int cf_corr_volume_bwd(const float*, const float*, const float*, int, int, int, int, float*, float*, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
int cf_vecint(const float*, int, int, int, int, float*, float*, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
int cf_sample_points(const float*, const float*, int, int, int, int, int, int, float*, void*) { return fail(CF_ERR_UNSUPPORTED, "not implemented"); }
}
