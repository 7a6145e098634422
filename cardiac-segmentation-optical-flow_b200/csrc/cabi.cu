// extern "C" entry points of libcardioflow.so (include/cardioflow.h): argument validation + dispatch.
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace cf {

char* error_buffer() {
    static thread_local char buf[512] = "";
    return buf;
}

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

bool corr_level_dims(int B, int H, int W, int L, LevelDims* d) {
    size_t off = 0;
    int h = H, w = W;
    for (int l = 0; l < L; ++l) {
        if (h < 1 || w < 1) return false;
        d->h[l] = h;
        d->w[l] = w;
        d->off[l] = off;
        off += round_up((size_t)B * H * W * h * w * sizeof(float), 256);
        h /= 2;
        w /= 2;
    }
    d->off[L] = off;
    return true;
}

namespace {
int check_corr_shape(int B, int H, int W, int L, LevelDims* d) {
    CF_REQUIRE(B > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "B, H, W must be positive (got %d, %d, %d)", B, H, W);
    CF_REQUIRE(H <= 4096 && W <= 4096, CF_ERR_UNSUPPORTED, "feature maps larger than 4096 are not supported (got %dx%d)", H, W);
    CF_REQUIRE(L >= 1 && L <= CF_MAX_LEVELS, CF_ERR_UNSUPPORTED, "num_levels %d not in [1,%d]", L, CF_MAX_LEVELS);
    // torchvision raft.py:368-383: the coarsest level must keep at least 2 values per axis
    const int min_size = 2 << (L - 1);
    CF_REQUIRE(H >= min_size && W >= min_size, CF_ERR_UNSUPPORTED,
               "feature maps %dx%d too small for %d pyramid levels (need >= %d)", H, W, L, min_size);
    CF_REQUIRE(corr_level_dims(B, H, W, L, d), CF_ERR_UNSUPPORTED, "empty pyramid level");
    return CF_OK;
}
}  // namespace

}  // namespace cf

using namespace cf;

#ifdef CF_BRINGUP
namespace cf { void corr_tcgen05_set_debug(float* p); void corr_tcgen05_set_timestamps(unsigned long long* p, int cta); void lookup_set_timestamps(unsigned long long* p); }
#endif

extern "C" {

#ifdef CF_BRINGUP   // bring-up hooks (CF_NVCC_EXTRA=-DCF_BRINGUP): not in the public header, not in the shipped library
CF_API void cf_debug_set_gemm_dump(float* p) { cf::corr_tcgen05_set_debug(p); }
CF_API void cf_debug_set_gemm_timestamps(unsigned long long* p, int cta) { cf::corr_tcgen05_set_timestamps(p, cta); }
CF_API void cf_debug_set_lookup_timestamps(unsigned long long* p) { cf::lookup_set_timestamps(p); }
#endif

int cf_version(void) { return 100; }

const char* cf_last_error(void) { return error_buffer(); }

int cf_pyramid_level_dims(int H, int W, int level, int* h_out, int* w_out) {
    CF_REQUIRE(level >= 0 && level < CF_MAX_LEVELS && h_out && w_out, CF_ERR_BAD_ARG, "bad level %d", level);
    *h_out = H >> level;
    *w_out = W >> level;
    return CF_OK;
}

size_t cf_corr_pyramid_level_offset(int B, int H, int W, int level) {
    LevelDims d;
    if (level < 0 || level > CF_MAX_LEVELS || !corr_level_dims(B, H, W, level, &d)) return 0;
    return d.off[level];
}

size_t cf_corr_pyramid_bytes(int B, int H, int W, int L) { return cf_corr_pyramid_level_offset(B, H, W, L); }

int cf_coords_grid(float* out, int B, int H, int W, void* stream) {
    CF_REQUIRE(out && B > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_coords_grid: bad argument");
    return launch_coords_grid(out, B, H, W, (cudaStream_t)stream);
}

size_t cf_corr_build_workspace_bytes(int B, int C, int H, int W, int precision) {
    if (precision == CF_PREC_FP32 || B < 1 || C < 1 || H < 1 || W < 1) return 0;
    if (corr_tc2_supported(B, C, H, W)) return corr_tc2_workspace_bytes(B, C, H, W, precision != CF_PREC_BF16);
    return (precision == CF_PREC_BF16 && corr_tcgen05_supported(B, C, H, W, false)) ? corr_tcgen05_workspace_bytes(B, C, H, W) : 0;
}

int cf_corr_build(const float* f1, const float* f2, int B, int C, int H, int W, int L, int precision,
                  float* pyramid, void* workspace, size_t workspace_bytes, int* used_precision, void* stream) {
    return cf_corr_build_ex(f1, f2, CF_DTYPE_F32, CF_LAYOUT_NCHW, B, C, H, W, L, precision, 0, pyramid, workspace, workspace_bytes,
                            used_precision, nullptr, stream);
}

int cf_corr_build_layout(const float* f1, const float* f2, int B, int C, int H, int W, int L, int precision, int flags,
                         float* pyramid, void* workspace, size_t workspace_bytes, int* used_precision,
                         unsigned* tiled_mask, void* stream) {
    return cf_corr_build_ex(f1, f2, CF_DTYPE_F32, CF_LAYOUT_NCHW, B, C, H, W, L, precision, flags, pyramid, workspace,
                            workspace_bytes, used_precision, tiled_mask, stream);
}

int cf_corr_build_ex(const void* f1, const void* f2, int in_dtype, int in_layout, int B, int C, int H, int W, int L,
                     int precision, int flags, float* pyramid, void* workspace, size_t workspace_bytes,
                     int* used_precision, unsigned* tiled_mask, void* stream) {
    if (tiled_mask) *tiled_mask = 0;
    CF_REQUIRE((flags & ~CF_BUILD_ALLOW_TILED) == 0, CF_ERR_BAD_ARG, "cf_corr_build: unknown flags 0x%x", flags);
    CF_REQUIRE(!(flags & CF_BUILD_ALLOW_TILED) || tiled_mask, CF_ERR_BAD_ARG,
               "cf_corr_build: CF_BUILD_ALLOW_TILED needs tiled_mask to report the layout");
    CF_REQUIRE(f1 && f2 && pyramid, CF_ERR_BAD_ARG, "cf_corr_build: null pointer");
    CF_REQUIRE(C > 0, CF_ERR_BAD_ARG, "cf_corr_build: C must be positive");
    CF_REQUIRE(((uintptr_t)pyramid & 255u) == 0, CF_ERR_BAD_ARG, "cf_corr_build: pyramid must be 256-byte aligned");
    CF_REQUIRE(precision == CF_PREC_BF16 || precision == CF_PREC_FP32 || precision == CF_PREC_AUTO || precision == CF_PREC_BF16X3,
               CF_ERR_BAD_ARG, "cf_corr_build: unknown precision %d", precision);
    CF_REQUIRE(in_dtype == CF_DTYPE_F32 || in_dtype == CF_DTYPE_BF16, CF_ERR_BAD_ARG, "cf_corr_build: unknown in_dtype %d", in_dtype);
    CF_REQUIRE(in_layout == CF_LAYOUT_NCHW || in_layout == CF_LAYOUT_NHWC, CF_ERR_BAD_ARG, "cf_corr_build: unknown in_layout %d", in_layout);
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    float* levels[CF_MAX_LEVELS] = {};
    for (int l = 0; l < L; ++l) levels[l] = reinterpret_cast<float*>(reinterpret_cast<char*>(pyramid) + d.off[l]);
    const bool plain = in_dtype == CF_DTYPE_F32 && in_layout == CF_LAYOUT_NCHW;   // what the fp32 SIMT kernel and version 1 read

    int from_level = 0;
    int used = CF_PREC_FP32;
    if (precision != CF_PREC_FP32) {
        const bool want_split = precision != CF_PREC_BF16;
        // plain bf16 at W = 32 (the cfg2 / cfg5 shape): version 1's register-resident epilogue pools inside a thread and is the
        // faster builder there (71 vs 76 us at cfg2); version 2 serves the fp32-grade split mode, every other width (tiled
        // levels 0 / 1) and non-fp32-NCHW inputs.  CF_BF16_W32_V1=0 forces version 2 (A/B knob).
        static const bool v1_knob = [] { const char* e = getenv("CF_BF16_W32_V1"); return e ? atoi(e) != 0 : true; }();
        const bool prefer_v1 = v1_knob && !want_split && plain && W == 32 && H % 4 == 0 && corr_tcgen05_supported(B, C, H, W, true) &&
                               aligned16(f1) && aligned16(f2);
        if (!prefer_v1 && corr_tc2_supported(B, C, H, W) && aligned16(f1) && aligned16(f2)) {
            int done = 1;
            rc = launch_corr_tc2(f1, f2, in_dtype, in_layout, B, C, H, W, levels, L, want_split, workspace, workspace_bytes, &done,
                                 (flags & CF_BUILD_ALLOW_TILED) != 0, tiled_mask, s);
            if (rc) return rc;
            used = (want_split && in_dtype == CF_DTYPE_F32) ? CF_PREC_BF16X3 : CF_PREC_BF16;
            from_level = done - 1;
        } else if (precision == CF_PREC_BF16 && plain && corr_tcgen05_supported(B, C, H, W, false) && aligned16(f1) && aligned16(f2)) {
            // version 1 (W = 32 register epilogue / generic): shapes version 2 rejects (H % 4 != 0 or W % 8 != 0)
            int done = 1;
            rc = launch_corr_tcgen05(reinterpret_cast<const float*>(f1), reinterpret_cast<const float*>(f2), B, C, H, W, levels, L,
                                     workspace, workspace_bytes, &done, (flags & CF_BUILD_ALLOW_TILED) != 0, tiled_mask, s);
            if (rc) return rc;
            used = CF_PREC_BF16;
            from_level = done - 1;
        } else if (precision == CF_PREC_BF16 || precision == CF_PREC_BF16X3 || !plain) {
            return fail(CF_ERR_UNSUPPORTED,
                        "cf_corr_build: shape B=%d C=%d H=%d W=%d does not fit the tcgen05 tiling (need H %% 4 == 0, W %% 8 == 0, "
                        "C %% 8 == 0, C <= 256, 16-byte aligned inputs); use fp32 NCHW maps with CF_PREC_AUTO or CF_PREC_FP32",
                        B, C, H, W);
        }
    } else if (!plain) {
        return fail(CF_ERR_UNSUPPORTED, "cf_corr_build: CF_PREC_FP32 reads fp32 NCHW feature maps only");
    }
    if (used == CF_PREC_FP32) {
        rc = launch_corr_simt(reinterpret_cast<const float*>(f1), reinterpret_cast<const float*>(f2), B, C, H * W, levels[0], s);
        if (rc) return rc;
    }
    if (used_precision) *used_precision = used;
    return launch_pool_chain(pyramid, d, (long)B * H * W, L, from_level, s, tiled_mask ? *tiled_mask : 0u);
}

int cf_corr_pool(float* pyramid, int B, int H, int W, int L, int from_level, void* stream) {
    CF_REQUIRE(pyramid, CF_ERR_BAD_ARG, "cf_corr_pool: null pointer");
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    CF_REQUIRE(from_level >= 0 && from_level < L, CF_ERR_BAD_ARG, "cf_corr_pool: from_level %d not in [0,%d)", from_level, L);
    return launch_pool_chain(pyramid, d, (long)B * H * W, L, from_level, (cudaStream_t)stream);
}

int cf_corr_lookup(const float* pyramid, const float* coords, int B, int H, int W, int L, int r, float* out,
                   void* stream) {
    return cf_corr_lookup_layout(pyramid, coords, B, H, W, L, r, 0u, out, stream);
}

int cf_corr_lookup_layout(const float* pyramid, const float* coords, int B, int H, int W, int L, int r, unsigned tiled_mask,
                          float* out, void* stream) {
    CF_REQUIRE(pyramid && coords && out, CF_ERR_BAD_ARG, "cf_corr_lookup: null pointer");
    CF_REQUIRE(r >= 1 && r <= CF_MAX_RADIUS, CF_ERR_UNSUPPORTED, "cf_corr_lookup: radius %d not in [1,%d]", r, CF_MAX_RADIUS);
    CF_REQUIRE(((uintptr_t)pyramid & 255u) == 0, CF_ERR_BAD_ARG, "cf_corr_lookup: pyramid must be 256-byte aligned");
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    CF_REQUIRE(L >= 32 || (tiled_mask >> L) == 0, CF_ERR_BAD_ARG, "cf_corr_lookup_layout: tiled_mask 0x%x names levels >= %d", tiled_mask, L);
    return launch_lookup(pyramid, d, coords, B, H, W, L, r, out, tiled_mask, (cudaStream_t)stream);
}

int cf_corr_lookup_taps(const float* coords, int B, int H, int W, int L, int r, float* taps, int32_t* idx,
                        void* stream) {
    CF_REQUIRE(coords && taps && idx, CF_ERR_BAD_ARG, "cf_corr_lookup_taps: null pointer");
    CF_REQUIRE(r >= 1 && r <= CF_MAX_RADIUS, CF_ERR_UNSUPPORTED, "cf_corr_lookup_taps: radius %d not in [1,%d]", r, CF_MAX_RADIUS);
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    return launch_lookup_taps(d, coords, B, H, W, L, r, taps, idx, (cudaStream_t)stream);
}

int cf_warp2d(const float* flow, const float* src, int B, int C, int H, int W, int mode, float* out, void* stream) {
    CF_REQUIRE(flow && src && out, CF_ERR_BAD_ARG, "cf_warp2d: null pointer");
    CF_REQUIRE(B > 0 && C > 0 && H > 1 && W > 1, CF_ERR_BAD_ARG, "cf_warp2d: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    CF_REQUIRE(mode == CF_MODE_BILINEAR || mode == CF_MODE_NEAREST, CF_ERR_UNSUPPORTED, "cf_warp2d: mode %d", mode);
    CF_REQUIRE(out != src && out != flow, CF_ERR_BAD_ARG, "cf_warp2d: out must not alias the inputs");
    return launch_warp2d(flow, src, B, C, H, W, mode, out, (cudaStream_t)stream);
}

int cf_warp2d_taps(const float* flow, int B, int H, int W, float* taps, void* stream) {
    CF_REQUIRE(flow && taps && B > 0 && H > 1 && W > 1, CF_ERR_BAD_ARG, "cf_warp2d_taps: bad argument");
    return launch_warp2d_taps(flow, B, H, W, taps, (cudaStream_t)stream);
}

int cf_warp2d_onehot_argmax(const float* flow, const int64_t* label, int num_classes, int B, int H, int W,
                            int64_t* out, void* stream) {
    CF_REQUIRE(flow && label && out, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax: null pointer");
    CF_REQUIRE(B > 0 && H > 1 && W > 1 && num_classes > 0, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax: bad shape");
    CF_REQUIRE(out != label, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax: out must not alias label");
    return launch_warp2d_onehot_argmax(flow, label, num_classes, B, H, W, out, nullptr, (cudaStream_t)stream);
}

int cf_warp2d_onehot_argmax_counts(const float* flow, const int64_t* label, int num_classes, int B, int H, int W,
                                   int64_t* out, int64_t* counts, void* stream) {
    CF_REQUIRE(flow && label && out && counts, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax_counts: null pointer");
    CF_REQUIRE(B > 0 && H > 1 && W > 1 && num_classes > 0, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax_counts: bad shape");
    CF_REQUIRE(out != label, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax_counts: out must not alias label");
    return launch_warp2d_onehot_argmax(flow, label, num_classes, B, H, W, out, counts, (cudaStream_t)stream);
}

int cf_warp2d_onehot_argmax_u8(const float* flow, const int64_t* label, int num_classes, int B, int H, int W, uint8_t* out,
                               int64_t* counts, int counts_total, void* stream) {
    CF_REQUIRE(flow && label && out, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax_u8: null pointer");
    CF_REQUIRE(B > 0 && H > 1 && W > 1 && num_classes > 0, CF_ERR_BAD_ARG, "cf_warp2d_onehot_argmax_u8: bad shape");
    return launch_warp2d_onehot_argmax_u8(flow, label, num_classes, B, H, W, out, counts, counts_total != 0, (cudaStream_t)stream);
}

// ---- on-the-fly lookup (K4) --------------------------------------------------------------------------------
size_t cf_fmap_pyramid_level_offset(int B, int C, int H, int W, int level) {
    LevelDims d;
    if (B < 1 || C < 1 || level < 0 || level > CF_MAX_LEVELS || !fmap_level_dims(B, C, H, W, level, &d)) return 0;
    return d.off[level];
}

size_t cf_fmap_pyramid_bytes(int B, int C, int H, int W, int L) { return cf_fmap_pyramid_level_offset(B, C, H, W, L); }

int cf_fmap_pyramid_build(const float* f2, int B, int C, int H, int W, int L, float* f2_pyramid, void* stream) {
    CF_REQUIRE(f2 && f2_pyramid, CF_ERR_BAD_ARG, "cf_fmap_pyramid_build: null pointer");
    CF_REQUIRE(C > 0, CF_ERR_BAD_ARG, "cf_fmap_pyramid_build: C must be positive");
    CF_REQUIRE(((uintptr_t)f2_pyramid & 255u) == 0, CF_ERR_BAD_ARG, "cf_fmap_pyramid_build: f2_pyramid must be 256-byte aligned");
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    return launch_fmap_pyramid(f2, B, C, H, W, L, f2_pyramid, (cudaStream_t)stream);
}

int cf_corr_lookup_fused(const float* f1, const float* f2_pyramid, const float* coords, int B, int C, int H, int W,
                         int L, int r, float* out, void* stream) {
    CF_REQUIRE(f1 && f2_pyramid && coords && out, CF_ERR_BAD_ARG, "cf_corr_lookup_fused: null pointer");
    CF_REQUIRE(C > 0, CF_ERR_BAD_ARG, "cf_corr_lookup_fused: C must be positive");
    CF_REQUIRE(r >= 1 && r <= CF_MAX_RADIUS, CF_ERR_UNSUPPORTED, "cf_corr_lookup_fused: radius %d not in [1,%d]", r, CF_MAX_RADIUS);
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    return launch_lookup_fused(f1, f2_pyramid, coords, B, C, H, W, L, r, out, (cudaStream_t)stream);
}

// ---- on-the-fly lookup on the tensor cores (K4, tcgen05) ----------------------------------------------------------
int cf_corr_otf_supported(int B, int C, int H, int W, int L) { return lookup_tc_supported(B, C, H, W, L) ? 1 : 0; }

size_t cf_corr_otf_operand_bytes(int B, int C, int H, int W, int L) {
    return lookup_tc_supported(B, C, H, W, L) ? lookup_tc_operand_bytes(B, C, H, W, L) : 0;
}

int cf_corr_otf_prepare(const float* f1, const float* f2, int B, int C, int H, int W, int L, void* operands, void* workspace,
                        size_t workspace_bytes, void* stream) {
    CF_REQUIRE(f1 && f2 && operands, CF_ERR_BAD_ARG, "cf_corr_otf_prepare: null pointer");
    CF_REQUIRE(((uintptr_t)operands & 255u) == 0, CF_ERR_BAD_ARG, "cf_corr_otf_prepare: operands must be 256-byte aligned");
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    CF_REQUIRE(C > 0 && lookup_tc_supported(B, C, H, W, L), CF_ERR_UNSUPPORTED,
               "cf_corr_otf_prepare: shape B=%d C=%d H=%d W=%d needs H %% 8 == 0, W %% 16 == 0, C %% 8 == 0, C <= 256 "
               "(cf_corr_lookup_fused serves every other shape)", B, C, H, W);
    return launch_lookup_tc_prepare(f1, f2, B, C, H, W, L, operands, workspace, workspace_bytes, (cudaStream_t)stream);
}

int cf_corr_lookup_otf(const void* operands, const float* coords, int B, int C, int H, int W, int L, int r, float* out,
                       void* stream) {
    CF_REQUIRE(operands && coords && out, CF_ERR_BAD_ARG, "cf_corr_lookup_otf: null pointer");
    CF_REQUIRE(r >= 1 && r <= CF_MAX_RADIUS, CF_ERR_UNSUPPORTED, "cf_corr_lookup_otf: radius %d not in [1,%d]", r, CF_MAX_RADIUS);
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    CF_REQUIRE(C > 0 && lookup_tc_supported(B, C, H, W, L), CF_ERR_UNSUPPORTED, "cf_corr_lookup_otf: unsupported shape");
    return launch_lookup_tc(operands, coords, B, C, H, W, L, r, out, (cudaStream_t)stream);
}

// ---- 8(f) rows: backward passes, VecInt, point sampling -------------------------------------------------------
int cf_warp2d_bwd(const float* flow, const float* src, const float* grad_out, int B, int C, int H, int W,
                  float* grad_src, float* grad_flow, void* stream) {
    CF_REQUIRE(flow && src && grad_out, CF_ERR_BAD_ARG, "cf_warp2d_bwd: null pointer");
    CF_REQUIRE(grad_src || grad_flow, CF_ERR_BAD_ARG, "cf_warp2d_bwd: neither grad_src nor grad_flow requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 1 && W > 1, CF_ERR_BAD_ARG, "cf_warp2d_bwd: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    return launch_warp2d_bwd(flow, src, grad_out, B, C, H, W, grad_src, grad_flow, (cudaStream_t)stream);
}

int cf_warp2d_nearest_bwd(const float* flow, const float* grad_out, int B, int C, int H, int W, float* grad_src, void* stream) {
    CF_REQUIRE(flow && grad_out && grad_src, CF_ERR_BAD_ARG, "cf_warp2d_nearest_bwd: null pointer");
    CF_REQUIRE(B > 0 && C > 0 && H > 1 && W > 1, CF_ERR_BAD_ARG, "cf_warp2d_nearest_bwd: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    return launch_warp2d_nearest_bwd(flow, grad_out, B, C, H, W, grad_src, (cudaStream_t)stream);
}

int cf_corr_lookup_bwd(const float* coords, const float* grad_out, int B, int H, int W, int L, int r,
                       float* grad_pyramid, void* stream) {
    CF_REQUIRE(coords && grad_out && grad_pyramid, CF_ERR_BAD_ARG, "cf_corr_lookup_bwd: null pointer");
    CF_REQUIRE(r >= 1 && r <= CF_MAX_RADIUS, CF_ERR_UNSUPPORTED, "cf_corr_lookup_bwd: radius %d not in [1,%d]", r, CF_MAX_RADIUS);
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    return launch_lookup_bwd(d, coords, grad_out, B, H, W, L, r, grad_pyramid, (cudaStream_t)stream);
}

int cf_corr_pool_bwd(float* grad_pyramid, int B, int H, int W, int L, void* stream) {
    CF_REQUIRE(grad_pyramid, CF_ERR_BAD_ARG, "cf_corr_pool_bwd: null pointer");
    LevelDims d;
    int rc = check_corr_shape(B, H, W, L, &d);
    if (rc) return rc;
    return launch_pool_chain_bwd(grad_pyramid, d, (long)B * H * W, L, (cudaStream_t)stream);
}

int cf_corr_volume_bwd(const float* f1, const float* f2, const float* grad_corr, int B, int C, int H, int W,
                       float* grad_f1, float* grad_f2, void* stream) {
    CF_REQUIRE(f1 && f2 && grad_corr, CF_ERR_BAD_ARG, "cf_corr_volume_bwd: null pointer");
    CF_REQUIRE(grad_f1 || grad_f2, CF_ERR_BAD_ARG, "cf_corr_volume_bwd: neither grad_f1 nor grad_f2 requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_corr_volume_bwd: bad shape");
    return launch_corr_volume_bwd(f1, f2, grad_corr, B, C, H * W, grad_f1, grad_f2, (cudaStream_t)stream);
}

size_t cf_corr_volume_bwd_workspace_bytes(int B, int C, int H, int W) {
    if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || !corr_bwd_tc_supported(B, C, H * W)) return 0;
    return corr_bwd_tc_workspace_bytes(B, C, H * W);
}

int cf_corr_volume_bwd_ex(const float* f1, const float* f2, const float* grad_corr, int B, int C, int H, int W, int precision,
                          float* grad_f1, float* grad_f2, void* workspace, size_t workspace_bytes, int* used_precision,
                          void* stream) {
    CF_REQUIRE(f1 && f2 && grad_corr, CF_ERR_BAD_ARG, "cf_corr_volume_bwd_ex: null pointer");
    CF_REQUIRE(grad_f1 || grad_f2, CF_ERR_BAD_ARG, "cf_corr_volume_bwd_ex: neither grad_f1 nor grad_f2 requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_corr_volume_bwd_ex: bad shape");
    CF_REQUIRE(precision == CF_PREC_AUTO || precision == CF_PREC_FP32 || precision == CF_PREC_BF16X3, CF_ERR_BAD_ARG,
               "cf_corr_volume_bwd_ex: precision %d (auto, fp32 or bf16x3)", precision);
    const bool tc_ok = corr_bwd_tc_supported(B, C, H * W);
    if (precision == CF_PREC_BF16X3 && !tc_ok)
        return fail(CF_ERR_UNSUPPORTED, "cf_corr_volume_bwd_ex: bf16x3 needs H*W %% 8 == 0 and a driver with tensor-map encoding");
    const bool use_tc = precision != CF_PREC_FP32 && tc_ok;
    if (used_precision) *used_precision = use_tc ? CF_PREC_BF16X3 : CF_PREC_FP32;
    if (use_tc)
        return launch_corr_volume_bwd_tc(f1, f2, grad_corr, B, C, H * W, grad_f1, grad_f2, workspace, workspace_bytes,
                                         (cudaStream_t)stream);
    return launch_corr_volume_bwd(f1, f2, grad_corr, B, C, H * W, grad_f1, grad_f2, (cudaStream_t)stream);
}

int cf_corr_volume_bwd_noalloc(const float* f1, const float* f2, const float* grad_corr, int B, int C, int H, int W,
                               float* grad_f1, float* grad_f2, void* stream) {
    CF_REQUIRE(f1 && f2 && grad_corr, CF_ERR_BAD_ARG, "cf_corr_volume_bwd_noalloc: null pointer");
    CF_REQUIRE(grad_f1 || grad_f2, CF_ERR_BAD_ARG, "cf_corr_volume_bwd_noalloc: neither grad_f1 nor grad_f2 requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_corr_volume_bwd_noalloc: bad shape");
    return launch_corr_volume_bwd_tc_noalloc(f1, f2, grad_corr, B, C, H * W, grad_f1, grad_f2, (cudaStream_t)stream);
}

int cf_vecint(const float* vec, int B, int H, int W, int nsteps, float* out, float* work, void* stream) {
    CF_REQUIRE(vec && out, CF_ERR_BAD_ARG, "cf_vecint: null pointer");
    CF_REQUIRE(B > 0 && H > 1 && W > 1, CF_ERR_BAD_ARG, "cf_vecint: bad shape B=%d H=%d W=%d", B, H, W);
    CF_REQUIRE(nsteps >= 0 && nsteps <= 30, CF_ERR_BAD_ARG, "cf_vecint: nsteps %d not in [0,30]", nsteps);
    CF_REQUIRE(nsteps < 2 || work, CF_ERR_BAD_ARG, "cf_vecint: work buffer required for nsteps >= 2");
    CF_REQUIRE(out != vec && work != vec && (work != out || nsteps < 2), CF_ERR_BAD_ARG, "cf_vecint: buffers must not alias");
    return launch_vecint(vec, B, H, W, nsteps, out, work, (cudaStream_t)stream);
}

int cf_sample_points(const float* locs, const float* src, int B, int C, int H, int W, int P, int size_h, int size_w,
                     int mode, float* out, void* stream) {
    CF_REQUIRE(locs && src && out, CF_ERR_BAD_ARG, "cf_sample_points: null pointer");
    CF_REQUIRE(B > 0 && C > 0 && H > 1 && W > 1 && P > 0 && size_h > 1 && size_w > 1, CF_ERR_BAD_ARG, "cf_sample_points: bad shape");
    CF_REQUIRE(mode == CF_MODE_BILINEAR || mode == CF_MODE_NEAREST, CF_ERR_UNSUPPORTED, "cf_sample_points: mode %d", mode);
    return launch_sample_points(locs, src, B, C, H, W, P, size_h, size_w, mode, out, (cudaStream_t)stream);
}

int cf_sample_points_bwd(const float* locs, const float* src, const float* grad_out, int B, int C, int H, int W, int P,
                         int size_h, int size_w, int mode, float* grad_src, float* grad_locs, void* stream) {
    CF_REQUIRE(locs && src && grad_out, CF_ERR_BAD_ARG, "cf_sample_points_bwd: null pointer");
    CF_REQUIRE(grad_src || grad_locs, CF_ERR_BAD_ARG, "cf_sample_points_bwd: neither grad_src nor grad_locs requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 1 && W > 1 && P > 0 && size_h > 1 && size_w > 1, CF_ERR_BAD_ARG, "cf_sample_points_bwd: bad shape");
    CF_REQUIRE(mode == CF_MODE_BILINEAR || mode == CF_MODE_NEAREST, CF_ERR_UNSUPPORTED, "cf_sample_points_bwd: mode %d", mode);
    return launch_sample_points_bwd(locs, src, grad_out, B, C, H, W, P, size_h, size_w, mode, grad_src, grad_locs,
                                    (cudaStream_t)stream);
}

int cf_local_corr(const float* cur, const float* prev, int B, int C, int H, int W, int radius, int stride, float scale, int x_major,
                  float* out, void* stream) {
    CF_REQUIRE(cur && prev && out, CF_ERR_BAD_ARG, "cf_local_corr: null pointer");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && stride >= 1, CF_ERR_BAD_ARG, "cf_local_corr: bad shape B=%d C=%d H=%d W=%d stride=%d", B, C, H, W, stride);
    CF_REQUIRE(out != cur && out != prev, CF_ERR_BAD_ARG, "cf_local_corr: out must not alias the inputs");
    return launch_local_corr(cur, prev, B, C, H, W, radius, stride, scale, x_major != 0, out, (cudaStream_t)stream);
}

int cf_local_corr_bwd(const float* cur, const float* prev, const float* grad_out, int B, int C, int H, int W, int radius, int stride,
                      float scale, int x_major, float* grad_cur, float* grad_prev, void* stream) {
    CF_REQUIRE(cur && prev && grad_out, CF_ERR_BAD_ARG, "cf_local_corr_bwd: null pointer");
    CF_REQUIRE(grad_cur || grad_prev, CF_ERR_BAD_ARG, "cf_local_corr_bwd: neither gradient requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && stride >= 1, CF_ERR_BAD_ARG, "cf_local_corr_bwd: bad shape");
    return launch_local_corr_bwd(cur, prev, grad_out, B, C, H, W, radius, stride, scale, x_major != 0, grad_cur, grad_prev,
                                 (cudaStream_t)stream);
}

int cf_warp2d_xy(const float* flow, const float* src, int B, int C, int H, int W, float* out, void* stream) {
    CF_REQUIRE(flow && src && out, CF_ERR_BAD_ARG, "cf_warp2d_xy: null pointer");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_warp2d_xy: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    CF_REQUIRE(out != src && out != flow, CF_ERR_BAD_ARG, "cf_warp2d_xy: out must not alias the inputs");
    return launch_warp2d_xy(flow, src, B, C, H, W, out, (cudaStream_t)stream);
}

int cf_warp2d_xy_bwd(const float* flow, const float* src, const float* grad_out, int B, int C, int H, int W, float* grad_src,
                     float* grad_flow, void* stream) {
    CF_REQUIRE(flow && src && grad_out, CF_ERR_BAD_ARG, "cf_warp2d_xy_bwd: null pointer");
    CF_REQUIRE(grad_src || grad_flow, CF_ERR_BAD_ARG, "cf_warp2d_xy_bwd: neither grad_src nor grad_flow requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_warp2d_xy_bwd: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    return launch_warp2d_xy_bwd(flow, src, grad_out, B, C, H, W, grad_src, grad_flow, (cudaStream_t)stream);
}

int cf_upsample_convex(const float* field, const float* mask, int B, int C, int H, int W, int factor, float scale, float* out,
                       void* stream) {
    CF_REQUIRE(field && mask && out, CF_ERR_BAD_ARG, "cf_upsample_convex: null pointer");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_upsample_convex: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    CF_REQUIRE(factor == 8, CF_ERR_UNSUPPORTED, "cf_upsample_convex: factor %d (the reference only has 8)", factor);
    CF_REQUIRE(aligned16(out), CF_ERR_BAD_ARG, "cf_upsample_convex: out must be 16-byte aligned");
    CF_REQUIRE(out != field && out != mask, CF_ERR_BAD_ARG, "cf_upsample_convex: out must not alias the inputs");
    return launch_upsample_convex(field, mask, B, C, H, W, scale, out, (cudaStream_t)stream);
}

int cf_upsample_convex_bwd(const float* field, const float* mask, const float* grad_out, int B, int C, int H, int W, int factor,
                           float scale, float* grad_field, float* grad_mask, void* stream) {
    CF_REQUIRE(field && mask && grad_out, CF_ERR_BAD_ARG, "cf_upsample_convex_bwd: null pointer");
    CF_REQUIRE(grad_field || grad_mask, CF_ERR_BAD_ARG, "cf_upsample_convex_bwd: neither grad_field nor grad_mask requested");
    CF_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_upsample_convex_bwd: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    CF_REQUIRE(factor == 8, CF_ERR_UNSUPPORTED, "cf_upsample_convex_bwd: factor %d (the reference only has 8)", factor);
    CF_REQUIRE(aligned16(grad_out), CF_ERR_BAD_ARG, "cf_upsample_convex_bwd: grad_out must be 16-byte aligned");
    return launch_upsample_convex_bwd(field, mask, grad_out, B, C, H, W, scale, grad_field, grad_mask, (cudaStream_t)stream);
}

size_t cf_ncc_workspace_bytes(int N, int H, int W, int backward) { return ncc_workspace_bytes(N, H, W, backward); }

int cf_ncc(const float* I, const float* J, int N, int H, int W, int win, float eps, float* loss_map, float* cc_sum, void* workspace,
           size_t ws_bytes, void* stream) {
    CF_REQUIRE(I && J, CF_ERR_BAD_ARG, "cf_ncc: null pointer");
    CF_REQUIRE(loss_map || cc_sum, CF_ERR_BAD_ARG, "cf_ncc: neither loss_map nor cc_sum requested");
    CF_REQUIRE(N > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_ncc: bad shape N=%d H=%d W=%d", N, H, W);
    CF_REQUIRE(win == 9, CF_ERR_UNSUPPORTED, "cf_ncc: win %d (every NCC in the reference is built with the default 9)", win);
    CF_REQUIRE(!cc_sum || (workspace && ws_bytes >= ncc_workspace_bytes(N, H, W, 0)), CF_ERR_BAD_ARG,
               "cf_ncc: workspace of %zu bytes needed, got %zu", ncc_workspace_bytes(N, H, W, 0), ws_bytes);
    CF_REQUIRE(loss_map != I && loss_map != J, CF_ERR_BAD_ARG, "cf_ncc: loss_map must not alias the inputs");
    return launch_ncc(I, J, N, H, W, eps, loss_map, cc_sum, (float*)workspace, (cudaStream_t)stream);
}

int cf_ncc_bwd(const float* I, const float* J, int N, int H, int W, int win, float eps, const float* grad, int per_pixel, float mult,
               float* grad_I, float* grad_J, void* workspace, size_t ws_bytes, void* stream) {
    CF_REQUIRE(I && J && grad, CF_ERR_BAD_ARG, "cf_ncc_bwd: null pointer");
    CF_REQUIRE(grad_I || grad_J, CF_ERR_BAD_ARG, "cf_ncc_bwd: neither grad_I nor grad_J requested");
    CF_REQUIRE(N > 0 && H > 0 && W > 0, CF_ERR_BAD_ARG, "cf_ncc_bwd: bad shape N=%d H=%d W=%d", N, H, W);
    CF_REQUIRE(win == 9, CF_ERR_UNSUPPORTED, "cf_ncc_bwd: win %d (every NCC in the reference is built with the default 9)", win);
    CF_REQUIRE(workspace && ws_bytes >= ncc_workspace_bytes(N, H, W, 1), CF_ERR_BAD_ARG,
               "cf_ncc_bwd: workspace of %zu bytes needed, got %zu", ncc_workspace_bytes(N, H, W, 1), ws_bytes);
    CF_REQUIRE(grad_I != I && grad_I != J && grad_J != I && grad_J != J, CF_ERR_BAD_ARG,
               "cf_ncc_bwd: gradients must not alias the inputs");
    return launch_ncc_bwd(I, J, N, H, W, eps, grad, per_pixel, mult, grad_I, grad_J, (float*)workspace, (cudaStream_t)stream);
}

size_t cf_sum_f64_workspace_bytes(void) { return sum_f64_workspace_bytes(); }

int cf_sum_f64(const float* x, size_t n, double* out, void* workspace, size_t ws_bytes, void* stream) {
    CF_REQUIRE(x && out, CF_ERR_BAD_ARG, "cf_sum_f64: null pointer");
    CF_REQUIRE(aligned16(x), CF_ERR_BAD_ARG, "cf_sum_f64: x must be 16-byte aligned");
    CF_REQUIRE(workspace && ws_bytes >= sum_f64_workspace_bytes(), CF_ERR_BAD_ARG, "cf_sum_f64: workspace of %zu bytes needed, got %zu",
               sum_f64_workspace_bytes(), ws_bytes);
    return launch_sum_f64(x, n, out, (double*)workspace, (cudaStream_t)stream);
}

}  // extern "C"
