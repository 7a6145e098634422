// Shared helpers for libcardioflow.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#include "cardioflow.h"

namespace cf {

// thread-local error message behind cf_last_error()
char* error_buffer();
int fail(int code, const char* fmt, ...);

#define CF_CHECK_CUDA(expr)                                                                        \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return cf::fail(CF_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),   \
                            __FILE__, __LINE__);                                                   \
    } while (0)

#define CF_CHECK_LAUNCH(name)                                                                      \
    do {                                                                                           \
        cudaError_t _e = cudaGetLastError();                                                       \
        if (_e != cudaSuccess)                                                                     \
            return cf::fail(CF_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e)); \
    } while (0)

#define CF_REQUIRE(cond, code, ...)                      \
    do {                                                 \
        if (!(cond)) return cf::fail(code, __VA_ARGS__); \
    } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline int cdiv(long a, long b) { return static_cast<int>((a + b - 1) / b); }

struct LevelDims {
    int h[CF_MAX_LEVELS];
    int w[CF_MAX_LEVELS];
    size_t off[CF_MAX_LEVELS + 1];  // byte offsets of the packed corr pyramid, off[L] = total
};
// Fills dims for L levels of a [B*H*W, 1, h_l, w_l] pyramid; returns false if a level would be empty.
bool corr_level_dims(int B, int H, int W, int L, LevelDims* d);

// ---- the exact fp32 coordinate sequences (no FMA contraction: *_rn intrinsics are never fused) ----
// Division by s1 = size - 1 (a small integer) is the one expensive step.  q = a * r corrected once with the exact
// remainder, r = RN(1/s1), equals the IEEE quotient __fdiv_rn(a, s1) bit for bit: checked EXHAUSTIVELY on B200 for
// every fp32 a with 2^-24 <= |a| < 2^14 and every s1 in 1..4095 (tools/div_probe.cu, 0 mismatches); exponent
// scaling does not change the rounding, so smaller |a| are covered too (and there every caller subtracts a constant
// >= 0.5 next, which absorbs q entirely).  |a| >= 2^14 takes the IEEE path.
__device__ __forceinline__ float div_small(float a, float s1, float r) {
    const float q0 = __fmul_rn(a, r);
    return __fmaf_rn(__fmaf_rn(-q0, s1, a), r, q0);
}
__device__ __forceinline__ float div_exact(float a, float s1, float r) {
    return (fabsf(a) < 16384.0f && s1 <= 4095.0f) ? div_small(a, s1, r) : __fdiv_rn(a, s1);
}
// lookup: RAFT bilinear_sampler n = 2*p/(S-1) - 1, then ATen unnormalize ((n+1)/2)*(S-1).  r = 1/s1 (host, IEEE).
// The lookup kernels call this only with |p| < 8192 (farther queries are dead: every tap is outside the image)
// and s1 <= 4095 (cabi.cu: level sizes are capped at 4096).
__device__ __forceinline__ float lookup_coord(float p, float s1, float r) {
    const float n = __fsub_rn(div_small(__fmul_rn(2.0f, p), s1, r), 1.0f);
    return __fmul_rn(__fmul_rn(__fadd_rn(n, 1.0f), 0.5f), s1);
}
__device__ __forceinline__ float lookup_coord_ieee(float p, float s1) {   // any |p|: IEEE division
    const float n = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, p), s1), 1.0f);
    return __fmul_rn(__fmul_rn(__fadd_rn(n, 1.0f), 0.5f), s1);
}
// warp: integration.py:68 n = 2*(p/(S-1) - 0.5), then the same ATen step
__device__ __forceinline__ float warp_coord(float p, float s1, float r) {
    const float n = __fmul_rn(2.0f, __fsub_rn(div_exact(p, s1, r), 0.5f));
    return __fmul_rn(__fmul_rn(__fadd_rn(n, 1.0f), 0.5f), s1);
}

__device__ __forceinline__ float4 ldg_f4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

}  // namespace cf

// kernels' host launchers (one per .cu file)
namespace cf {
int launch_coords_grid(float* out, int B, int H, int W, cudaStream_t s);
int launch_warp2d(const float* flow, const float* src, int B, int C, int H, int W, int mode, float* out,
                  cudaStream_t s);
int launch_warp2d_taps(const float* flow, int B, int H, int W, float* taps, cudaStream_t s);
int launch_warp2d_onehot_argmax(const float* flow, const int64_t* label, int ncls, int B, int H, int W,
                                int64_t* out, int64_t* counts, cudaStream_t s);   // counts: nullptr or [B, ncls]
int launch_warp2d_onehot_argmax_u8(const float* flow, const int64_t* label, int ncls, int B, int H, int W, uint8_t* out,
                                   int64_t* counts, int counts_total, cudaStream_t s);   // counts: nullptr, [B, ncls] or (total) [ncls]
int launch_warp2d_bwd(const float* flow, const float* src, const float* go, int B, int C, int H, int W,
                      float* gsrc, float* gflow, cudaStream_t s);
int launch_warp2d_nearest_bwd(const float* flow, const float* go, int B, int C, int H, int W, float* gsrc, cudaStream_t s);
int launch_vecint(const float* vec, int B, int H, int W, int nsteps, float* out, float* work, cudaStream_t s);
int launch_sample_points(const float* locs, const float* src, int B, int C, int H, int W, int P, int SH, int SW,
                         int mode, float* out, cudaStream_t s);
int launch_sample_points_bwd(const float* locs, const float* src, const float* go, int B, int C, int H, int W, int P, int SH,
                             int SW, int mode, float* gsrc, float* glocs, cudaStream_t s);

int launch_pool2x2(const float* in, float* out, long n_img, int h, int w, cudaStream_t s);
int launch_pool_chain(float* pyramid, const LevelDims& d, long n_img, int L, int from_level, cudaStream_t s,
                      unsigned tiled_mask = 0 /* bit l: INPUT level l is stored as 4x4 tiles; outputs are row-major */);
int launch_pool_chain_bwd(float* gpyr, const LevelDims& d, long n_img, int L, cudaStream_t s);

int launch_corr_simt(const float* f1, const float* f2, int B, int C, int HW, float* level0, cudaStream_t s);
// returns CF_ERR_UNSUPPORTED (without touching the stream) when the shape does not fit the TMA/UMMA tiling
int launch_corr_tcgen05(const float* f1, const float* f2, int B, int C, int H, int W, float* const* levels, int L,
                        void* workspace, size_t ws_bytes, int* levels_done /* levels written by the GEMM */,
                        bool allow_tiled, unsigned* tiled_mask /* out: bit l set when level l was written as 4x4 tiles */,
                        cudaStream_t s);
size_t corr_tcgen05_workspace_bytes(int B, int C, int H, int W);
bool corr_tcgen05_supported(int B, int C, int H, int W, bool fuse_pool);
// version 2 of the tensor-core build (corr_tc2.cu): persistent, swapped roles, bf16 or bf16x3, any input dtype / layout
bool corr_tc2_supported(int B, int C, int H, int W);
size_t corr_tc2_workspace_bytes(int B, int C, int H, int W, bool split);
int launch_corr_tc2(const void* f1, const void* f2, int in_dtype, int in_layout, int B, int C, int H, int W,
                    float* const* levels, int L, bool split, void* workspace, size_t ws_bytes, int* levels_done,
                    bool allow_tiled, unsigned* tiled_mask, cudaStream_t s);
int launch_corr_volume_bwd(const float* f1, const float* f2, const float* gcorr, int B, int C, int HW,
                           float* gf1, float* gf2, cudaStream_t s);
// the same on tcgen05 with the bf16x3 split (corr_bwd_tc.cu); workspace = 4 bf16 planes of the feature maps
bool corr_bwd_tc_supported(int B, int C, int HW);
size_t corr_bwd_tc_workspace_bytes(int B, int C, int HW);
int launch_corr_volume_bwd_tc(const float* f1, const float* f2, const float* gcorr, int B, int C, int HW, float* gf1,
                              float* gf2, void* ws, size_t ws_bytes, cudaStream_t s);
// the same with the planes in a static device array (corr_bwd_tc_noalloc.inl): no allocation, sizes up to CF_NOALLOC_SCRATCH_BYTES
int launch_corr_volume_bwd_tc_noalloc(const float* f1, const float* f2, const float* gcorr, int B, int C, int HW, float* gf1,
                                      float* gf2, cudaStream_t s);

int launch_lookup(const float* pyramid, const LevelDims& d, const float* coords, int B, int H, int W, int L,
                  int r, float* out, unsigned tiled_mask /* bit l: level l stored as 4x4 tiles */, cudaStream_t s);
int launch_lookup_taps(const LevelDims& d, const float* coords, int B, int H, int W, int L, int r, float* taps,
                       int32_t* idx, cudaStream_t s);
int launch_lookup_bwd(const LevelDims& d, const float* coords, const float* go, int B, int H, int W, int L,
                      int r, float* gpyr, cudaStream_t s);

int launch_warp2d_xy_bwd(const float* flow, const float* src, const float* go, int B, int C, int H, int W, float* gsrc,
                         float* gflow, cudaStream_t s);
int launch_warp2d_xy(const float* flow, const float* src, int B, int C, int H, int W, float* out, cudaStream_t s);
int launch_upsample_convex_bwd(const float* field, const float* mask, const float* gout, int B, int C, int H, int W, float scale,
                               float* gfield, float* gmask, cudaStream_t s);
int launch_upsample_convex(const float* field, const float* mask, int B, int C, int H, int W, float scale, float* out,
                           cudaStream_t s);

size_t ncc_workspace_bytes(int N, int H, int W, int backward);
int launch_ncc(const float* I, const float* J, int N, int H, int W, float eps, float* loss_map, float* cc_sum, float* partial,
               cudaStream_t s);
int launch_ncc_bwd(const float* I, const float* J, int N, int H, int W, float eps, const float* grad, int per_pixel, float mult,
                   float* gI, float* gJ, float* ws, cudaStream_t s);

size_t sum_f64_workspace_bytes();
int launch_sum_f64(const float* x, size_t n, double* out, double* partial, cudaStream_t s);

// pooled-fmap2 pyramid of the on-the-fly lookup: level l = [B, C, h_l, w_l], 256-byte aligned offsets
bool fmap_level_dims(int B, int C, int H, int W, int L, LevelDims* d);
int launch_fmap_pyramid(const float* f2, int B, int C, int H, int W, int L, float* f2pyr, cudaStream_t s);
int launch_lookup_fused(const float* f1, const float* f2pyr, const float* coords, int B, int C, int H, int W,
                        int L, int r, float* out, cudaStream_t s);

// on-the-fly lookup on the tensor cores (lookup_tc.cu): operands = bf16 channels-last f1 + pooled f2 levels
bool lookup_tc_supported(int B, int C, int H, int W, int L);
size_t lookup_tc_operand_bytes(int B, int C, int H, int W, int L);
int launch_lookup_tc_prepare(const float* f1, const float* f2, int B, int C, int H, int W, int L, void* operands, void* workspace,
                             size_t ws_bytes, cudaStream_t s);
int launch_lookup_tc(const void* operands, const float* coords, int B, int C, int H, int W, int L, int r, float* out,
                     cudaStream_t s);
// local correlation layer (local_corr.cu; parity unpinned)
int launch_local_corr(const float* cur, const float* prev, int B, int C, int H, int W, int r, int stride, float scale,
                      int x_major, float* out, cudaStream_t s);
int launch_local_corr_bwd(const float* cur, const float* prev, const float* go, int B, int C, int H, int W, int r, int stride,
                          float scale, int x_major, float* gcur, float* gprev, cudaStream_t s);
}  // namespace cf
