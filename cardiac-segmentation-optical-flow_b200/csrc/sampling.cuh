// Bilinear sampling helpers shared by the warp / lookup kernels and their backward passes.
// Weights and tap order follow ATen's grid_sampler_2d (nw, ne, sw, se; align_corners=True; zero padding).
#pragma once
#include "common.cuh"

namespace cf {

struct Tap {
    int x0, y0;
    float w00, w01, w10, w11;  // nw, ne, sw, se (ATen order)
};

__device__ __forceinline__ Tap make_tap(float uy, float ux) {
    Tap t;
    const float fx = floorf(ux), fy = floorf(uy);
    t.x0 = __float2int_rd(ux);
    t.y0 = __float2int_rd(uy);
    const float wxh = __fsub_rn(ux, fx), wxl = __fsub_rn(__fadd_rn(fx, 1.0f), ux);
    const float wyh = __fsub_rn(uy, fy), wyl = __fsub_rn(__fadd_rn(fy, 1.0f), uy);
    t.w00 = __fmul_rn(wxl, wyl);
    t.w01 = __fmul_rn(wxh, wyl);
    t.w10 = __fmul_rn(wxl, wyh);
    t.w11 = __fmul_rn(wxh, wyh);
    return t;
}

__device__ __forceinline__ float fetch(const float* __restrict__ img, int H, int W, int y, int x) {
    return ((unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W) ? __ldg(img + (size_t)y * W + x) : 0.0f;
}

__device__ __forceinline__ float sample_bilinear(const float* __restrict__ img, int H, int W, const Tap& t) {
    float acc = __fmul_rn(fetch(img, H, W, t.y0, t.x0), t.w00);
    acc = fmaf(fetch(img, H, W, t.y0, t.x0 + 1), t.w01, acc);
    acc = fmaf(fetch(img, H, W, t.y0 + 1, t.x0), t.w10, acc);
    acc = fmaf(fetch(img, H, W, t.y0 + 1, t.x0 + 1), t.w11, acc);
    return acc;
}

// one axis tap: sampling coordinate -> floor index, weight of tap i0 and of tap i0 + 1
__device__ __forceinline__ void axis_tap(float c, int d, float s1, float rcp, int& i0, float& w_lo, float& w_hi) {
    const float u = lookup_coord(__fadd_rn(c, (float)d), s1, rcp);
    const float f = floorf(u);
    i0 = __float2int_rd(u);
    w_hi = __fsub_rn(u, f);
    w_lo = __fsub_rn(__fadd_rn(f, 1.0f), u);
}

// Queries this far out have every tap outside any level image (sizes <= 4096): zeros.  Also catches NaN / inf and
// keeps |2 * (c + d)| < 2^14 for div_small.
constexpr float kDeadCoord = 8000.0f;

}  // namespace cf
