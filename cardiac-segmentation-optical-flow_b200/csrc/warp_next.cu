// SURVEY.md 8(f) rows 1-2 for the warp: backward of SpatialTransformer.forward, VecInt scaling-and-squaring,
// SpatialTransformerContour point sampling.
//
//   cf_warp2d_bwd     autograd of /root/reference/nnunet/network_architecture/integration.py:61-79 as the
//                     trainer differentiates it (`l.mean().backward()`, training/network_training/
//                     SegFlowGaussian.py:2225): ATen grid_sampler_2d_backward (bilinear, zeros,
//                     align_corners=True) chained through the reference's `2 * (p / (S - 1) - 0.5)`.
//   cf_vecint         integration.py:82-99: vec * 2^-n, then n x (vec = vec + warp(vec, vec)).  The reference runs
//                     7 x 6 launches; here n launches (the first one folds the scaling in), ping-pong buffers.
//   cf_sample_points  integration.py:5-34: grid_sample at P sparse points (contour tracking, get_strain.py:81).
#include "common.cuh"
#include "sampling.cuh"

namespace cf {
namespace {

// ---------------------------------------------------------------------------------------------- warp backward
// lib/utils.warp coordinates (utils.py:49-50 + grid_sample align_corners=False): see warp2d_xy_kernel below
__device__ __forceinline__ float utils_warp_coord_bwd(float p, float d1 /* max(S-1,1) */, float size) {
    const float n = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, p), d1), 1.0f);
    return __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(n, 1.0f), size), 1.0f), 0.5f);
}

// One thread per output pixel, all C channels: d(out)/d(src) is scattered with atomics (several output pixels
// may sample the same source texel), d(out)/d(flow) is a per-pixel reduction over channels (no atomics).
// XY = false: SpatialTransformer (flow ch0 displaces rows, align_corners=True); XY = true: lib/utils.warp (flow ch0 displaces
// columns, align_corners=False: d(u)/d(n) = S/2, d(n)/d(p) = 2/max(S-1,1)).
template <bool XY>
__global__ void __launch_bounds__(256) warp2d_bwd_kernel(const float* __restrict__ flow, const float* __restrict__ src,
                                                         const float* __restrict__ go, float* __restrict__ gsrc,
                                                         float* __restrict__ gflow, int C, int H, int W, float ry, float rx,
                                                         long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int x = (int)(t % W);
    const long row = t / W;
    const int y = (int)(row % H);
    const long b = row / H;
    const size_t HW = (size_t)H * W, pix = (size_t)y * W + x;
    const float s1y = (float)(H - 1), s1x = (float)(W - 1);
    const float fl0 = __ldg(flow + (b * 2 + 0) * HW + pix), fl1 = __ldg(flow + (b * 2 + 1) * HW + pix);
    const float d1y = (float)max(H - 1, 1), d1x = (float)max(W - 1, 1);
    const float uy = XY ? utils_warp_coord_bwd(__fadd_rn((float)y, fl1), d1y, (float)H) : warp_coord(__fadd_rn((float)y, fl0), s1y, ry);
    const float ux = XY ? utils_warp_coord_bwd(__fadd_rn((float)x, fl0), d1x, (float)W) : warp_coord(__fadd_rn((float)x, fl1), s1x, rx);
    const Tap tp = make_tap(uy, ux);
    const float fx = floorf(ux), fy = floorf(uy);
    // ATen names: (ix_nw, iy_nw) = (fx, fy); se = nw + 1
    const float dx_lo = __fsub_rn(__fadd_rn(fx, 1.0f), ux), dx_hi = __fsub_rn(ux, fx);
    const float dy_lo = __fsub_rn(__fadd_rn(fy, 1.0f), uy), dy_hi = __fsub_rn(uy, fy);
    const bool y0in = (unsigned)tp.y0 < (unsigned)H, y1in = (unsigned)(tp.y0 + 1) < (unsigned)H;
    const bool x0in = (unsigned)tp.x0 < (unsigned)W, x1in = (unsigned)(tp.x0 + 1) < (unsigned)W;
    const size_t o00 = (size_t)tp.y0 * W + tp.x0;

    float gix = 0.0f, giy = 0.0f;
    for (int c = 0; c < C; ++c) {
        const size_t chan = (b * C + c) * HW;
        const float g = __ldg(go + chan + pix);
        if (gsrc) {
            float* gs = gsrc + chan;
            if (y0in && x0in) atomicAdd(gs + o00, tp.w00 * g);
            if (y0in && x1in) atomicAdd(gs + o00 + 1, tp.w01 * g);
            if (y1in && x0in) atomicAdd(gs + o00 + W, tp.w10 * g);
            if (y1in && x1in) atomicAdd(gs + o00 + W + 1, tp.w11 * g);
        }
        if (gflow) {
            const float* img = src + chan;
            const float nw = (y0in && x0in) ? __ldg(img + o00) : 0.0f;
            const float ne = (y0in && x1in) ? __ldg(img + o00 + 1) : 0.0f;
            const float sw = (y1in && x0in) ? __ldg(img + o00 + W) : 0.0f;
            const float se = (y1in && x1in) ? __ldg(img + o00 + W + 1) : 0.0f;
            // ATen GridSampler backward, bilinear
            gix -= nw * dy_lo * g;  giy -= nw * dx_lo * g;
            gix += ne * dy_lo * g;  giy -= ne * dx_hi * g;
            gix -= sw * dy_hi * g;  giy += sw * dx_lo * g;
            gix += se * dy_hi * g;  giy += se * dx_hi * g;
        }
    }
    if (gflow) {
        if (XY) {
            // grid_sampler un-normalise (align_corners=False): d(u)/d(n) = S/2 ; utils.warp normalise: d(n)/d(p) = 2/max(S-1,1)
            const float gx = __fdiv_rn(__fmul_rn(__fmul_rn(gix, __fmul_rn((float)W, 0.5f)), 2.0f), d1x);
            const float gy = __fdiv_rn(__fmul_rn(__fmul_rn(giy, __fmul_rn((float)H, 0.5f)), 2.0f), d1y);
            gflow[(b * 2 + 0) * HW + pix] = gx;   // flow channel 0 displaces columns
            gflow[(b * 2 + 1) * HW + pix] = gy;
        } else {
            // grid_sampler un-normalise: d(u)/d(n) = (S-1)/2 ; reference normalise: d(n)/d(p) = 2/(S-1)
            const float gx = __fdiv_rn(__fmul_rn(__fmul_rn(gix, __fmul_rn(s1x, 0.5f)), 2.0f), s1x);
            const float gy = __fdiv_rn(__fmul_rn(__fmul_rn(giy, __fmul_rn(s1y, 0.5f)), 2.0f), s1y);
            gflow[(b * 2 + 0) * HW + pix] = gy;   // flow channel 0 displaces rows
            gflow[(b * 2 + 1) * HW + pix] = gx;
        }
    }
}

// Nearest-mode backward (ATen grid_sampler_2d_backward, nearest): the whole output gradient goes to the nearest texel,
// the gradient w.r.t. the sampling grid (hence the flow) is zero.
__global__ void __launch_bounds__(256) warp2d_nearest_bwd_kernel(const float* __restrict__ flow, const float* __restrict__ go,
                                                                 float* __restrict__ gsrc, int C, int H, int W, float ry,
                                                                 float rx, long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int x = (int)(t % W);
    const long row = t / W;
    const int y = (int)(row % H);
    const long b = row / H;
    const size_t HW = (size_t)H * W, pix = (size_t)y * W + x;
    const float uy = warp_coord(__fadd_rn((float)y, __ldg(flow + (b * 2 + 0) * HW + pix)), (float)(H - 1), ry);
    const float ux = warp_coord(__fadd_rn((float)x, __ldg(flow + (b * 2 + 1) * HW + pix)), (float)(W - 1), rx);
    const int ny = __float2int_rn(uy), nx = __float2int_rn(ux);
    if ((unsigned)ny >= (unsigned)H || (unsigned)nx >= (unsigned)W) return;
    for (int c = 0; c < C; ++c) atomicAdd(gsrc + (b * C + c) * HW + (size_t)ny * W + nx, __ldg(go + (b * C + c) * HW + pix));
}

// ---------------------------------------------------------------------------------------------- VecInt
// out = s*v + warp(s*v, flow = s*v), s a power of two (exact), both channels per thread.
__global__ void __launch_bounds__(256) vecint_step_kernel(const float* __restrict__ v, float* __restrict__ out,
                                                          float s, int H, int W, float ry, float rx, long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int x = (int)(t % W);
    const long row = t / W;
    const int y = (int)(row % H);
    const long b = row / H;
    const size_t HW = (size_t)H * W, pix = (size_t)y * W + x;
    const float* v0 = v + (b * 2 + 0) * HW;
    const float* v1 = v + (b * 2 + 1) * HW;
    const float dy = __fmul_rn(__ldg(v0 + pix), s), dx = __fmul_rn(__ldg(v1 + pix), s);
    const float uy = warp_coord(__fadd_rn((float)y, dy), (float)(H - 1), ry);
    const float ux = warp_coord(__fadd_rn((float)x, dx), (float)(W - 1), rx);
    const Tap tp = make_tap(uy, ux);
    const bool y0in = (unsigned)tp.y0 < (unsigned)H, y1in = (unsigned)(tp.y0 + 1) < (unsigned)H;
    const bool x0in = (unsigned)tp.x0 < (unsigned)W, x1in = (unsigned)(tp.x0 + 1) < (unsigned)W;
    const size_t o00 = (size_t)tp.y0 * W + tp.x0;
    float r[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const float* img = c == 0 ? v0 : v1;
        const float nw = (y0in && x0in) ? __fmul_rn(__ldg(img + o00), s) : 0.0f;
        const float ne = (y0in && x1in) ? __fmul_rn(__ldg(img + o00 + 1), s) : 0.0f;
        const float sw = (y1in && x0in) ? __fmul_rn(__ldg(img + o00 + W), s) : 0.0f;
        const float se = (y1in && x1in) ? __fmul_rn(__ldg(img + o00 + W + 1), s) : 0.0f;
        float acc = __fmul_rn(nw, tp.w00);
        acc = fmaf(ne, tp.w01, acc);
        acc = fmaf(sw, tp.w10, acc);
        acc = fmaf(se, tp.w11, acc);
        r[c] = acc;
    }
    out[(b * 2 + 0) * HW + pix] = __fadd_rn(dy, r[0]);
    out[(b * 2 + 1) * HW + pix] = __fadd_rn(dx, r[1]);
}

__global__ void __launch_bounds__(256) scale_kernel(const float* __restrict__ v, float* __restrict__ out, float s,
                                                    long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t < total) out[t] = __fmul_rn(__ldg(v + t), s);
}

// ---------------------------------------------------------------------------------------------- point sampling
// locs [B,2,P]: ch0 = x normalised by (SW-1), ch1 = y normalised by (SH-1) (the reference's `shape[~i]`), then
// grid_sample on the [H,W] image.  One thread per (b, point), all channels.
template <int MODE>
__global__ void __launch_bounds__(128) sample_points_kernel(const float* __restrict__ locs, const float* __restrict__ src,
                                                            float* __restrict__ out, int C, int H, int W, int P,
                                                            float sn_x, float sn_y, long total) {
    const long t = (long)blockIdx.x * 128 + threadIdx.x;
    if (t >= total) return;
    const long b = t / P;
    const int p = (int)(t - b * P);
    const float px = __ldg(locs + (b * 2 + 0) * P + p), py = __ldg(locs + (b * 2 + 1) * P + p);
    // normalise with the transformer's size, un-normalise with the image's (ATen)
    const float nx = __fmul_rn(2.0f, __fsub_rn(__fdiv_rn(px, sn_x), 0.5f));
    const float ny = __fmul_rn(2.0f, __fsub_rn(__fdiv_rn(py, sn_y), 0.5f));
    const float ux = __fmul_rn(__fmul_rn(__fadd_rn(nx, 1.0f), 0.5f), (float)(W - 1));
    const float uy = __fmul_rn(__fmul_rn(__fadd_rn(ny, 1.0f), 0.5f), (float)(H - 1));
    const size_t HW = (size_t)H * W;
    Tap tp;
    int qy = 0, qx = 0;
    if (MODE == CF_MODE_BILINEAR) {
        tp = make_tap(uy, ux);
    } else {
        qy = __float2int_rn(uy);
        qx = __float2int_rn(ux);
    }
    for (int c = 0; c < C; ++c) {
        const float* img = src + (b * C + c) * HW;
        out[(b * C + c) * P + p] = (MODE == CF_MODE_BILINEAR) ? sample_bilinear(img, H, W, tp) : fetch(img, H, W, qy, qx);
    }
}

// Backward of the point sampling: the reference back-propagates an MSE point loss through SpatialTransformerContour into
// `original` = the predicted flow (training/network_training/SegFlowGaussian.py:1497-1503).  One thread per (b, point):
// d/d(src) scattered with atomics (two points may share a texel), d/d(locs) written (ATen grid_sampler_2d_backward chained
// through the reference's 2 * (p / (S - 1) - 0.5): d(u)/d(p) = ((size - 1) / 2) * 2 / (S - 1)).
template <int MODE>
__global__ void __launch_bounds__(128) sample_points_bwd_kernel(const float* __restrict__ locs, const float* __restrict__ src,
                                                                const float* __restrict__ go, float* __restrict__ gsrc,
                                                                float* __restrict__ glocs, int C, int H, int W, int P,
                                                                float sn_x, float sn_y, long total) {
    const long t = (long)blockIdx.x * 128 + threadIdx.x;
    if (t >= total) return;
    const long b = t / P;
    const int p = (int)(t - b * P);
    const float px = __ldg(locs + (b * 2 + 0) * P + p), py = __ldg(locs + (b * 2 + 1) * P + p);
    const float nx = __fmul_rn(2.0f, __fsub_rn(__fdiv_rn(px, sn_x), 0.5f));
    const float ny = __fmul_rn(2.0f, __fsub_rn(__fdiv_rn(py, sn_y), 0.5f));
    const float ux = __fmul_rn(__fmul_rn(__fadd_rn(nx, 1.0f), 0.5f), (float)(W - 1));
    const float uy = __fmul_rn(__fmul_rn(__fadd_rn(ny, 1.0f), 0.5f), (float)(H - 1));
    const size_t HW = (size_t)H * W;
    if (MODE == CF_MODE_NEAREST) {
        const int qy = __float2int_rn(uy), qx = __float2int_rn(ux);
        const bool in = (unsigned)qy < (unsigned)H && (unsigned)qx < (unsigned)W;
        if (gsrc && in)
            for (int c = 0; c < C; ++c) atomicAdd(gsrc + (b * C + c) * HW + (size_t)qy * W + qx, __ldg(go + (b * C + c) * P + p));
        if (glocs) { glocs[(b * 2 + 0) * P + p] = 0.0f; glocs[(b * 2 + 1) * P + p] = 0.0f; }
        return;
    }
    const Tap tp = make_tap(uy, ux);
    const float fx = floorf(ux), fy = floorf(uy);
    const float dx_lo = __fsub_rn(__fadd_rn(fx, 1.0f), ux), dx_hi = __fsub_rn(ux, fx);
    const float dy_lo = __fsub_rn(__fadd_rn(fy, 1.0f), uy), dy_hi = __fsub_rn(uy, fy);
    const bool y0in = (unsigned)tp.y0 < (unsigned)H, y1in = (unsigned)(tp.y0 + 1) < (unsigned)H;
    const bool x0in = (unsigned)tp.x0 < (unsigned)W, x1in = (unsigned)(tp.x0 + 1) < (unsigned)W;
    const size_t o00 = (size_t)tp.y0 * W + tp.x0;
    float gix = 0.0f, giy = 0.0f;
    for (int c = 0; c < C; ++c) {
        const size_t chan = (b * C + c) * HW;
        const float g = __ldg(go + (b * C + c) * P + p);
        if (gsrc) {
            float* gs = gsrc + chan;
            if (y0in && x0in) atomicAdd(gs + o00, tp.w00 * g);
            if (y0in && x1in) atomicAdd(gs + o00 + 1, tp.w01 * g);
            if (y1in && x0in) atomicAdd(gs + o00 + W, tp.w10 * g);
            if (y1in && x1in) atomicAdd(gs + o00 + W + 1, tp.w11 * g);
        }
        if (glocs) {
            const float* img = src + chan;
            const float nw = (y0in && x0in) ? __ldg(img + o00) : 0.0f;
            const float ne = (y0in && x1in) ? __ldg(img + o00 + 1) : 0.0f;
            const float sw = (y1in && x0in) ? __ldg(img + o00 + W) : 0.0f;
            const float se = (y1in && x1in) ? __ldg(img + o00 + W + 1) : 0.0f;
            gix -= nw * dy_lo * g;  giy -= nw * dx_lo * g;
            gix += ne * dy_lo * g;  giy -= ne * dx_hi * g;
            gix -= sw * dy_hi * g;  giy += sw * dx_lo * g;
            gix += se * dy_hi * g;  giy += se * dx_hi * g;
        }
    }
    if (glocs) {
        glocs[(b * 2 + 0) * P + p] = __fdiv_rn(__fmul_rn(__fmul_rn(gix, __fmul_rn((float)(W - 1), 0.5f)), 2.0f), sn_x);
        glocs[(b * 2 + 1) * P + p] = __fdiv_rn(__fmul_rn(__fmul_rn(giy, __fmul_rn((float)(H - 1), 0.5f)), 2.0f), sn_y);
    }
}

}  // namespace

int launch_sample_points_bwd(const float* locs, const float* src, const float* go, int B, int C, int H, int W, int P, int SH,
                             int SW, int mode, float* gsrc, float* glocs, cudaStream_t s) {
    const long total = (long)B * P;
    const float snx = (float)(SW - 1), sny = (float)(SH - 1);
    if (mode == CF_MODE_BILINEAR)
        sample_points_bwd_kernel<CF_MODE_BILINEAR><<<cdiv(total, 128), 128, 0, s>>>(locs, src, go, gsrc, glocs, C, H, W, P, snx, sny, total);
    else
        sample_points_bwd_kernel<CF_MODE_NEAREST><<<cdiv(total, 128), 128, 0, s>>>(locs, src, go, gsrc, glocs, C, H, W, P, snx, sny, total);
    CF_CHECK_LAUNCH("sample_points_bwd_kernel");
    return CF_OK;
}

int launch_warp2d_bwd(const float* flow, const float* src, const float* go, int B, int C, int H, int W, float* gsrc,
                      float* gflow, cudaStream_t s) {
    const long total = (long)B * H * W;
    warp2d_bwd_kernel<false><<<cdiv(total, 256), 256, 0, s>>>(flow, src, go, gsrc, gflow, C, H, W, 1.0f / (float)(H - 1), 1.0f / (float)(W - 1), total);
    CF_CHECK_LAUNCH("warp2d_bwd_kernel");
    return CF_OK;
}

int launch_warp2d_nearest_bwd(const float* flow, const float* go, int B, int C, int H, int W, float* gsrc, cudaStream_t s) {
    const long total = (long)B * H * W;
    warp2d_nearest_bwd_kernel<<<cdiv(total, 256), 256, 0, s>>>(flow, go, gsrc, C, H, W, 1.0f / (float)(H - 1), 1.0f / (float)(W - 1), total);
    CF_CHECK_LAUNCH("warp2d_nearest_bwd_kernel");
    return CF_OK;
}

int launch_warp2d_xy_bwd(const float* flow, const float* src, const float* go, int B, int C, int H, int W, float* gsrc,
                         float* gflow, cudaStream_t s) {
    const long total = (long)B * H * W;
    warp2d_bwd_kernel<true><<<cdiv(total, 256), 256, 0, s>>>(flow, src, go, gsrc, gflow, C, H, W, 0.0f, 0.0f, total);
    CF_CHECK_LAUNCH("warp2d_bwd_kernel<xy>");
    return CF_OK;
}

int launch_vecint(const float* vec, int B, int H, int W, int nsteps, float* out, float* work, cudaStream_t s) {
    const long total = (long)B * H * W;
    const float scale = ldexpf(1.0f, -nsteps);
    if (nsteps == 0) {
        scale_kernel<<<cdiv(total * 2, 256), 256, 0, s>>>(vec, out, 1.0f, total * 2);
        CF_CHECK_LAUNCH("scale_kernel");
        return CF_OK;
    }
    // ping-pong so that the last step lands in `out`: step k (0-based) writes out when (nsteps - 1 - k) is even
    const float* cur = vec;
    for (int k = 0; k < nsteps; ++k) {
        float* dst = ((nsteps - 1 - k) % 2 == 0) ? out : work;
        vecint_step_kernel<<<cdiv(total, 256), 256, 0, s>>>(cur, dst, k == 0 ? scale : 1.0f, H, W, 1.0f / (float)(H - 1), 1.0f / (float)(W - 1), total);
        CF_CHECK_LAUNCH("vecint_step_kernel");
        cur = dst;
    }
    return CF_OK;
}

int launch_sample_points(const float* locs, const float* src, int B, int C, int H, int W, int P, int SH, int SW,
                         int mode, float* out, cudaStream_t s) {
    const long total = (long)B * P;
    const float snx = (float)(SW - 1), sny = (float)(SH - 1);
    if (mode == CF_MODE_BILINEAR)
        sample_points_kernel<CF_MODE_BILINEAR><<<cdiv(total, 128), 128, 0, s>>>(locs, src, out, C, H, W, P, snx, sny, total);
    else
        sample_points_kernel<CF_MODE_NEAREST><<<cdiv(total, 128), 128, 0, s>>>(locs, src, out, C, H, W, P, snx, sny, total);
    CF_CHECK_LAUNCH("sample_points_kernel");
    return CF_OK;
}

// ---------------------------------------------------------------------------------------------- lib/utils.warp
// SURVEY.md 8(f) row 4 (second half): /root/reference/nnunet/lib/utils.py:29-59 `warp(x, flo)`, used at
// network_architecture/SegFlowGaussian.py:1044.  Unlike SpatialTransformer: flow channel 0 displaces COLUMNS, channel 1
// rows; n = 2*(p + f)/max(S-1, 1) - 1; grid_sample with its defaults (bilinear, zeros, align_corners=False), whose
// un-normalisation is ((n + 1) * S - 1) / 2 -- coordinates land half a pixel off the align_corners=True warp.
// Same fp32 operation order as the reference -> bit-exact taps.  lane = column, blockIdx.y = row, blockIdx.z = item.
namespace {
__device__ __forceinline__ float utils_warp_coord(float p, float d1 /* max(S-1,1) */, float size) {
    const float n = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, p), d1), 1.0f);
    return __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(n, 1.0f), size), 1.0f), 0.5f);
}

__global__ void __launch_bounds__(128) warp2d_xy_kernel(const float* __restrict__ flow, const float* __restrict__ src,
                                                        float* __restrict__ out, int B, int C, int H, int W) {
    const int x = blockIdx.x * 128 + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const unsigned HW = (unsigned)(H * W), pix = (unsigned)y * (unsigned)W + (unsigned)x;
    const float dx1 = (float)max(W - 1, 1), dy1 = (float)max(H - 1, 1);
    for (unsigned b = blockIdx.z; b < (unsigned)B; b += gridDim.z) {
        const float* fl = flow + (size_t)b * 2 * HW + pix;
        const float ux = utils_warp_coord(__fadd_rn((float)x, __ldg(fl)), dx1, (float)W);
        const float uy = utils_warp_coord(__fadd_rn((float)y, __ldg(fl + HW)), dy1, (float)H);
        const Tap tp = make_tap(uy, ux);
        const float* img = src + (size_t)b * C * HW;
        float* o = out + (size_t)b * C * HW + pix;
        for (int c = 0; c < C; ++c, img += HW, o += HW) *o = sample_bilinear(img, H, W, tp);
    }
}
}  // namespace

int launch_warp2d_xy(const float* flow, const float* src, int B, int C, int H, int W, float* out, cudaStream_t s) {
    if ((long)H * W >= (1L << 30) || H > 65535) return fail(CF_ERR_UNSUPPORTED, "warp2d_xy: image of %dx%d pixels too large", H, W);
    const dim3 grid((unsigned)cdiv(W, 128), (unsigned)H, (unsigned)(B < 65535 ? B : 65535));
    warp2d_xy_kernel<<<grid, 128, 0, s>>>(flow, src, out, B, C, H, W);
    CF_CHECK_LAUNCH("warp2d_xy_kernel");
    return CF_OK;
}

}  // namespace cf
