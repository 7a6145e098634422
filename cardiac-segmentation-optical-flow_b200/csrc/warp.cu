// K5: dense bilinear warp (SpatialTransformer.forward), its fused one-hot/argmax variant, coords_grid.
//
// Replaces /root/reference/nnunet/network_architecture/integration.py:61-79, which on a GPU runs as
// add + 2 normalise kernels + permute + channel-swap gather + grid_sampler_2d (6 launches, ~4x the
// compulsory traffic on the coordinate tensor).  Here: one launch, flow read once (128-bit), all C channels
// sampled per pixel from the taps computed once, 128-bit stores.  HBM-bound: 4*H*W*(2C+2) bytes per item.
#include "common.cuh"
#include "sampling.cuh"

namespace cf {
namespace {

template <int VEC, int MODE>
__global__ void __launch_bounds__(256) warp2d_kernel(const float* __restrict__ flow, const float* __restrict__ src,
                                                     float* __restrict__ out, int C, int H, int W, float ry, float rx,
                                                     long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int wv = W / VEC;
    const int x = (int)(t % wv) * VEC;
    const long row = t / wv;
    const int y = (int)(row % H);
    const long b = row / H;
    const size_t HW = (size_t)H * W, pix = (size_t)y * W + x;
    const float s1y = (float)(H - 1), s1x = (float)(W - 1);

    float dy[VEC], dx[VEC];
    const float* fyp = flow + (b * 2 + 0) * HW + pix;
    const float* fxp = flow + (b * 2 + 1) * HW + pix;
    if (VEC == 4) {
        const float4 a = ldg_f4(fyp), c = ldg_f4(fxp);
        dy[0] = a.x; dy[1 % VEC] = a.y; dy[2 % VEC] = a.z; dy[3 % VEC] = a.w;
        dx[0] = c.x; dx[1 % VEC] = c.y; dx[2 % VEC] = c.z; dx[3 % VEC] = c.w;
    } else {
        dy[0] = __ldg(fyp);
        dx[0] = __ldg(fxp);
    }

    Tap tap[VEC];
    int ny[VEC], nx[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
        const float uy = warp_coord(__fadd_rn((float)y, dy[v]), s1y, ry);
        const float ux = warp_coord(__fadd_rn((float)(x + v), dx[v]), s1x, rx);
        if (MODE == CF_MODE_BILINEAR) {
            tap[v] = make_tap(uy, ux);
        } else {  // ATen nearest: nearbyint = round half to even
            ny[v] = __float2int_rn(uy);
            nx[v] = __float2int_rn(ux);
        }
    }

    for (int c = 0; c < C; ++c) {
        const float* img = src + (b * C + c) * HW;
        float r[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v)
            r[v] = (MODE == CF_MODE_BILINEAR) ? sample_bilinear(img, H, W, tap[v]) : fetch(img, H, W, ny[v], nx[v]);
        float* o = out + (b * C + c) * HW + pix;
        if (VEC == 4) {
            *reinterpret_cast<float4*>(o) = make_float4(r[0], r[1 % VEC], r[2 % VEC], r[3 % VEC]);
        } else {
            o[0] = r[0];
        }
    }
}

__global__ void __launch_bounds__(256) warp2d_taps_kernel(const float* __restrict__ flow, float* __restrict__ taps,
                                                          int H, int W, float ry, float rx, long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int x = (int)(t % W);
    const long row = t / W;
    const int y = (int)(row % H);
    const long b = row / H;
    const size_t HW = (size_t)H * W, pix = (size_t)y * W + x;
    const float dy = flow[(b * 2 + 0) * HW + pix], dx = flow[(b * 2 + 1) * HW + pix];
    taps[(b * 2 + 0) * HW + pix] = warp_coord(__fadd_rn((float)y, dy), (float)(H - 1), ry);
    taps[(b * 2 + 1) * HW + pix] = warp_coord(__fadd_rn((float)x, dx), (float)(W - 1), rx);
}

// a6: argmax_c warp(one_hot(label))[c] without materialising the one-hot tensor.  The per-class value is
// the sum, in ATen order, of the weights of the in-image taps whose label equals c; first maximum wins.
__global__ void __launch_bounds__(256) warp2d_onehot_argmax_kernel(const float* __restrict__ flow,
                                                                    const int64_t* __restrict__ label,
                                                                    int64_t* __restrict__ out, int ncls, int H, int W,
                                                                    float ry, float rx, long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int x = (int)(t % W);
    const long row = t / W;
    const int y = (int)(row % H);
    const long b = row / H;
    const size_t HW = (size_t)H * W, pix = (size_t)y * W + x;
    const float uy = warp_coord(__fadd_rn((float)y, __ldg(flow + (b * 2 + 0) * HW + pix)), (float)(H - 1), ry);
    const float ux = warp_coord(__fadd_rn((float)x, __ldg(flow + (b * 2 + 1) * HW + pix)), (float)(W - 1), rx);
    const Tap tp = make_tap(uy, ux);
    const float w[4] = {tp.w00, tp.w01, tp.w10, tp.w11};
    const int ty[4] = {tp.y0, tp.y0, tp.y0 + 1, tp.y0 + 1};
    const int tx[4] = {tp.x0, tp.x0 + 1, tp.x0, tp.x0 + 1};
    long lab[4];
    const int64_t* lb = label + b * HW;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        lab[k] = ((unsigned)ty[k] < (unsigned)H && (unsigned)tx[k] < (unsigned)W) ? __ldg(lb + (size_t)ty[k] * W + tx[k])
                                                                               : -1;
    long best = 0;
    float bestv = -INFINITY;
    for (int c = 0; c < ncls; ++c) {
        float acc = 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (lab[k] == c) acc = __fadd_rn(acc, w[k]);
        if (acc > bestv) {
            bestv = acc;
            best = c;
        }
    }
    out[b * HW + pix] = best;
}

__global__ void __launch_bounds__(256) coords_grid_kernel(float* __restrict__ out, int H, int W, long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int x = (int)(t % W);
    const long r = t / W;
    const int y = (int)(r % H);
    const long ch = (r / H) & 1;  // layout [B,2,H,W]: channel index is the next-slower axis
    out[t] = ch == 0 ? (float)x : (float)y;
}

}  // namespace

int launch_coords_grid(float* out, int B, int H, int W, cudaStream_t s) {
    const long total = (long)B * 2 * H * W;
    coords_grid_kernel<<<cdiv(total, 256), 256, 0, s>>>(out, H, W, total);
    CF_CHECK_LAUNCH("coords_grid_kernel");
    return CF_OK;
}

int launch_warp2d(const float* flow, const float* src, int B, int C, int H, int W, int mode, float* out,
                  cudaStream_t s) {
    const bool vec = (W % 4 == 0) && aligned16(flow) && aligned16(src) && aligned16(out);
    const long total = (long)B * H * (vec ? W / 4 : W);
    const int grid = cdiv(total, 256);
    const float ry = 1.0f / (float)(H - 1), rx = 1.0f / (float)(W - 1);   // IEEE-rounded reciprocals for div_small
    if (mode == CF_MODE_BILINEAR) {
        if (vec) warp2d_kernel<4, CF_MODE_BILINEAR><<<grid, 256, 0, s>>>(flow, src, out, C, H, W, ry, rx, total);
        else warp2d_kernel<1, CF_MODE_BILINEAR><<<grid, 256, 0, s>>>(flow, src, out, C, H, W, ry, rx, total);
    } else {
        if (vec) warp2d_kernel<4, CF_MODE_NEAREST><<<grid, 256, 0, s>>>(flow, src, out, C, H, W, ry, rx, total);
        else warp2d_kernel<1, CF_MODE_NEAREST><<<grid, 256, 0, s>>>(flow, src, out, C, H, W, ry, rx, total);
    }
    CF_CHECK_LAUNCH("warp2d_kernel");
    return CF_OK;
}

int launch_warp2d_taps(const float* flow, int B, int H, int W, float* taps, cudaStream_t s) {
    const long total = (long)B * H * W;
    warp2d_taps_kernel<<<cdiv(total, 256), 256, 0, s>>>(flow, taps, H, W, 1.0f / (float)(H - 1), 1.0f / (float)(W - 1), total);
    CF_CHECK_LAUNCH("warp2d_taps_kernel");
    return CF_OK;
}

int launch_warp2d_onehot_argmax(const float* flow, const int64_t* label, int ncls, int B, int H, int W,
                                int64_t* out, cudaStream_t s) {
    const long total = (long)B * H * W;
    warp2d_onehot_argmax_kernel<<<cdiv(total, 256), 256, 0, s>>>(flow, label, out, ncls, H, W, 1.0f / (float)(H - 1), 1.0f / (float)(W - 1), total);
    CF_CHECK_LAUNCH("warp2d_onehot_argmax_kernel");
    return CF_OK;
}

}  // namespace cf
