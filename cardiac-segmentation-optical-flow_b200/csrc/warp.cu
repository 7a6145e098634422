// K5: dense bilinear warp (SpatialTransformer.forward), its fused one-hot/argmax variant, coords_grid.
//
// Replaces /root/reference/nnunet/network_architecture/integration.py:61-79, which on a GPU runs as
// add + 2 normalise kernels + permute + channel-swap gather + grid_sampler_2d (6 launches, ~4x the
// compulsory traffic on the coordinate tensor).  Here: one launch, flow read once (128-bit), all C channels
// sampled per pixel from the taps computed once, 128-bit stores.  HBM-bound: 4*H*W*(2C+2) bytes per item.
#include <cstdlib>

#include "common.cuh"
#include "sampling.cuh"

namespace cf {
namespace {

// One thread = VEC horizontally adjacent pixels of one image, all C channels.  Everything that does not depend on the
// channel (coordinates, weights, tap offsets, in-image masks) is computed once per pixel; the channel loop is 4
// predicated loads + 1 multiply + 3 FMAs per pixel, 32-bit index arithmetic throughout.
template <int VEC, int MODE>
__global__ void __launch_bounds__(256, VEC == 2 ? 5 : (VEC == 4 ? 3 : 1)) warp2d_kernel(const float* __restrict__ flow, const float* __restrict__ src,
                                                     float* __restrict__ out, int B, int C, int H, int W, float ry, float rx,
                                                     int wv_shift) {
    const unsigned wv = (unsigned)W / VEC;
    const unsigned t = blockIdx.x * 256u + threadIdx.x;          // position inside one image
    if (t >= (unsigned)H * wv) return;
    const unsigned y = wv_shift >= 0 ? (t >> wv_shift) : t / wv;
    const unsigned x = (t - y * wv) * VEC;
    const unsigned HW = (unsigned)(H * W), pix = y * (unsigned)W + x;
    const float s1y = (float)(H - 1), s1x = (float)(W - 1);

    for (unsigned b = blockIdx.y; b < (unsigned)B; b += gridDim.y) {
        float dy[VEC], dx[VEC];
        const float* fyp = flow + ((size_t)b * 2 + 0) * HW + pix;
        const float* fxp = fyp + HW;
        if (VEC == 4) {
            const float4 a = ldg_f4(fyp), c = ldg_f4(fxp);
            dy[0] = a.x; dy[1 % VEC] = a.y; dy[2 % VEC] = a.z; dy[3 % VEC] = a.w;
            dx[0] = c.x; dx[1 % VEC] = c.y; dx[2 % VEC] = c.z; dx[3 % VEC] = c.w;
        } else if (VEC == 2) {
            const float2 a = __ldg(reinterpret_cast<const float2*>(fyp)), c = __ldg(reinterpret_cast<const float2*>(fxp));
            dy[0] = a.x; dy[1 % VEC] = a.y;
            dx[0] = c.x; dx[1 % VEC] = c.y;
        } else {
            dy[0] = __ldg(fyp);
            dx[0] = __ldg(fxp);
        }

        // Tap offsets are clamped into the image and loaded unconditionally (straight-line code, every load of a pixel in
        // flight at once); a tap outside the image is replaced by zero afterwards (zero padding), never multiplied in.
        int off[VEC][4];       // nw, ne, sw, se (bilinear) / [0] = nearest texel
        unsigned msk[VEC];     // bit k: tap k lies inside the image
        float w00[VEC], w01[VEC], w10[VEC], w11[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            const float uy = warp_coord(__fadd_rn((float)y, dy[v]), s1y, ry);
            const float ux = warp_coord(__fadd_rn((float)(x + v), dx[v]), s1x, rx);
            if (MODE == CF_MODE_BILINEAR) {
                const Tap tp = make_tap(uy, ux);
                const bool y0 = (unsigned)tp.y0 < (unsigned)H, y1 = (unsigned)(tp.y0 + 1) < (unsigned)H;
                const bool x0 = (unsigned)tp.x0 < (unsigned)W, x1 = (unsigned)(tp.x0 + 1) < (unsigned)W;
                msk[v] = (y0 && x0 ? 1u : 0u) | (y0 && x1 ? 2u : 0u) | (y1 && x0 ? 4u : 0u) | (y1 && x1 ? 8u : 0u);
                const int r0 = (y0 ? tp.y0 : 0) * W, r1 = (y1 ? tp.y0 + 1 : 0) * W;
                const int c0 = x0 ? tp.x0 : 0, c1 = x1 ? tp.x0 + 1 : 0;
                off[v][0] = r0 + c0; off[v][1] = r0 + c1; off[v][2] = r1 + c0; off[v][3] = r1 + c1;
                w00[v] = tp.w00; w01[v] = tp.w01; w10[v] = tp.w10; w11[v] = tp.w11;
            } else {  // ATen nearest: nearbyint = round half to even
                const int ny = __float2int_rn(uy), nx = __float2int_rn(ux);
                const bool in = (unsigned)ny < (unsigned)H && (unsigned)nx < (unsigned)W;
                msk[v] = in ? 1u : 0u;
                off[v][0] = in ? ny * W + nx : 0;
                off[v][1] = off[v][2] = off[v][3] = 0;
                w00[v] = w01[v] = w10[v] = w11[v] = 0.0f;
            }
        }

        const float* img = src + (size_t)b * C * HW;
        float* o = out + (size_t)b * C * HW + pix;
        for (int c = 0; c < C; ++c, img += HW, o += HW) {
            float r[VEC];
#pragma unroll
            for (int v = 0; v < VEC; ++v) {
                if (MODE == CF_MODE_BILINEAR) {
                    const float l00 = __ldg(img + off[v][0]), l01 = __ldg(img + off[v][1]);
                    const float l10 = __ldg(img + off[v][2]), l11 = __ldg(img + off[v][3]);
                    const float v00 = (msk[v] & 1u) ? l00 : 0.0f, v01 = (msk[v] & 2u) ? l01 : 0.0f;
                    const float v10 = (msk[v] & 4u) ? l10 : 0.0f, v11 = (msk[v] & 8u) ? l11 : 0.0f;
                    float acc = __fmul_rn(v00, w00[v]);      // ATen order: nw, ne, sw, se
                    acc = fmaf(v01, w01[v], acc);
                    acc = fmaf(v10, w10[v], acc);
                    acc = fmaf(v11, w11[v], acc);
                    r[v] = acc;
                } else {
                    const float l00 = __ldg(img + off[v][0]);
                    r[v] = (msk[v] & 1u) ? l00 : 0.0f;
                }
            }
            if (VEC == 4) *reinterpret_cast<float4*>(o) = make_float4(r[0], r[1 % VEC], r[2 % VEC], r[3 % VEC]);
            else if (VEC == 2) *reinterpret_cast<float2*>(o) = make_float2(r[0], r[1 % VEC]);
            else o[0] = r[0];
        }
    }
}

// ---- bilinear, tiled (the production path) ----
// A warp covers 32 consecutive pixels of 2 image rows (lane = column, a thread samples (x, y) and (x, y + 1)), a CTA
// 32 x 16 pixels.  The bound is the L1 wavefront rate of the gathers (ncu: l1tex data-pipe 75-80 % with a 16 x 4 warp
// tile, 20 sectors per request), so lanes run along a row: a tap load of a warp touches 1-2 lines.  The kernel votes
// once per warp: when every tap of every lane is inside the image and every coordinate is in div_small's checked range,
// the channel loop is 4 unpredicated loads at o, o+1, o+W, o+W+1 + 1 FMUL + 3 FFMA per pixel with ONE offset per
// pixel; any other warp (image border, flow pointing outside) takes the general path (bounds-checked taps, IEEE
// division).
// The coordinate round trip is shortened without changing a bit: with t = RN(p/s1 - 0.5), the reference's
// ((2t + 1) * 0.5) * s1 has exact doublings / halvings, so u = RN(RN(2t + 1) * (s1/2)) = fmul(fma(t, 2, 1), s1/2).
__device__ __forceinline__ float warp_coord_lean(float p, float s1, float r, float hs1) {
    const float t = __fsub_rn(div_small(p, s1, r), 0.5f);
    return __fmul_rn(__fmaf_rn(t, 2.0f, 1.0f), hs1);
}

// general path of the tiled kernel (rare: border warps), kept out of line so that it does not set the register budget
__device__ __noinline__ void warp2d_general(const float* __restrict__ img, float* __restrict__ o, int C, int H, int W,
                                            float py, float px, float ry, float rx) {
    const unsigned HW = (unsigned)(H * W);
    const Tap t = make_tap(warp_coord(py, (float)(H - 1), ry), warp_coord(px, (float)(W - 1), rx));
#pragma unroll 1
    for (int c = 0; c < C; ++c, img += HW, o += HW) *o = sample_bilinear(img, H, W, t);
}

constexpr int kTileW = 32, kTileH = 16;   // CTA tile: 8 warps, each 32 pixels x 2 rows (lane = column)

// UNROLL / MINB measured at cfg3 (300 x {1,5} x 224^2): (2,4) 77 / 210 us, (4,4) 102 / 255, (5,3) 96 / 241, (2,5) 73 / 206,
// (1,6) 71 / 196, (1,8) 69 / 175 (round 2, with 9800 CTAs): the kernel wants resident warps (32 registers, 8 CTAs per SM = the
// 64-warp limit), not deeper unrolling.
template <bool FASTDIV, int UNROLL = 1, int MINB = 6>
__global__ void __launch_bounds__(256, MINB) warp2d_tile_kernel(const float* __restrict__ flow, const float* __restrict__ src,
                                                             float* __restrict__ out, int B, int C, int H, int W, float ry,
                                                             float rx, int pf_rows) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int x = blockIdx.x * kTileW + lane;
    const int y = blockIdx.y * kTileH + wid * 2;
    const bool act0 = x < W && y < H, act1 = x < W && y + 1 < H;
    const unsigned HW = (unsigned)(H * W), pix = (unsigned)y * (unsigned)W + (unsigned)x;
    const float s1y = (float)(H - 1), s1x = (float)(W - 1), hs1y = 0.5f * s1y, hs1x = 0.5f * s1x;
    const float yf0 = (float)y, yf1 = (float)(y + 1), xf = (float)x;
    const int wid_u = __shfl_sync(0xffffffffu, wid, 0);      // the warp index as a value the compiler knows to be warp-uniform

    // The chain flow load -> coordinates -> gathers -> store is two DRAM latencies long and a thread keeps only 16 B of
    // flow in flight: the flow of the NEXT image of this thread is loaded before the current one is sampled.
    float nfy0 = 0.f, nfy1 = 0.f, nfx0 = 0.f, nfx1 = 0.f;
    if (blockIdx.z < (unsigned)B) {
        const float* fyp = flow + ((size_t)blockIdx.z * 2) * HW + pix;
        if (act0) { nfy0 = __ldcs(fyp); nfx0 = __ldcs(fyp + HW); }
        if (act1) { nfy1 = __ldcs(fyp + W); nfx1 = __ldcs(fyp + HW + W); }
    }
    for (unsigned b = blockIdx.z; b < (unsigned)B; b += gridDim.z) {
        const float fy0 = nfy0, fy1 = nfy1, fx0 = nfx0, fx1 = nfx1;
        if (b + gridDim.z < (unsigned)B) {
            const float* fyp = flow + ((size_t)(b + gridDim.z) * 2) * HW + pix;
            if (act0) { nfy0 = __ldcs(fyp); nfx0 = __ldcs(fyp + HW); }
            if (act1) { nfy1 = __ldcs(fyp + W); nfx1 = __ldcs(fyp + HW + W); }
        }
        // L2 prefetch of the source rows around this tile for the NEXT image of this CTA (a hint: 128 bytes per row and channel,
        // the tile's own columns -- the margins are the neighbouring CTAs' columns): the gathers below then wait for L2, not DRAM.
        // (Also prefetching the first image of the CTA that follows in this slot was measured and dropped: no gain, more registers.)
        if (pf_rows > 0 && b + gridDim.z < (unsigned)B) {
            // warp w issues rows w, w + 8, ... of every channel: addresses are warp-uniform (the bulk prefetch takes uniform
            // registers; per-lane addresses made the compiler serialise the lanes in a loop)
            const float* nimg = src + (size_t)(b + gridDim.z) * C * HW + blockIdx.x * kTileW;
            const int r0 = (int)(blockIdx.y * kTileH) - (pf_rows - kTileH) / 2;
            const unsigned bytes = (unsigned)min(kTileW, W - (int)(blockIdx.x * kTileW)) * 4u;
            for (int c = 0; c < C; ++c, nimg += HW)
                for (int rr = wid_u; rr < pf_rows; rr += 8) {
                    const int r = r0 + rr;
                    if ((unsigned)r < (unsigned)H)
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nimg + r * W), "r"(bytes) : "memory");
                }
        }
        const float py0 = __fadd_rn(yf0, fy0), py1 = __fadd_rn(yf1, fy1);
        const float px0 = __fadd_rn(xf, fx0), px1 = __fadd_rn(xf, fx1);
        bool ok = FASTDIV && fabsf(py0) < 16384.0f && fabsf(py1) < 16384.0f && fabsf(px0) < 16384.0f && fabsf(px1) < 16384.0f;
        const float uy0 = warp_coord_lean(py0, s1y, ry, hs1y), uy1 = warp_coord_lean(py1, s1y, ry, hs1y);
        const float ux0 = warp_coord_lean(px0, s1x, rx, hs1x), ux1 = warp_coord_lean(px1, s1x, rx, hs1x);
        const float gy0 = floorf(uy0), gy1 = floorf(uy1), gx0 = floorf(ux0), gx1 = floorf(ux1);
        const int iy0 = (int)gy0, iy1 = (int)gy1, ix0 = (int)gx0, ix1 = (int)gx1;
        ok = ok && (unsigned)iy0 < (unsigned)(H - 1) && (unsigned)iy1 < (unsigned)(H - 1) &&
             (unsigned)ix0 < (unsigned)(W - 1) && (unsigned)ix1 < (unsigned)(W - 1);
        const float* img = src + (size_t)b * C * HW;
        float* o = out + (size_t)b * C * HW + pix;
        if (__all_sync(0xffffffffu, ok && act1)) {
            // weights in make_tap's order of operations
            const float ah0 = __fsub_rn(ux0, gx0), al0 = __fsub_rn(__fadd_rn(gx0, 1.0f), ux0);
            const float bh0 = __fsub_rn(uy0, gy0), bl0 = __fsub_rn(__fadd_rn(gy0, 1.0f), uy0);
            const float ah1 = __fsub_rn(ux1, gx1), al1 = __fsub_rn(__fadd_rn(gx1, 1.0f), ux1);
            const float bh1 = __fsub_rn(uy1, gy1), bl1 = __fsub_rn(__fadd_rn(gy1, 1.0f), uy1);
            const float w00a = __fmul_rn(al0, bl0), w01a = __fmul_rn(ah0, bl0), w10a = __fmul_rn(al0, bh0), w11a = __fmul_rn(ah0, bh0);
            const float w00b = __fmul_rn(al1, bl1), w01b = __fmul_rn(ah1, bl1), w10b = __fmul_rn(al1, bh1), w11b = __fmul_rn(ah1, bh1);
            const float* pa = img + (iy0 * W + ix0);
            const float* pb = img + (iy1 * W + ix1);
#pragma unroll UNROLL
            for (int c = 0; c < C; ++c, pa += HW, pb += HW, o += HW) {
                const float a00 = __ldg(pa), a01 = __ldg(pa + 1), a10 = __ldg(pa + W), a11 = __ldg(pa + W + 1);
                const float b00 = __ldg(pb), b01 = __ldg(pb + 1), b10 = __ldg(pb + W), b11 = __ldg(pb + W + 1);
                float ra = __fmul_rn(a00, w00a);          // ATen order: nw, ne, sw, se
                ra = fmaf(a01, w01a, ra);
                ra = fmaf(a10, w10a, ra);
                ra = fmaf(a11, w11a, ra);
                float rb = __fmul_rn(b00, w00b);
                rb = fmaf(b01, w01b, rb);
                rb = fmaf(b10, w10b, rb);
                rb = fmaf(b11, w11b, rb);
                __stcs(o, ra);
                __stcs(o + W, rb);
            }
        } else {
            // general path: the reference's sequence with the range-checked division and per-tap bounds checks
            if (act0) warp2d_general(img, o, C, H, W, py0, px0, ry, rx);
            if (act1) warp2d_general(img, o + W, C, H, W, py1, px1, ry, rx);
        }
    }
}

// Measured and dropped (round 2, profiles/README.md "warp: staged patch"): staging each channel's source patch of the tile in shared
// memory (bounding box of the CTA's taps, 16-byte cp.async, three images in flight per CTA, taps from LDS) is bit-identical but
// slower at cfg3 C = 5: 237-325 us against 184-196 us for the direct gathers below -- the box of a 32 x 16 tile under a smooth 4-pixel
// flow is about 22 x 46 words, the two patch sets cost the resident warps (4 CTAs per SM) and every image costs two CTA barriers.

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

// a6, tiled: same thread mapping as warp2d_tile_kernel (lane = column, two rows per thread, 3-D grid, 32-bit index math,
// lean coordinate round trip when the coordinate is in div_small's range).  The first version spent ~280 instructions per
// pixel on 64-bit div / mod index arithmetic (ncu: 65 % issue-slot utilisation at 30 % of HBM).
// COUNT: also accumulate counts[b][c] = number of pixels of image b whose warped label is c (the per-frame class areas the
// cohort scripts derive from the warped masks): one ballot per class per warp, one global atomic per (warp, class present).
// Whole warps stay alive in that variant (columns beyond W are predicated off instead of returning).
// OutT: int64 (the reference's argmax dtype) or uint8 (registered label maps as the cohort scripts store and gather them).
template <bool FASTDIV, bool COUNT, typename OutT = int64_t>
__global__ void __launch_bounds__(256) warp2d_onehot_argmax_tile_kernel(const float* __restrict__ flow,
                                                                         const int64_t* __restrict__ label,
                                                                         OutT* __restrict__ out, int ncls, int B, int H,
                                                                         int W, float ry, float rx,
                                                                         unsigned long long* __restrict__ counts,
                                                                         int counts_total = 0 /* 1: one [ncls] row for the whole batch */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int x = blockIdx.x * kTileW + lane;
    const int y0 = blockIdx.y * kTileH + wid * 2;
    const bool xin = x < W;
    if (!COUNT && (y0 >= H || !xin)) return;      // COUNT: the whole CTA stays alive for the barriers of the count flush
    __shared__ unsigned cta_cnt[32];               // COUNT, ncls <= 32: per-CTA class counts, one global atomic per class and image
    const bool cta_agg = COUNT && ncls <= 32;
    if (cta_agg) {
        if (threadIdx.x < 32) cta_cnt[threadIdx.x] = 0u;
        __syncthreads();
    }
    const unsigned HW = (unsigned)(H * W);
    const float s1y = (float)(H - 1), s1x = (float)(W - 1), hs1y = 0.5f * s1y, hs1x = 0.5f * s1x;
    const float xf = (float)x;
    const int rows = y0 >= H ? 0 : ((y0 + 1 < H) ? 2 : 1);
    for (unsigned b = blockIdx.z; b < (unsigned)B; b += gridDim.z) {
        const float* fl = flow + (size_t)b * 2 * HW;
        const int64_t* lb = label + (size_t)b * HW;
        OutT* ob = out + (size_t)b * HW;
        int won[2] = {-1, -1};
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            if (rr >= rows) break;
            if (COUNT && !xin) continue;
            const int y = y0 + rr;
            const unsigned pix = (unsigned)y * (unsigned)W + (unsigned)x;
            const float py = __fadd_rn((float)y, __ldcs(fl + pix)), px = __fadd_rn(xf, __ldcs(fl + HW + pix));
            const bool lean = FASTDIV && fabsf(py) < 16384.0f && fabsf(px) < 16384.0f;
            const float uy = lean ? warp_coord_lean(py, s1y, ry, hs1y) : warp_coord(py, s1y, ry);
            const float ux = lean ? warp_coord_lean(px, s1x, rx, hs1x) : warp_coord(px, s1x, rx);
            const Tap tp = make_tap(uy, ux);
            const float w[4] = {tp.w00, tp.w01, tp.w10, tp.w11};
            const bool yin0 = (unsigned)tp.y0 < (unsigned)H, yin1 = (unsigned)(tp.y0 + 1) < (unsigned)H;
            const bool xin0 = (unsigned)tp.x0 < (unsigned)W, xin1 = (unsigned)(tp.x0 + 1) < (unsigned)W;
            const int o = tp.y0 * W + tp.x0;
            long lab[4];
            lab[0] = (yin0 && xin0) ? __ldg(lb + o) : -1;
            lab[1] = (yin0 && xin1) ? __ldg(lb + o + 1) : -1;
            lab[2] = (yin1 && xin0) ? __ldg(lb + o + W) : -1;
            lab[3] = (yin1 && xin1) ? __ldg(lb + o + W + 1) : -1;
            int cls[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) cls[k] = (lab[k] >= 0 && lab[k] < (long)ncls) ? (int)lab[k] : -1;
            int best = 0;
            float bestv = -INFINITY;
            for (int c = 0; c < ncls; ++c) {
                float acc = 0.0f;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (cls[k] == c) acc = __fadd_rn(acc, w[k]);
                if (acc > bestv) {
                    bestv = acc;
                    best = c;
                }
            }
            if (sizeof(OutT) == 8) __stcs(reinterpret_cast<long long*>(ob + pix), (long long)best);
            else ob[pix] = (OutT)best;
            won[rr] = best;
        }
        if (COUNT) {
            for (int c = 0; c < ncls; ++c) {
                const int n = __popc(__ballot_sync(0xffffffffu, won[0] == c)) + __popc(__ballot_sync(0xffffffffu, won[1] == c));
                if (lane == 0 && n) {
                    if (cta_agg) atomicAdd(&cta_cnt[c], (unsigned)n);
                    else atomicAdd(counts + (size_t)(counts_total ? 0 : b) * ncls + c, (unsigned long long)n);
                }
            }
            if (cta_agg) {      // 8 warps -> one global atomic per class present in this CTA's tile (was: one per warp)
                __syncthreads();
                if (threadIdx.x < ncls) {
                    const unsigned n = cta_cnt[threadIdx.x];
                    cta_cnt[threadIdx.x] = 0u;
                    if (n) atomicAdd(counts + (size_t)(counts_total ? 0 : b) * ncls + threadIdx.x, (unsigned long long)n);
                }
                __syncthreads();
            }
        }
    }
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
    if ((long)H * W >= (1L << 30)) return fail(CF_ERR_UNSUPPORTED, "warp2d: image of %dx%d pixels too large", H, W);
    static const int want = [] { const char* e = getenv("CF_WARP_VEC"); return e ? atoi(e) : 0; }();   // tuning knob: 0 = tiled kernel
    const float ry_ = 1.0f / (float)(H - 1), rx_ = 1.0f / (float)(W - 1);   // IEEE-rounded reciprocals for div_small
    if (want == 0 && mode == CF_MODE_BILINEAR) {
        const int tiles_x = cdiv(W, kTileW), tiles_y = cdiv(H, kTileH);
        // a thread's setup (tile position, constants) is amortised over a few images once the grid fills the machine
        static const int target = [] { const char* e = getenv("CF_WARP_CTAS"); return e ? atoi(e) : 9800; }();   // CTAs of about 3 images each; measured at cfg3 C = 5 (6 resident CTAs per SM): 882 CTAs 202 us, 2450 198, 4900 190, 9800 186, 14700 184, 29400 192; C = 1: 81 / 71 / 67 / 67 / 69 / 77
        long gz = (target + (long)tiles_x * tiles_y - 1) / ((long)tiles_x * tiles_y);
        gz = gz < 1 ? 1 : (gz > B ? B : (gz > 65535 ? 65535 : gz));
        const dim3 grid((unsigned)tiles_x, (unsigned)tiles_y, (unsigned)gz);
        if (tiles_y > 65535) return fail(CF_ERR_UNSUPPORTED, "warp2d: image height %d too large", H);
        // L2 prefetch rows per tile and channel (0 = off); needs 16-byte aligned rows.  Measured at cfg3 (profiles/README.md, round 2
        // "warp"): C = 5 175.5 us without, 173.5-174.5 with 16-24 rows (167 on another box); C = 1 69.0 without, 75 with: one channel
        // issues on ~85 % of the cycles already and the prefetch instructions cost more than the latency they hide -> C >= 2 only.
        static const int pf_knob = [] { const char* e = getenv("CF_WARP_PF"); return e ? atoi(e) : 16; }();
        const int pf = (C >= 2 && W % 4 == 0 && aligned16(src)) ? pf_knob : 0;
        // 8 resident CTAs per SM (32 registers, no spills) against 6 (40): C = 5 187.6 -> 175.5 us, C = 4 154.0 -> 144.8, C = 1 69.0 = 69.0
        static const int minb8 = [] { const char* e = getenv("CF_WARP_MINB8"); return e ? atoi(e) : 1; }();
        if (H - 1 <= 4095 && W - 1 <= 4095 && minb8) warp2d_tile_kernel<true, 1, 8><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry_, rx_, pf);
        else if (H - 1 <= 4095 && W - 1 <= 4095) warp2d_tile_kernel<true><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry_, rx_, pf);
        else warp2d_tile_kernel<false><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry_, rx_, pf);
        CF_CHECK_LAUNCH("warp2d_tile_kernel");
        return CF_OK;
    }
    const bool vec = (W % 4 == 0) && aligned16(flow) && aligned16(src) && aligned16(out);
    const int nv = vec ? (want == 4 ? 4 : 2) : 1;
    const unsigned wv = (unsigned)(W / nv);
    int wv_shift = -1;
    for (int k = 0; k < 31; ++k)
        if ((1u << k) == wv) wv_shift = k;
    const dim3 grid((unsigned)cdiv((long)H * wv, 256), (unsigned)(B < 65535 ? B : 65535));
    const float ry = 1.0f / (float)(H - 1), rx = 1.0f / (float)(W - 1);   // IEEE-rounded reciprocals for div_small
    if (mode == CF_MODE_BILINEAR) {
        if (nv == 4) warp2d_kernel<4, CF_MODE_BILINEAR><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry, rx, wv_shift);
        else if (nv == 2) warp2d_kernel<2, CF_MODE_BILINEAR><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry, rx, wv_shift);
        else warp2d_kernel<1, CF_MODE_BILINEAR><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry, rx, wv_shift);
    } else {
        if (nv == 4) warp2d_kernel<4, CF_MODE_NEAREST><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry, rx, wv_shift);
        else if (nv == 2) warp2d_kernel<2, CF_MODE_NEAREST><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry, rx, wv_shift);
        else warp2d_kernel<1, CF_MODE_NEAREST><<<grid, 256, 0, s>>>(flow, src, out, B, C, H, W, ry, rx, wv_shift);
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

int launch_warp2d_onehot_argmax_u8(const float* flow, const int64_t* label, int ncls, int B, int H, int W, uint8_t* out,
                                   int64_t* counts, int counts_total, cudaStream_t s) {
    const float ry = 1.0f / (float)(H - 1), rx = 1.0f / (float)(W - 1);
    const int tx = cdiv(W, kTileW), ty = cdiv(H, kTileH);
    if (!((long)H * W < (1L << 30) && ty <= 65535)) return fail(CF_ERR_UNSUPPORTED, "warp2d_onehot_argmax_u8: image of %dx%d pixels too large", H, W);
    if (ncls > 256) return fail(CF_ERR_UNSUPPORTED, "warp2d_onehot_argmax_u8: %d classes do not fit uint8", ncls);
    long gz = (4736 + (long)tx * ty - 1) / ((long)tx * ty);
    gz = gz < 1 ? 1 : (gz > B ? B : (gz > 65535 ? 65535 : gz));
    const dim3 grid((unsigned)tx, (unsigned)ty, (unsigned)gz);
    const bool fast = H - 1 <= 4095 && W - 1 <= 4095;
    unsigned long long* cn = reinterpret_cast<unsigned long long*>(counts);
    if (counts) {
        CF_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * (size_t)(counts_total ? 1 : B) * ncls, s));
        if (fast) warp2d_onehot_argmax_tile_kernel<true, true, uint8_t><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, cn, counts_total);
        else warp2d_onehot_argmax_tile_kernel<false, true, uint8_t><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, cn, counts_total);
    } else {
        if (fast) warp2d_onehot_argmax_tile_kernel<true, false, uint8_t><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, nullptr);
        else warp2d_onehot_argmax_tile_kernel<false, false, uint8_t><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, nullptr);
    }
    CF_CHECK_LAUNCH("warp2d_onehot_argmax_tile_kernel<uint8>");
    return CF_OK;
}

int launch_warp2d_onehot_argmax(const float* flow, const int64_t* label, int ncls, int B, int H, int W,
                                int64_t* out, int64_t* counts, cudaStream_t s) {
    const float ry = 1.0f / (float)(H - 1), rx = 1.0f / (float)(W - 1);
    if (counts) {
        const int tx = cdiv(W, kTileW), ty = cdiv(H, kTileH);
        if (!((long)H * W < (1L << 30) && ty <= 65535)) return fail(CF_ERR_UNSUPPORTED, "warp2d_onehot_argmax: counts need an image below 2^30 pixels");
        CF_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * (size_t)B * ncls, s));
        long gz = (4736 + (long)tx * ty - 1) / ((long)tx * ty);
        gz = gz < 1 ? 1 : (gz > B ? B : (gz > 65535 ? 65535 : gz));
        const dim3 grid((unsigned)tx, (unsigned)ty, (unsigned)gz);
        unsigned long long* cn = reinterpret_cast<unsigned long long*>(counts);
        if (H - 1 <= 4095 && W - 1 <= 4095) warp2d_onehot_argmax_tile_kernel<true, true><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, cn);
        else warp2d_onehot_argmax_tile_kernel<false, true><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, cn);
        CF_CHECK_LAUNCH("warp2d_onehot_argmax_tile_kernel<count>");
        return CF_OK;
    }
    static const bool legacy = [] { const char* e = getenv("CF_ARGMAX_LEGACY"); return e && atoi(e) != 0; }();   // tuning knob
    const int tiles_x = cdiv(W, kTileW), tiles_y = cdiv(H, kTileH);
    if (!legacy && (long)H * W < (1L << 30) && tiles_y <= 65535) {
        long gz = (4736 + (long)tiles_x * tiles_y - 1) / ((long)tiles_x * tiles_y);
        gz = gz < 1 ? 1 : (gz > B ? B : (gz > 65535 ? 65535 : gz));
        const dim3 grid((unsigned)tiles_x, (unsigned)tiles_y, (unsigned)gz);
        if (H - 1 <= 4095 && W - 1 <= 4095) warp2d_onehot_argmax_tile_kernel<true, false><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, nullptr);
        else warp2d_onehot_argmax_tile_kernel<false, false><<<grid, 256, 0, s>>>(flow, label, out, ncls, B, H, W, ry, rx, nullptr);
        CF_CHECK_LAUNCH("warp2d_onehot_argmax_tile_kernel");
        return CF_OK;
    }
    const long total = (long)B * H * W;
    warp2d_onehot_argmax_kernel<<<cdiv(total, 256), 256, 0, s>>>(flow, label, out, ncls, H, W, ry, rx, total);
    CF_CHECK_LAUNCH("warp2d_onehot_argmax_kernel");
    return CF_OK;
}

}  // namespace cf
