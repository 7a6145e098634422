// SURVEY.md 8(f) row 4, first half: the local correlation layer `CorrVolume(radius, stride)(cur, prev)` the reference puts
// on its skip connections (/root/reference/nnunet/network_architecture/SegFlowGaussian.py:41 import, :259 construction with
// radius[idx] / stride[idx] (config `stride: [4, 2, 1, 1]`), :1376 call `corr = layer(skip_co_current[s], previous_skip_co[s])`,
// result fed to a conv block with (2r+1)^2 input channels and concatenated at the skip connection's own resolution).
//
// PARITY UNPINNED: the defining file nnunet/lib/raft.py was never committed and no third-party twin is called by the
// reference, so only the call-site contract is known: [B,C,H,W] x [B,C,H,W] -> [B,(2r+1)^2,H,W].  This kernel implements the
// standard local cost volume of that shape,
//     out[b, (dy+r)*(2r+1) + (dx+r), y, x] = scale * sum_c cur[b,c,y,x] * prev[b,c, y + dy*stride, x + dx*stride]
// with zero padding; `scale` (1/sqrt(C) like RAFT's CorrBlock by default, 1/C or 1 on request) and the channel order
// (dy-major by default) are arguments so that a user holding the missing file can match it.
//
// One thread = one output pixel x one displacement row dy: 2r+1 accumulators over C; lanes run along x, so the cur load and
// every prev load of a warp are contiguous 128-byte requests; the 2r+1 displacement rows of a pixel sit in the same CTA and
// share their prev lines through L1.  Compute-bound on the fp32 pipes (2*C*(2r+1)^2 FLOP per pixel); not tuned further.
#include "common.cuh"

namespace cf {
namespace {

template <int R>
__global__ void __launch_bounds__(32 * (2 * R + 1)) local_corr_kernel(const float* __restrict__ cur, const float* __restrict__ prev,
                                                                      float* __restrict__ out, int C, int H, int W, int stride,
                                                                      float scale, int x_major) {
    constexpr int SIDE = 2 * R + 1;
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y, b = blockIdx.z, i = threadIdx.y;   // i: displacement row
    if (x >= W) return;
    const size_t HW = (size_t)H * W;
    const int yy = y + (i - R) * stride;
    float acc[SIDE];
#pragma unroll
    for (int j = 0; j < SIDE; ++j) acc[j] = 0.0f;
    if ((unsigned)yy < (unsigned)H) {
        const float* cp = cur + (size_t)b * C * HW + (size_t)y * W + x;
        const float* pp = prev + (size_t)b * C * HW + (size_t)yy * W;
        bool in[SIDE];
        int xo[SIDE];
#pragma unroll
        for (int j = 0; j < SIDE; ++j) {
            const int xx = x + (j - R) * stride;
            in[j] = (unsigned)xx < (unsigned)W;
            xo[j] = in[j] ? xx : 0;
        }
        for (int c = 0; c < C; ++c, cp += HW, pp += HW) {
            const float a = __ldg(cp);
#pragma unroll
            for (int j = 0; j < SIDE; ++j) acc[j] = fmaf(a, in[j] ? __ldg(pp + xo[j]) : 0.0f, acc[j]);
        }
    }
    float* o = out + (size_t)b * SIDE * SIDE * HW + (size_t)y * W + x;
#pragma unroll
    for (int j = 0; j < SIDE; ++j) o[(size_t)(x_major ? j * SIDE + i : i * SIDE + j) * HW] = acc[j] * scale;
}

// d/d(cur)[b,c,y,x]  = scale * sum_d go[b,d,y,x] * prev[b,c,y+dy*s,x+dx*s]
// d/d(prev)[b,c,y,x] = scale * sum_d go[b,d,y-dy*s,x-dx*s] * cur[b,c,y-dy*s,x-dx*s]
// One thread per (b, c, y, x); lanes along x.  Both gradients are written (no atomics).
template <int R>
__global__ void __launch_bounds__(128) local_corr_bwd_kernel(const float* __restrict__ cur, const float* __restrict__ prev,
                                                             const float* __restrict__ go, float* __restrict__ gcur,
                                                             float* __restrict__ gprev, int C, int H, int W, int stride,
                                                             float scale, int x_major) {
    constexpr int SIDE = 2 * R + 1;
    const int x = blockIdx.x * 128 + threadIdx.x, y = blockIdx.y;
    const int c = blockIdx.z % C, b = blockIdx.z / C;
    if (x >= W) return;
    const size_t HW = (size_t)H * W;
    const float* gob = go + (size_t)b * SIDE * SIDE * HW;
    const float* curc = cur + ((size_t)b * C + c) * HW;
    const float* prevc = prev + ((size_t)b * C + c) * HW;
    float a_cur = 0.0f, a_prev = 0.0f;
    for (int i = 0; i < SIDE; ++i) {
        const int dy = (i - R) * stride;
#pragma unroll
        for (int j = 0; j < SIDE; ++j) {
            const int dx = (j - R) * stride;
            const size_t ch = (size_t)(x_major ? j * SIDE + i : i * SIDE + j) * HW;
            const int yp = y + dy, xp = x + dx;       // prev position this output pixel looked at
            if (gcur && (unsigned)yp < (unsigned)H && (unsigned)xp < (unsigned)W)
                a_cur = fmaf(__ldg(gob + ch + (size_t)y * W + x), __ldg(prevc + (size_t)yp * W + xp), a_cur);
            const int ym = y - dy, xm = x - dx;       // output pixel that looked at this prev position
            if (gprev && (unsigned)ym < (unsigned)H && (unsigned)xm < (unsigned)W)
                a_prev = fmaf(__ldg(gob + ch + (size_t)ym * W + xm), __ldg(curc + (size_t)ym * W + xm), a_prev);
        }
    }
    const size_t o = ((size_t)b * C + c) * HW + (size_t)y * W + x;
    if (gcur) gcur[o] = a_cur * scale;
    if (gprev) gprev[o] = a_prev * scale;
}

}  // namespace

int launch_local_corr(const float* cur, const float* prev, int B, int C, int H, int W, int r, int stride, float scale,
                      int x_major, float* out, cudaStream_t s) {
    if (H > 65535 || B > 65535) return fail(CF_ERR_UNSUPPORTED, "local_corr: H or B above 65535");
    const dim3 grid((unsigned)cdiv(W, 32), (unsigned)H, (unsigned)B);
    switch (r) {
        case 1: local_corr_kernel<1><<<grid, dim3(32, 3), 0, s>>>(cur, prev, out, C, H, W, stride, scale, x_major); break;
        case 2: local_corr_kernel<2><<<grid, dim3(32, 5), 0, s>>>(cur, prev, out, C, H, W, stride, scale, x_major); break;
        case 3: local_corr_kernel<3><<<grid, dim3(32, 7), 0, s>>>(cur, prev, out, C, H, W, stride, scale, x_major); break;
        case 4: local_corr_kernel<4><<<grid, dim3(32, 9), 0, s>>>(cur, prev, out, C, H, W, stride, scale, x_major); break;
        default: return fail(CF_ERR_UNSUPPORTED, "local_corr radius %d not in [1,%d]", r, CF_MAX_RADIUS);
    }
    CF_CHECK_LAUNCH("local_corr_kernel");
    return CF_OK;
}

int launch_local_corr_bwd(const float* cur, const float* prev, const float* go, int B, int C, int H, int W, int r, int stride,
                          float scale, int x_major, float* gcur, float* gprev, cudaStream_t s) {
    if (H > 65535 || (long)B * C > 65535) return fail(CF_ERR_UNSUPPORTED, "local_corr_bwd: H or B*C above 65535");
    const dim3 grid((unsigned)cdiv(W, 128), (unsigned)H, (unsigned)(B * C));
    switch (r) {
        case 1: local_corr_bwd_kernel<1><<<grid, 128, 0, s>>>(cur, prev, go, gcur, gprev, C, H, W, stride, scale, x_major); break;
        case 2: local_corr_bwd_kernel<2><<<grid, 128, 0, s>>>(cur, prev, go, gcur, gprev, C, H, W, stride, scale, x_major); break;
        case 3: local_corr_bwd_kernel<3><<<grid, 128, 0, s>>>(cur, prev, go, gcur, gprev, C, H, W, stride, scale, x_major); break;
        case 4: local_corr_bwd_kernel<4><<<grid, 128, 0, s>>>(cur, prev, go, gcur, gprev, C, H, W, stride, scale, x_major); break;
        default: return fail(CF_ERR_UNSUPPORTED, "local_corr radius %d not in [1,%d]", r, CF_MAX_RADIUS);
    }
    CF_CHECK_LAUNCH("local_corr_bwd_kernel");
    return CF_OK;
}

}  // namespace cf
