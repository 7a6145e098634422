// K6: RAFT convex 8x upsampling of a low-resolution field with a learned mask.
//
//   out[n, c, 8y + i, 8x + j] = sum_k softmax_k(mask[n, k*64 + i*8 + j, y, x]) * (8 * field[n, c, y + ky - 1, x + kx - 1])
//   k = ky*3 + kx over the 3x3 neighbourhood, zeros outside the field (F.unfold padding=1)
//
// Replaces /root/reference/nnunet/network_architecture/SegFlowGaussian.py:846-857 (`upsample_flow`, C = 2) and
// :860-871 (`upsample_seg`, C = 4; it scales by 8 as well), the step right after every lookup / update iteration:
// view + softmax over a strided dim + `8 * flow` + unfold + broadcast multiply + sum + permute + reshape copy, i.e. the
// [B,576,H,W] mask is read, a softmaxed copy written and re-read, a [B,C,9,8,8,H,W] product written and re-read.
//
// Here: one launch, HBM-bound on the compulsory traffic 4*HW*(576 + 64*C + C) bytes per item.
//   * lane <-> low-resolution column x, warp <-> sub-row i (8 warps per CTA), CTA <-> (32 columns, row y, item n)
//   * for every (k, j) the warp reads mask channel k*64 + i*8 + j at 32 consecutive x: one 128 B line per request,
//     72 independent requests per thread in flight
//   * a lane accumulates its 8 sub-columns j in registers: they are 8 CONSECUTIVE floats of output row 8y + i, so a warp
//     stores 1 KB contiguous per (channel, row) with two 128-bit streaming stores per lane -- no shared memory, no transpose
//   * softmax in ATen's order (x - max, exp, sum, normalise) with ex2.approx and one reciprocal per sub-pixel (the
//     kernel was issue-bound with expf + IEEE division: 34 % -> see profiles/); products rounded before the sequential
//     sum over k; parity bar 1e-5 relative, measured ~1e-6
#include "common.cuh"

namespace cf {
namespace {

constexpr int kUpF = 8;          // upsampling factor (the reference hard-codes 8)
constexpr int kUpMaxC = 4;       // channels per launch (2 = flow, 4 = segmentation logits)

template <int NC>
__global__ void __launch_bounds__(256, NC >= 3 ? 3 : 4) upsample_convex_kernel(const float* __restrict__ field, const float* __restrict__ mask,
                                                              float* __restrict__ out, int C, int c0, int H, int W, float scale) {
    const int lane = threadIdx.x & 31, i = threadIdx.x >> 5;
    const int x = blockIdx.x * 32 + lane, y = blockIdx.y, n = blockIdx.z;
    if (x >= W) return;
    const size_t HW = (size_t)H * W;

    // the 3x3 neighbourhood of the scaled field, zero outside (unfold padding)
    float nb[NC][9];
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const float* f = field + ((size_t)n * C + c0 + c) * HW;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                const int yy = y + ky - 1, xx = x + kx - 1;
                const bool in = (unsigned)yy < (unsigned)H && (unsigned)xx < (unsigned)W;
                nb[c][ky * 3 + kx] = in ? __fmul_rn(scale, __ldg(f + (size_t)yy * W + xx)) : 0.0f;
            }
    }

    const float* m0 = mask + ((size_t)n * (9 * kUpF * kUpF) + (size_t)i * kUpF) * HW + (size_t)y * W + x;
    float acc[NC][kUpF];
#pragma unroll
    for (int j = 0; j < kUpF; ++j) {
        float m[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) m[k] = __ldcs(m0 + ((size_t)k * (kUpF * kUpF) + j) * HW);
        float mx = m[0];
#pragma unroll
        for (int k = 1; k < 9; ++k) mx = fmaxf(mx, m[k]);
        float sum = 0.0f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            m[k] = __expf(__fsub_rn(m[k], mx));      // arguments <= 0: ex2.approx is within ~2 ulp where the weight matters
            sum = __fadd_rn(sum, m[k]);
        }
        const float inv = __frcp_rn(sum);            // sum in [1, 9]; one reciprocal instead of nine divisions (<= 1 ulp apart)
#pragma unroll
        for (int k = 0; k < 9; ++k) m[k] = __fmul_rn(m[k], inv);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            float a = __fmul_rn(m[0], nb[c][0]);
#pragma unroll
            for (int k = 1; k < 9; ++k) a = __fadd_rn(a, __fmul_rn(m[k], nb[c][k]));
            acc[c][j] = a;
        }
    }

    const size_t OW = (size_t)W * kUpF;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        float4* o = reinterpret_cast<float4*>(out + (((size_t)n * C + c0 + c) * H * kUpF + (size_t)y * kUpF + i) * OW + (size_t)x * kUpF);
        __stcs(o, make_float4(acc[c][0], acc[c][1], acc[c][2], acc[c][3]));
        __stcs(o + 1, make_float4(acc[c][4], acc[c][5], acc[c][6], acc[c][7]));
    }
}

// Backward of the convex upsampling (autograd of SegFlowGaussian.py:846-871 as the trainer differentiates it):
//   ds_k = sum_c g[c] * nb[c][k],  dmask_k = s_k * (ds_k - sum_k' s_k' ds_k'),  dfield[c, neighbour k] += scale * s_k * g[c]
// Same mapping as the forward kernel.  grad_mask is written coalesced (128 B per warp and channel); the field gradient is
// accumulated over the 8 sub-columns in registers, over the 8 sub-rows (warps) in shared memory, and then scattered to the
// 3x3 neighbourhood with one atomicAdd per (channel, neighbour) and low-resolution pixel.
template <int NC>
__global__ void __launch_bounds__(256) upsample_convex_bwd_kernel(const float* __restrict__ field, const float* __restrict__ mask,
                                                                  const float* __restrict__ gout, float* __restrict__ gfield,
                                                                  float* __restrict__ gmask, int C, int c0, int H, int W,
                                                                  float scale) {
    __shared__ float red[kUpF][NC * 9][32];
    const int lane = threadIdx.x & 31, i = threadIdx.x >> 5;
    const int x = blockIdx.x * 32 + lane, y = blockIdx.y, n = blockIdx.z;
    const bool active = x < W;
    const size_t HW = (size_t)H * W;
    float dnb[NC][9];
#pragma unroll
    for (int c = 0; c < NC; ++c)
#pragma unroll
        for (int k = 0; k < 9; ++k) dnb[c][k] = 0.0f;

    if (active) {
        float nb[NC][9];
        float g[NC][kUpF];
        const size_t OW = (size_t)W * kUpF;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const float* f = field + ((size_t)n * C + c0 + c) * HW;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const int yy = y + ky - 1, xx = x + kx - 1;
                    const bool in = (unsigned)yy < (unsigned)H && (unsigned)xx < (unsigned)W;
                    nb[c][ky * 3 + kx] = in ? __fmul_rn(scale, __ldg(f + (size_t)yy * W + xx)) : 0.0f;
                }
            const float4* gp = reinterpret_cast<const float4*>(gout + (((size_t)n * C + c0 + c) * H * kUpF + (size_t)y * kUpF + i) * OW +
                                                                (size_t)x * kUpF);
            const float4 a = __ldcs(gp), b = __ldcs(gp + 1);
            g[c][0] = a.x; g[c][1] = a.y; g[c][2] = a.z; g[c][3] = a.w;
            g[c][4] = b.x; g[c][5] = b.y; g[c][6] = b.z; g[c][7] = b.w;
        }
        const size_t moff = ((size_t)n * (9 * kUpF * kUpF) + (size_t)i * kUpF) * HW + (size_t)y * W + x;
#pragma unroll
        for (int j = 0; j < kUpF; ++j) {
            float m[9];
#pragma unroll
            for (int k = 0; k < 9; ++k) m[k] = __ldcs(mask + moff + ((size_t)k * (kUpF * kUpF) + j) * HW);
            float mx = m[0];
#pragma unroll
            for (int k = 1; k < 9; ++k) mx = fmaxf(mx, m[k]);
            float sum = 0.0f;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                m[k] = __expf(m[k] - mx);
                sum += m[k];
            }
            const float inv = __frcp_rn(sum);
            float ds[9], dot = 0.0f;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                m[k] *= inv;
                float d = 0.0f;
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    d = fmaf(g[c][j], nb[c][k], d);
                    dnb[c][k] = fmaf(m[k], g[c][j], dnb[c][k]);
                }
                ds[k] = d;
                dot = fmaf(m[k], d, dot);
            }
            if (gmask != nullptr) {
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    float* dst = gmask + moff + ((size_t)k * (kUpF * kUpF) + j) * HW;
                    float v = m[k] * (ds[k] - dot);
                    if (c0 > 0) v += *dst;                  // later channel chunks add to the first one's result
                    *dst = v;
                }
            }
        }
    }
    if (gfield == nullptr) return;
    // sum over the 8 sub-rows, then scatter
#pragma unroll
    for (int c = 0; c < NC; ++c)
#pragma unroll
        for (int k = 0; k < 9; ++k) red[i][c * 9 + k][lane] = dnb[c][k];
    __syncthreads();
    for (int e = i; e < NC * 9; e += kUpF) {
        float v = 0.0f;
#pragma unroll
        for (int r = 0; r < kUpF; ++r) v += red[r][e][lane];
        const int c = e / 9, k = e - c * 9;
        const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
        if (active && (unsigned)yy < (unsigned)H && (unsigned)xx < (unsigned)W)
            atomicAdd(gfield + ((size_t)n * C + c0 + c) * HW + (size_t)yy * W + xx, scale * v);
    }
}

}  // namespace

int launch_upsample_convex(const float* field, const float* mask, int B, int C, int H, int W, float scale, float* out,
                           cudaStream_t s) {
    if (H > 65535 || B > 65535) return fail(CF_ERR_UNSUPPORTED, "upsample_convex: H=%d or B=%d above 65535", H, B);
    const dim3 grid((unsigned)cdiv(W, 32), (unsigned)H, (unsigned)B);
    for (int c0 = 0; c0 < C; c0 += kUpMaxC) {
        const int nc = C - c0 < kUpMaxC ? C - c0 : kUpMaxC;
        switch (nc) {
            case 1: upsample_convex_kernel<1><<<grid, 256, 0, s>>>(field, mask, out, C, c0, H, W, scale); break;
            case 2: upsample_convex_kernel<2><<<grid, 256, 0, s>>>(field, mask, out, C, c0, H, W, scale); break;
            case 3: upsample_convex_kernel<3><<<grid, 256, 0, s>>>(field, mask, out, C, c0, H, W, scale); break;
            default: upsample_convex_kernel<4><<<grid, 256, 0, s>>>(field, mask, out, C, c0, H, W, scale); break;
        }
        CF_CHECK_LAUNCH("upsample_convex_kernel");
    }
    return CF_OK;
}

int launch_upsample_convex_bwd(const float* field, const float* mask, const float* gout, int B, int C, int H, int W, float scale,
                               float* gfield, float* gmask, cudaStream_t s) {
    if (H > 65535 || B > 65535) return fail(CF_ERR_UNSUPPORTED, "upsample_convex_bwd: H=%d or B=%d above 65535", H, B);
    const dim3 grid((unsigned)cdiv(W, 32), (unsigned)H, (unsigned)B);
    for (int c0 = 0; c0 < C; c0 += kUpMaxC) {
        const int nc = C - c0 < kUpMaxC ? C - c0 : kUpMaxC;
        switch (nc) {
            case 1: upsample_convex_bwd_kernel<1><<<grid, 256, 0, s>>>(field, mask, gout, gfield, gmask, C, c0, H, W, scale); break;
            case 2: upsample_convex_bwd_kernel<2><<<grid, 256, 0, s>>>(field, mask, gout, gfield, gmask, C, c0, H, W, scale); break;
            case 3: upsample_convex_bwd_kernel<3><<<grid, 256, 0, s>>>(field, mask, gout, gfield, gmask, C, c0, H, W, scale); break;
            default: upsample_convex_bwd_kernel<4><<<grid, 256, 0, s>>>(field, mask, gout, gfield, gmask, C, c0, H, W, scale); break;
        }
        CF_CHECK_LAUNCH("upsample_convex_bwd_kernel");
    }
    return CF_OK;
}

}  // namespace cf
