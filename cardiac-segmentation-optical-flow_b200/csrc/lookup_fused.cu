// K4 (optional path of north_star; not in the reference): CorrBlock lookup WITHOUT a materialised volume.
//
// avg_pool over q of <f1[p], f2[q]> == <f1[p], avg_pool(f2)[q]> (linearity), so level l of the correlation
// pyramid restricted to the (2r+4)^2 window a query can touch is a set of dot products between the query's
// f1 column and the l-times pooled fmap2.  Memory: B*C*HW*(1 + 1/4 + ...) floats instead of B*HW^2*(...) --
// cfg4 (64x64 fmaps, 64 pairs): 0.36 GB instead of 5.70 GB.  Cost: the window correlations are recomputed by
// every lookup (sum_l 2*C*HW*min(100, h_l*w_l) FLOP per pair per call).
//
//   cf_fmap_pyramid_build   level 0 = fmap2, level l = avg_pool2d(level l-1, 2, 2)   [B, C, h_l, w_l]
//   cf_corr_lookup_fused    CTA = 32 consecutive queries x one level, 4 warps:
//       1. the queries' f1 columns are staged in shared memory ([C][32], coalesced)
//       2. thread (lane = query, warp = window row) accumulates one window row of dot products over C with
//          the f1 value shared in a register; pooled fmap2 is read through L1/L2 (adjacent queries have
//          adjacent windows -> the 32 lanes of a request hit the same few lines)
//       3. windows (scaled by 1/sqrt(C), zero outside the image) are interpolated from shared memory with
//          the same exact tap arithmetic as the materialised lookup (sampling.cuh: axis_tap)
// Results differ from the materialised fp32 path only by fp32 summation order (pool-then-dot vs dot-then-pool).
#include "sampling.cuh"
#include "tc_common.cuh"

namespace cf {

bool fmap_level_dims(int B, int C, int H, int W, int L, LevelDims* d) {
    size_t off = 0;
    int h = H, w = W;
    for (int l = 0; l < L; ++l) {
        if (h < 1 || w < 1) return false;
        d->h[l] = h;
        d->w[l] = w;
        d->off[l] = off;
        off += round_up((size_t)B * C * h * w * sizeof(float), 256);
        h /= 2;
        w /= 2;
    }
    d->off[L] = off;
    return true;
}

int launch_fmap_pyramid(const float* f2, int B, int C, int H, int W, int L, float* f2pyr, cudaStream_t s) {
    LevelDims d;
    if (!fmap_level_dims(B, C, H, W, L, &d)) return fail(CF_ERR_UNSUPPORTED, "empty fmap pyramid level");
    CF_CHECK_CUDA(cudaMemcpyAsync(f2pyr, f2, (size_t)B * C * H * W * sizeof(float), cudaMemcpyDeviceToDevice, s));
    char* base = reinterpret_cast<char*>(f2pyr);
    for (int l = 0; l + 1 < L; ++l) {
        const int rc = launch_pool2x2(reinterpret_cast<const float*>(base + d.off[l]),
                                      reinterpret_cast<float*>(base + d.off[l + 1]), (long)B * C, d.h[l], d.w[l], s);
        if (rc) return rc;
    }
    return CF_OK;
}

namespace {

constexpr int kQF = 32;   // queries per CTA
constexpr int kNWF = 4;   // warps per CTA

struct FusedParams {
    const float* f1;
    const float* f2l[CF_MAX_LEVELS];
    int lh[CF_MAX_LEVELS];
    int lw[CF_MAX_LEVELS];
    const float* coords;
    float* out;
    long N;
    int HW;
    int C;
    int L;
    float div;   // sqrt(C)
};

template <int R>
__global__ void __launch_bounds__(32 * kNWF) lookup_fused_kernel(const __grid_constant__ FusedParams P) {
    constexpr int SIDE = 2 * R + 1, PR = 2 * R + 4, CELLS = PR * PR, K = SIDE * SIDE;
    constexpr int PITCH = CELLS + 1;                 // odd: lane-private windows are conflict-free
    constexpr int JPW = (SIDE + kNWF - 1) / kNWF;    // output rows per warp
    extern __shared__ __align__(16) float smem[];
    float* const f1s = smem;                         // [C][32]
    float* const win = smem + (size_t)P.C * kQF;     // [32][PITCH]

    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned task = blockIdx.x / (unsigned)P.L;
    const int l = (int)(blockIdx.x - task * (unsigned)P.L);
    const long n = (long)task * kQF + lane;
    const bool valid = n < P.N;
    const int hl = P.lh[l], wl = P.lw[l];
    const int C = P.C;

    long b = 0;
    int p = 0;
    float cx = 0.0f, cy = 0.0f;
    if (valid) {
        b = n / P.HW;
        p = (int)(n - b * P.HW);
        const float inv = 1.0f / (float)(1 << l);
        cx = __fmul_rn(__ldg(P.coords + (b * 2 + 0) * P.HW + p), inv);
        cy = __fmul_rn(__ldg(P.coords + (b * 2 + 1) * P.HW + p), inv);
    }
    const bool dead = !valid || !(fabsf(cx) < kDeadCoord) || !(fabsf(cy) < kDeadCoord);
    const int ax = dead ? 0 : __float2int_rd(cx) - R - 1;
    const int ay = dead ? 0 : __float2int_rd(cy) - R - 1;

    // ---- 1. f1 columns of the 32 queries ----
    {
        const float* f1p = P.f1 + (size_t)b * C * P.HW + p;
        for (int c = w; c < C; c += kNWF) f1s[c * kQF + lane] = valid ? __ldg(f1p + (size_t)c * P.HW) : 0.0f;
    }
    __syncthreads();

    // ---- 2. window correlations: warp w owns window rows w, w + 4, ... ----
    {
        const size_t img = (size_t)hl * wl;
        const float* f2b = P.f2l[l] + (size_t)b * C * img;
        int xc[PR];
#pragma unroll
        for (int k = 0; k < PR; ++k) xc[k] = min(max(ax + k, 0), wl - 1);
        float* my = win + lane * PITCH;
        for (int row = w; row < PR; row += kNWF) {
            const int y = ay + row;
            const bool row_in = !dead && (unsigned)y < (unsigned)hl;
            float acc[PR];
#pragma unroll
            for (int k = 0; k < PR; ++k) acc[k] = 0.0f;
            if (row_in) {
                const float* rp = f2b + (size_t)y * wl;
#pragma unroll 2
                for (int c = 0; c < C; ++c, rp += img) {
                    const float a = f1s[c * kQF + lane];
#pragma unroll
                    for (int k = 0; k < PR; ++k) acc[k] = fmaf(a, __ldg(rp + xc[k]), acc[k]);
                }
            }
#pragma unroll
            for (int k = 0; k < PR; ++k) {
                const bool in = row_in && (unsigned)(ax + k) < (unsigned)wl;
                my[row * PR + k] = in ? __fdiv_rn(acc[k], P.div) : 0.0f;
            }
        }
    }
    __syncthreads();
    if (!valid) return;

    // ---- 3. bilinear interpolation of the window (same tap arithmetic as lookup.cu) ----
    const size_t HW = (size_t)P.HW;
    float* o = P.out + ((size_t)b * P.L * K + (size_t)l * K) * HW + p;
    const float* sq = win + lane * PITCH;
    const float s1x = (float)(wl - 1), s1y = (float)(hl - 1);
        const float rcx = __frcp_rn(s1x), rcy = __frcp_rn(s1y);   // IEEE-rounded reciprocals for div_small
    int yo[JPW];
    float yl[JPW], yh[JPW];
#pragma unroll
    for (int k = 0; k < JPW; ++k) {
        const int j = w * JPW + k;
        int j0 = 0;
        yl[k] = yh[k] = 0.0f;
        if (j < SIDE) axis_tap(cy, j - R, s1y, rcy, j0, yl[k], yh[k]);
        yo[k] = min(max(j0 - ay, 0), PR - 2) * PR;
    }
#pragma unroll
    for (int i = 0; i < SIDE; ++i) {
        int i0;
        float xl, xh;
        axis_tap(cx, i - R, s1x, rcx, i0, xl, xh);
        const int xo = min(max(i0 - ax, 0), PR - 2);
#pragma unroll
        for (int k = 0; k < JPW; ++k) {
            const int j = w * JPW + k;
            if (j < SIDE) {
                float acc = 0.0f;
                if (!dead) {
                    const float* t = sq + yo[k] + xo;
                    acc = __fmul_rn(t[0], __fmul_rn(xl, yl[k]));
                    acc = fmaf(t[1], __fmul_rn(xh, yl[k]), acc);
                    acc = fmaf(t[PR], __fmul_rn(xl, yh[k]), acc);
                    acc = fmaf(t[PR + 1], __fmul_rn(xh, yh[k]), acc);
                }
                __stcs(o + (size_t)(i * SIDE + j) * HW, acc);
            }
        }
    }
}

template <int R>
int launch_fused_r(const FusedParams& P, cudaStream_t s) {
    constexpr int PR = 2 * R + 4;
    const long grid = (P.N + kQF - 1) / kQF * P.L;
    if (grid > 0x7fffffffL) return fail(CF_ERR_UNSUPPORTED, "fused lookup grid too large (%ld)", grid);
    const size_t smem = ((size_t)P.C * kQF + (size_t)kQF * (PR * PR + 1)) * sizeof(float);
    if (smem > 200 * 1024) return fail(CF_ERR_UNSUPPORTED, "fused lookup: C = %d needs %zu B of shared memory", P.C, smem);
    {   // a per-device attribute: once per device and kernel (tc_common.cuh)
        const int rc = smem_optin(reinterpret_cast<const void*>(&lookup_fused_kernel<R>), 200 * 1024, 20 + R);
        if (rc) return rc;
    }
    lookup_fused_kernel<R><<<(unsigned)grid, 32 * kNWF, smem, s>>>(P);
    CF_CHECK_LAUNCH("lookup_fused_kernel");
    return CF_OK;
}

}  // namespace

int launch_lookup_fused(const float* f1, const float* f2pyr, const float* coords, int B, int C, int H, int W, int L,
                        int r, float* out, cudaStream_t s) {
    LevelDims d;
    if (!fmap_level_dims(B, C, H, W, L, &d)) return fail(CF_ERR_UNSUPPORTED, "empty fmap pyramid level");
    FusedParams P{};
    P.f1 = f1;
    for (int l = 0; l < L; ++l) {
        P.f2l[l] = reinterpret_cast<const float*>(reinterpret_cast<const char*>(f2pyr) + d.off[l]);
        P.lh[l] = d.h[l];
        P.lw[l] = d.w[l];
    }
    P.coords = coords;
    P.out = out;
    P.N = (long)B * H * W;
    P.HW = H * W;
    P.C = C;
    P.L = L;
    P.div = sqrtf((float)C);
    switch (r) {
        case 1: return launch_fused_r<1>(P, s);
        case 2: return launch_fused_r<2>(P, s);
        case 3: return launch_fused_r<3>(P, s);
        case 4: return launch_fused_r<4>(P, s);
    }
    return fail(CF_ERR_UNSUPPORTED, "lookup radius %d not in [1,%d]", r, CF_MAX_RADIUS);
}

}  // namespace cf
