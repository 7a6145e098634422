// K3: CorrBlock.__call__ -- radius-r bilinear window lookup over all pyramid levels in one launch.
//
// Replaces torchvision raft.py:394-422 (call site /root/reference/nnunet/network_architecture/
// SegFlowGaussian.py:935, 12x per frame pair): per level a delta build, add, 2 normalise kernels, cat,
// grid_sampler_2d over B*HW tiny images, then a final cat + permute + contiguous copy (~25-30 launches).
//
// Design (HBM-bound, SURVEY.md 8(d): out 4*L*K*HW + coords + window footprint per call):
//   * one warp = one task = 32 consecutive query pixels x one pyramid level; lane <-> query
//   * each lane derives its 2*(2r+1) sampling coordinates with the reference's exact fp32 sequence
//     (normalise -> un-normalise, no FMA contraction), so tap indices / weights are bit-exact
//   * the (2r+4) x (2r+4) window each query touches is staged into shared memory with 128-bit read-only
//     loads (16 B-aligned window, only the rows / 16 B chunks the taps need), zero-filled outside the
//     image -> zero padding costs nothing in the inner loop; query pitch is odd -> conflict-free LDS
//   * the 81 outputs of a lane are written straight into the final [B, L*K, H, W] layout: for a fixed
//     channel the 32 lanes of the warp write 128 contiguous bytes (no cat / permute pass)
#include "common.cuh"

namespace cf {
namespace {

constexpr int kQT = 32;  // queries per task

template <int R>
struct Geo {
    static constexpr int SIDE = 2 * R + 1;
    static constexpr int PR = 2 * R + 4;                    // window rows (and logical columns)
    static constexpr int PC = ((PR + 3 + 3) / 4) * 4;       // 16 B-aligned window columns (shift <= 3)
    static constexpr int CH = PC / 4;                       // 16 B chunks per window row
    static constexpr int QPITCH = PR * PC + 1;              // words per query, odd
    static constexpr int IPQ = PR * CH;                     // staging items (row, chunk) per query
    static constexpr int gcd_(int a, int b) { return b == 0 ? a : gcd_(b, a % b); }
    static constexpr int LCM = 32 / gcd_(32, IPQ) * IPQ;
    static constexpr int PERIOD = LCM / 32;                 // staging passes per period
    static constexpr int QPP = LCM / IPQ;                   // queries covered per period
    static_assert(32 % QPP == 0, "period must tile the 32 queries of a task");
    static_assert(PR + CH <= 32, "row/chunk masks must fit one 32-bit word twice");
};

struct LookupParams {
    const float* level[CF_MAX_LEVELS];
    int lh[CF_MAX_LEVELS];
    int lw[CF_MAX_LEVELS];
    const float* coords;
    float* out;
    float* taps;    // taps kernel only
    int32_t* idx;   // taps kernel only
    long N;         // B*H*W
    int HW;
    int L;
};

template <int R>
struct QueryTaps {
    float wl[Geo<R>::SIDE], wh[Geo<R>::SIDE];  // weight of tap i0 and of tap i0 + 1
    int i0[Geo<R>::SIDE];
};

template <int R>
__device__ __forceinline__ void axis_taps(float c, float s1, QueryTaps<R>& t) {
#pragma unroll
    for (int i = 0; i < Geo<R>::SIDE; ++i) {
        const float u = lookup_coord(__fadd_rn(c, (float)(i - R)), s1);
        const float f = floorf(u);
        t.i0[i] = __float2int_rd(u);
        t.wh[i] = __fsub_rn(u, f);
        t.wl[i] = __fsub_rn(__fadd_rn(f, 1.0f), u);
    }
}

template <int R>
__global__ void __launch_bounds__(32) lookup_kernel(const __grid_constant__ LookupParams P) {
    using G = Geo<R>;
    extern __shared__ float smem[];  // [kQT][QPITCH]
    const int lane = threadIdx.x;
    const int l = blockIdx.x % P.L;
    const long n0 = (long)(blockIdx.x / P.L) * kQT;
    const long n = n0 + lane;
    const bool valid = n < P.N;
    const int hl = P.lh[l], wl = P.lw[l];

    // ---- per-lane query: coordinates, taps, window anchor ----
    long b = 0;
    int p = 0;
    float cx = 0.0f, cy = 0.0f;
    if (valid) {
        b = n / P.HW;
        p = (int)(n - b * P.HW);
        const float sc = (float)(1 << l);
        cx = __fdiv_rn(__ldg(P.coords + (b * 2 + 0) * P.HW + p), sc);
        cy = __fdiv_rn(__ldg(P.coords + (b * 2 + 1) * P.HW + p), sc);
    }
    // every tap of a query this far out is outside any image -> zeros (also catches NaN / inf)
    const bool dead = !valid || !(fabsf(cx) < 1.0e6f) || !(fabsf(cy) < 1.0e6f);

    QueryTaps<R> tx, ty;
    axis_taps<R>(cx, (float)(wl - 1), tx);
    axis_taps<R>(cy, (float)(hl - 1), ty);

    // window anchor: nominal first tap minus one (the round trip can move a floor by +-1)
    const int ax = dead ? 0 : __float2int_rd(cx) - R - 1;
    const int ay = dead ? 0 : __float2int_rd(cy) - R - 1;
    const int ax4 = ax & ~3;       // 16 B-aligned window start (two's complement floor)
    const int sh = ax - ax4;       // 0..3

    int offx[G::SIDE], offy[G::SIDE];
    int clo = G::PR, chi = 0, rlo = G::PR, rhi = 0;
#pragma unroll
    for (int i = 0; i < G::SIDE; ++i) {
        const int lx = min(max(tx.i0[i] - ax, 0), G::PR - 2);
        const int ly = min(max(ty.i0[i] - ay, 0), G::PR - 2);
        clo = min(clo, lx); chi = max(chi, lx + 1);
        rlo = min(rlo, ly); rhi = max(rhi, ly + 1);
        offx[i] = lx + sh;
        offy[i] = ly * G::PC;
    }
    // staging masks: bit r (rows) / bit PR+c (chunks); low half = needed, high half = needed and inside image
    unsigned need = 0, inimg = 0;
    if (!dead) {
#pragma unroll
        for (int r = 0; r < G::PR; ++r) {
            const bool nd = (r >= rlo) && (r <= rhi);
            need |= (unsigned)nd << r;
            inimg |= (unsigned)(nd && (unsigned)(ay + r) < (unsigned)hl) << r;
        }
#pragma unroll
        for (int c = 0; c < G::CH; ++c) {
            const bool nd = (4 * c + 3 >= clo + sh) && (4 * c <= chi + sh);
            need |= (unsigned)nd << (G::PR + c);
            inimg |= (unsigned)(nd && (unsigned)(ax4 + 4 * c) < (unsigned)wl) << (G::PR + c);
        }
    }
    // element offset of this query's window origin relative to the task's first image (wraps harmlessly
    // when the row is outside the image: such offsets are never dereferenced)
    const unsigned my_goff = (unsigned)lane * (unsigned)(hl * wl) + (unsigned)ay * (unsigned)wl + (unsigned)ax4;

    // ---- stage the windows: item = (query, row, chunk); 32 items per pass ----
    const float* __restrict__ img0 = P.level[l] + (size_t)n0 * hl * wl;
    const bool vec_ok = (wl % 4) == 0;
    int it_q[G::PERIOD], it_row[G::PERIOD], it_ch[G::PERIOD];
#pragma unroll
    for (int k = 0; k < G::PERIOD; ++k) {
        const int item = k * 32 + lane;
        it_q[k] = item / G::IPQ;
        const int rem = item - it_q[k] * G::IPQ;
        it_row[k] = rem / G::CH;
        it_ch[k] = rem - it_row[k] * G::CH;
    }
    constexpr int UNROLL = (G::PERIOD * 4 <= 12) ? 4 : (G::PERIOD * 2 <= 12 ? 2 : 1);  // periods in flight
#pragma unroll 1
    for (int qb = 0; qb < kQT; qb += G::QPP * UNROLL) {
        float4 v[UNROLL * G::PERIOD];
        unsigned st[UNROLL * G::PERIOD];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
            for (int k = 0; k < G::PERIOD; ++k) {
                const int q = qb + u * G::QPP + it_q[k];
                const int src_lane = q & 31;
                const unsigned q_need = __shfl_sync(0xffffffffu, need, src_lane);
                const unsigned q_in = __shfl_sync(0xffffffffu, inimg, src_lane);
                const unsigned q_goff = __shfl_sync(0xffffffffu, my_goff, src_lane);
                // first column of this chunk in the image (only the unaligned path needs it; warp-uniform branch)
                const int q_gx = vec_ok ? 0 : __shfl_sync(0xffffffffu, ax4, src_lane) + it_ch[k] * 4;
                const bool nd = (q_need >> it_row[k]) & (q_need >> (G::PR + it_ch[k])) & 1u;
                const bool ld = (q_in >> it_row[k]) & (q_in >> (G::PR + it_ch[k])) & 1u;
                float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ld) {
                    const float* g = img0 + (q_goff + (unsigned)(it_row[k] * wl + it_ch[k] * 4));
                    if (vec_ok) {
                        val = ldg_f4(g);
                    } else {  // level width not a multiple of 4: rows are unaligned and chunks can straddle the edge
                        val.x = __ldg(g);
                        if (q_gx + 1 < wl) val.y = __ldg(g + 1);
                        if (q_gx + 2 < wl) val.z = __ldg(g + 2);
                        if (q_gx + 3 < wl) val.w = __ldg(g + 3);
                    }
                }
                v[u * G::PERIOD + k] = val;
                st[u * G::PERIOD + k] = nd ? (unsigned)(q * G::QPITCH + it_row[k] * G::PC + it_ch[k] * 4) : 0xffffffffu;
            }
        }
#pragma unroll
        for (int j = 0; j < UNROLL * G::PERIOD; ++j) {
            if (st[j] != 0xffffffffu) {
                float* d = smem + st[j];
                d[0] = v[j].x; d[1] = v[j].y; d[2] = v[j].z; d[3] = v[j].w;
            }
        }
    }
    __syncwarp();

    // ---- 81 outputs per lane, channel-major stores (128 B per warp per channel) ----
    if (valid) {
        constexpr int K = G::SIDE * G::SIDE;
        float* __restrict__ o = P.out + ((size_t)b * P.L * K + (size_t)l * K) * P.HW + p;
        const float* sq = smem + lane * G::QPITCH;
#pragma unroll
        for (int i = 0; i < G::SIDE; ++i) {
#pragma unroll
            for (int j = 0; j < G::SIDE; ++j) {
                const float* t = sq + offy[j] + offx[i];
                float acc = __fmul_rn(t[0], __fmul_rn(tx.wl[i], ty.wl[j]));
                acc = fmaf(t[1], __fmul_rn(tx.wh[i], ty.wl[j]), acc);
                acc = fmaf(t[G::PC], __fmul_rn(tx.wl[i], ty.wh[j]), acc);
                acc = fmaf(t[G::PC + 1], __fmul_rn(tx.wh[i], ty.wh[j]), acc);
                o[(size_t)(i * G::SIDE + j) * P.HW] = dead ? 0.0f : acc;
            }
        }
    }
}

template <int R>
__global__ void __launch_bounds__(128) lookup_taps_kernel(const __grid_constant__ LookupParams P) {
    using G = Geo<R>;
    const long n = (long)blockIdx.x * 128 + threadIdx.x;
    const int l = blockIdx.y;
    if (n >= P.N) return;
    const long b = n / P.HW;
    const int p = (int)(n - b * P.HW);
    const float sc = (float)(1 << l);
    const float cx = __fdiv_rn(__ldg(P.coords + (b * 2 + 0) * P.HW + p), sc);
    const float cy = __fdiv_rn(__ldg(P.coords + (b * 2 + 1) * P.HW + p), sc);
    const float s1x = (float)(P.lw[l] - 1), s1y = (float)(P.lh[l] - 1);
    const size_t base = (((size_t)l * P.N + n) * 2) * G::SIDE;
#pragma unroll
    for (int i = 0; i < G::SIDE; ++i) {
        const float ux = lookup_coord(__fadd_rn(cx, (float)(i - R)), s1x);
        const float uy = lookup_coord(__fadd_rn(cy, (float)(i - R)), s1y);
        P.taps[base + i] = ux;
        P.taps[base + G::SIDE + i] = uy;
        P.idx[base + i] = __float2int_rd(ux);
        P.idx[base + G::SIDE + i] = __float2int_rd(uy);
    }
}

template <int R>
int launch_lookup_r(const LookupParams& P, cudaStream_t s) {
    using G = Geo<R>;
    const size_t smem = (size_t)kQT * G::QPITCH * sizeof(float);
    const long ngroups = (P.N + kQT - 1) / kQT;
    const long grid = ngroups * P.L;
    if (grid > 0x7fffffffL) return fail(CF_ERR_UNSUPPORTED, "lookup grid too large (%ld tasks)", grid);
    lookup_kernel<R><<<(unsigned)grid, 32, smem, s>>>(P);
    CF_CHECK_LAUNCH("lookup_kernel");
    return CF_OK;
}

template <int R>
int launch_taps_r(const LookupParams& P, cudaStream_t s) {
    dim3 grid((unsigned)((P.N + 127) / 128), (unsigned)P.L);
    lookup_taps_kernel<R><<<grid, 128, 0, s>>>(P);
    CF_CHECK_LAUNCH("lookup_taps_kernel");
    return CF_OK;
}

LookupParams make_params(const float* pyramid, const LevelDims& d, const float* coords, int B, int H, int W, int L) {
    LookupParams P{};
    for (int l = 0; l < L; ++l) {
        P.level[l] = pyramid ? reinterpret_cast<const float*>(reinterpret_cast<const char*>(pyramid) + d.off[l]) : nullptr;
        P.lh[l] = d.h[l];
        P.lw[l] = d.w[l];
    }
    P.coords = coords;
    P.N = (long)B * H * W;
    P.HW = H * W;
    P.L = L;
    return P;
}

}  // namespace

int launch_lookup(const float* pyramid, const LevelDims& d, const float* coords, int B, int H, int W, int L, int r,
                  float* out, cudaStream_t s) {
    LookupParams P = make_params(pyramid, d, coords, B, H, W, L);
    P.out = out;
    switch (r) {
        case 1: return launch_lookup_r<1>(P, s);
        case 2: return launch_lookup_r<2>(P, s);
        case 3: return launch_lookup_r<3>(P, s);
        case 4: return launch_lookup_r<4>(P, s);
    }
    return fail(CF_ERR_UNSUPPORTED, "lookup radius %d not in [1,%d]", r, CF_MAX_RADIUS);
}

int launch_lookup_taps(const LevelDims& d, const float* coords, int B, int H, int W, int L, int r, float* taps,
                       int32_t* idx, cudaStream_t s) {
    LookupParams P = make_params(nullptr, d, coords, B, H, W, L);
    P.taps = taps;
    P.idx = idx;
    switch (r) {
        case 1: return launch_taps_r<1>(P, s);
        case 2: return launch_taps_r<2>(P, s);
        case 3: return launch_taps_r<3>(P, s);
        case 4: return launch_taps_r<4>(P, s);
    }
    return fail(CF_ERR_UNSUPPORTED, "lookup radius %d not in [1,%d]", r, CF_MAX_RADIUS);
}

}  // namespace cf
