// K3: CorrBlock.__call__ -- radius-r bilinear window lookup over all pyramid levels in one launch.
//
// Replaces torchvision raft.py:394-422 (call site /root/reference/nnunet/network_architecture/
// SegFlowGaussian.py:935, 12x per frame pair): per level a delta build, add, 2 normalise kernels, cat,
// grid_sampler_2d over B*HW tiny images, then a final cat + permute + contiguous copy (~25-30 launches).
//
// Design (HBM-bound, SURVEY.md 8(d): out 4*L*K*HW + coords + window footprint per call):
//   * one CTA = 32 consecutive query pixels x one pyramid level, NW warps; lane <-> query in every warp
//   * every tap coordinate goes through the reference's exact fp32 sequence (normalise -> un-normalise, no
//     FMA contraction), so tap indices / weights are bit-exact.  The 2*(2r+1) per-axis taps of a query are
//     computed ONCE per CTA (split over the warps) and the x taps are shared through shared memory
//   * the (2r+4)^2 window each query can touch is staged into shared memory with 128-bit read-only loads:
//     a warp pass covers 8 queries x 4 sixteen-byte chunks of ONE window row (8 distinct 128 B lines per
//     request instead of 32), stored shift-corrected so that all queries see the same logical column ->
//     bank = f(query) only: conflict-free loads in the compute phase (query pitch is odd); out-of-image
//     positions are stored as zeros, so zero padding costs nothing in the inner loop
//   * warp w evaluates the window rows j = w*JPW .. of the (2r+1)^2 outputs from shared memory in ATen's
//     nw, ne, sw, se order.  When no lane of the warp sits within 2^-8 of an integer coordinate (where the
//     round trip can flip a floor) the taps are provably the nominal ones and a register-window fast path
//     with static indexing runs; otherwise the general path with per-lane offsets
//   * outputs go straight into the final [B, L*K, H, W] layout: for a fixed channel the 32 lanes write 128
//     contiguous bytes (no cat / permute pass)
#include <cstdlib>

#include "common.cuh"

namespace cf {
namespace {

constexpr int kQT = 32;  // queries per CTA

template <int R>
struct Geo {
    static constexpr int SIDE = 2 * R + 1;
    static constexpr int PR = 2 * R + 4;        // window rows = logical columns: nominal 2r+2, +-1 for floor flips
    static constexpr int QP = PR * PR + 1;      // words per query, odd
    static constexpr int XT = SIDE * 3 * kQT;   // shared x taps: [SIDE][offset, w_lo, w_hi][query]
    static constexpr int TR = PR * PR + kQT;   // trash words for dropped chunk components
    static constexpr size_t SMEM = (size_t)(kQT * QP + XT + TR) * sizeof(float);
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

// one axis tap: sampling coordinate -> floor index, weight of tap i0 and of tap i0 + 1
__device__ __forceinline__ void axis_tap(float c, int d, float s1, int& i0, float& w_lo, float& w_hi) {
    const float u = lookup_coord(__fadd_rn(c, (float)d), s1);
    const float f = floorf(u);
    i0 = __float2int_rd(u);
    w_hi = __fsub_rn(u, f);
    w_lo = __fsub_rn(__fadd_rn(f, 1.0f), u);
}

// The round trip moves a coordinate p = c + k (|k| <= r) by less than 2^-24 * (4|p| + S) (four roundings of
// magnitude <= |p| resp. S/2).  If frac(c) keeps twice that margin from both integers, floor(u_k) == floor(c) + k
// for every tap: the nominal window, no flipped floors.
template <int R>
__device__ __forceinline__ bool nominal_taps(float c, int size) {
    const float fr = __fsub_rn(c, floorf(c));
    const float margin = fmaf(4.0f, fabsf(c) + (float)(R + 1), (float)size) * 1.1920929e-7f;   // 2 * bound
    return fabsf(c) <= 65536.0f && fr >= margin && fr <= 1.0f - margin;
}

// predicated read-only loads (forced predication: the staging loop must not turn into branches, so that all
// loads of a group are in flight together)
// The pyramid is re-read by every lookup of a CorrBlock (12x per frame pair) while each lookup streams its output
// once: pyramid loads carry an L2 evict_last policy, output stores are streaming (st.global.cs).
__device__ __forceinline__ uint64_t l2_evict_last_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ float4 ldg_f4_if(const float* p, bool pred, uint64_t pol) {
    float4 v;
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\tmov.b32 %2, 0;\n\tmov.b32 %3, 0;\n\t"
        "@p ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %6;\n\t}"
        : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
        : "l"(p), "r"((int)pred), "l"(pol));
    return v;
}
__device__ __forceinline__ float ldg_f1_if(const float* p, bool pred, uint64_t pol) {
    float v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.b32 %0, 0;\n\t"
                 "@p ld.global.nc.L2::cache_hint.f32 %0, [%1], %3;\n\t}"
                 : "=f"(v) : "l"(p), "r"((int)pred), "l"(pol));
    return v;
}

// Stage the (2r+4)^2 windows of the CTA's 32 queries.  A pass = one window row of 8 queries x 4 sixteen-byte
// chunks; warp w takes rows w, w+NW, ...  VEC: level width is a multiple of 4, chunks are 16 B-aligned in global
// memory and stored shift-corrected; otherwise chunks start at the anchor column and are read element-wise.
template <int R, int NW, bool VEC>
__device__ __forceinline__ void stage_windows(float* __restrict__ smem, const float* __restrict__ img0, int hl, int wl,
                                              int w, int lane, int ax, int ay, unsigned flags) {
    using G = Geo<R>;
    constexpr int PR = G::PR, QP = G::QP;
    constexpr int RPW = (PR + NW - 1) / NW;
    constexpr int TRASH = kQT * QP + G::XT;   // dropped components land here (lane-private word + row offset)
    const int qq = lane >> 2, ch = lane & 3;
    const int img_words = hl * wl;
    const uint64_t pol = l2_evict_last_policy();
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const int src = 4 * g + (qq & 3) + 16 * (qq >> 2);   // query served by this lane in group g
        const int q_ax = __shfl_sync(0xffffffffu, ax, src);
        const int q_ay = __shfl_sync(0xffffffffu, ay, src);
        const unsigned q_fl = __shfl_sync(0xffffffffu, flags, src);
        const bool nomx = q_fl & 2u, nomy = q_fl & 4u;
        const int clo = nomx ? 1 : 0, chi = nomx ? PR - 2 : PR - 1;   // logical columns the taps can read
        const int gx = VEC ? (q_ax & ~3) + 4 * ch : q_ax + 4 * ch;  // first global column of the chunk
        const int c0 = VEC ? 4 * ch - (q_ax & 3) : 4 * ch;           // logical column of component 0
        const bool ch_need = !(q_fl & 1u) && (c0 + 3 >= clo) && (c0 <= chi);
        const bool col_in = VEC ? (unsigned)gx < (unsigned)wl : true;
        const char* gq = reinterpret_cast<const char*>(img0 + ((long)src * img_words + (long)(q_ay + w) * wl + gx));
        const unsigned row_bytes = (unsigned)(NW * wl) * 4u;
        int sidx[4];   // shared-memory word of each chunk component at this warp's first row
#pragma unroll
        for (int e = 0; e < 4; ++e)
            sidx[e] = ((unsigned)(c0 + e) < (unsigned)PR ? src * QP + c0 + e : TRASH + lane) + w * PR;
        float4 v[RPW];
        bool wr[RPW];
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
            const int r = w + NW * k;
            const bool edge = (r == 0) || (r == PR - 1);            // rows only a flipped floor can reach
            wr[k] = ch_need && r < PR && !(edge && nomy);
            const bool ld = wr[k] && col_in && (unsigned)(q_ay + r) < (unsigned)hl;
            const float* gp = reinterpret_cast<const float*>(gq + (unsigned)k * row_bytes);
            if (VEC) {
                v[k] = ldg_f4_if(gp, ld, pol);
            } else {
                v[k].x = ldg_f1_if(gp + 0, ld && (unsigned)(gx + 0) < (unsigned)wl, pol);
                v[k].y = ldg_f1_if(gp + 1, ld && (unsigned)(gx + 1) < (unsigned)wl, pol);
                v[k].z = ldg_f1_if(gp + 2, ld && (unsigned)(gx + 2) < (unsigned)wl, pol);
                v[k].w = ldg_f1_if(gp + 3, ld && (unsigned)(gx + 3) < (unsigned)wl, pol);
            }
        }
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
            if (wr[k]) {
                smem[sidx[0] + NW * k * PR] = v[k].x;
                smem[sidx[1] + NW * k * PR] = v[k].y;
                smem[sidx[2] + NW * k * PR] = v[k].z;
                smem[sidx[3] + NW * k * PR] = v[k].w;
            }
        }
    }
}

template <int R, int NW>
__global__ void __launch_bounds__(32 * NW) lookup_kernel(const __grid_constant__ LookupParams P) {
    using G = Geo<R>;
    constexpr int SIDE = G::SIDE, PR = G::PR, QP = G::QP;
    constexpr int JPW = (SIDE + NW - 1) / NW;   // output rows (and taps per axis) per warp
    extern __shared__ float smem[];
    float* const win = smem;                    // [kQT][QP]
    float* const xt = smem + kQT * QP;          // [SIDE][3][kQT]

    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned task = blockIdx.x / (unsigned)P.L;
    const int l = (int)(blockIdx.x - task * (unsigned)P.L);
    const unsigned n0 = task * kQT;
    const unsigned n = n0 + lane;
    const bool valid = n < (unsigned)P.N;
    const int hl = P.lh[l], wl = P.lw[l];

    // ---- per-lane query (every warp holds the same 32 queries) ----
    unsigned b = 0, p = 0;
    float cx = 0.0f, cy = 0.0f;
    if (valid) {
        b = n / (unsigned)P.HW;
        p = n - b * (unsigned)P.HW;
        const float inv = 1.0f / (float)(1 << l);   // division by 2^l == multiplication by 2^-l, exactly
        const float* cp = P.coords + (size_t)b * 2 * P.HW + p;
        cx = __fmul_rn(__ldg(cp), inv);
        cy = __fmul_rn(__ldg(cp + P.HW), inv);
    }
    // every tap of a query this far out is outside any image -> zeros (also catches NaN / inf)
    const bool dead = !valid || !(fabsf(cx) < 1.0e6f) || !(fabsf(cy) < 1.0e6f);
    // window anchor: nominal first tap minus one (the round trip can move a floor by +-1)
    const int ax = dead ? 0 : __float2int_rd(cx) - R - 1;
    const int ay = dead ? 0 : __float2int_rd(cy) - R - 1;
    const bool nom_x = dead || nominal_taps<R>(cx, wl);
    const bool nom_y = dead || nominal_taps<R>(cy, hl);
    const bool fast = __all_sync(0xffffffffu, nom_x && nom_y);

    // ---- taps: warp w computes taps w*JPW.. of both axes; x taps are shared, y taps stay in registers ----
    int yoff[JPW];
    float ywl[JPW], ywh[JPW];
    {
        const float s1x = (float)(wl - 1), s1y = (float)(hl - 1);
#pragma unroll
        for (int k = 0; k < JPW; ++k) {
            const int t = w * JPW + k;
            yoff[k] = 0; ywl[k] = 0.0f; ywh[k] = 0.0f;
            if (t < SIDE) {
                int i0;
                float lo, hi;
                axis_tap(cx, t - R, s1x, i0, lo, hi);
                xt[(t * 3 + 0) * kQT + lane] = __int_as_float(min(max(i0 - ax, 0), PR - 2));
                xt[(t * 3 + 1) * kQT + lane] = lo;
                xt[(t * 3 + 2) * kQT + lane] = hi;
                axis_tap(cy, t - R, s1y, i0, ywl[k], ywh[k]);
                yoff[k] = min(max(i0 - ay, 0), PR - 2) * PR;
            }
        }
    }

    // ---- stage the windows ----
    {
        const float* __restrict__ img0 = P.level[l] + (size_t)n0 * hl * wl;
        const unsigned flags = (dead ? 1u : 0u) | (nom_x ? 2u : 0u) | (nom_y ? 4u : 0u);
        if ((wl & 3) == 0) stage_windows<R, NW, true>(smem, img0, hl, wl, w, lane, ax, ay, flags);
        else stage_windows<R, NW, false>(smem, img0, hl, wl, w, lane, ax, ay, flags);
    }
    __syncthreads();

    // ---- (2r+1) x JPW outputs per lane, channel-major stores (128 B per warp per channel) ----
    if (!valid) return;
    constexpr int K = SIDE * SIDE;
    const unsigned HW = (unsigned)P.HW;
    const int j0 = w * JPW;
    float* __restrict__ o = P.out + ((size_t)b * P.L * K + (size_t)l * K + j0) * HW + p;
    const float* sq = win + lane * QP;
    if (dead) {
#pragma unroll
        for (int k = 0; k < JPW; ++k)
            if (j0 + k < SIDE)
#pragma unroll
                for (int i = 0; i < SIDE; ++i) __stcs(o + (i * SIDE + k) * HW, 0.0f);
        return;
    }
    if (fast) {
        // nominal taps: x tap i reads logical columns i+1, i+2; output row j reads window rows j+1, j+2
        float xl[SIDE], xh[SIDE];
#pragma unroll
        for (int i = 0; i < SIDE; ++i) {
            xl[i] = xt[(i * 3 + 1) * kQT + lane];
            xh[i] = xt[(i * 3 + 2) * kQT + lane];
        }
        const float* rp = sq + (j0 + 1) * PR + 1;
        float ra[SIDE + 1], rb[SIDE + 1];
#pragma unroll
        for (int c = 0; c <= SIDE; ++c) ra[c] = rp[c];
#pragma unroll
        for (int k = 0; k < JPW; ++k) {
            if (j0 + k < SIDE) {
#pragma unroll
                for (int c = 0; c <= SIDE; ++c) rb[c] = rp[(k + 1) * PR + c];
#pragma unroll
                for (int i = 0; i < SIDE; ++i) {
                    float acc = __fmul_rn(ra[i], __fmul_rn(xl[i], ywl[k]));
                    acc = fmaf(ra[i + 1], __fmul_rn(xh[i], ywl[k]), acc);
                    acc = fmaf(rb[i], __fmul_rn(xl[i], ywh[k]), acc);
                    acc = fmaf(rb[i + 1], __fmul_rn(xh[i], ywh[k]), acc);
                    __stcs(o + (i * SIDE + k) * HW, acc);
                }
#pragma unroll
                for (int c = 0; c <= SIDE; ++c) ra[c] = rb[c];
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < SIDE; ++i) {
            const int xo = __float_as_int(xt[(i * 3 + 0) * kQT + lane]);
            const float xl = xt[(i * 3 + 1) * kQT + lane], xh = xt[(i * 3 + 2) * kQT + lane];
#pragma unroll
            for (int k = 0; k < JPW; ++k) {
                if (j0 + k < SIDE) {
                    const float* t = sq + yoff[k] + xo;
                    float acc = __fmul_rn(t[0], __fmul_rn(xl, ywl[k]));
                    acc = fmaf(t[1], __fmul_rn(xh, ywl[k]), acc);
                    acc = fmaf(t[PR], __fmul_rn(xl, ywh[k]), acc);
                    acc = fmaf(t[PR + 1], __fmul_rn(xh, ywh[k]), acc);
                    __stcs(o + (i * SIDE + k) * HW, acc);
                }
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

template <int R, int NW>
int launch_lookup_rw(const LookupParams& P, long grid, cudaStream_t s) {
    lookup_kernel<R, NW><<<(unsigned)grid, 32 * NW, Geo<R>::SMEM, s>>>(P);
    CF_CHECK_LAUNCH("lookup_kernel");
    return CF_OK;
}

template <int R>
int launch_lookup_r(const LookupParams& P, cudaStream_t s) {
    const long ngroups = (P.N + kQT - 1) / kQT;
    const long grid = ngroups * P.L;
    if (P.N > 0x7fffffffL || grid > 0x7fffffffL) return fail(CF_ERR_UNSUPPORTED, "lookup grid too large (%ld tasks)", grid);
    static const int nw = [] { const char* e = getenv("CF_LOOKUP_NW"); return e ? atoi(e) : 3; }();   // tuning knob
    switch (nw) {
        case 2: return launch_lookup_rw<R, 2>(P, grid, s);
        case 4: return launch_lookup_rw<R, 4>(P, grid, s);
        default: return launch_lookup_rw<R, 3>(P, grid, s);
    }
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
