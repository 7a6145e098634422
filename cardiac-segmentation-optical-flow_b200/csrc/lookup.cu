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
//   * the (2r+3)^2 window each query can touch (anchored by the side on which its floors can flip) is staged into
//     shared memory with asynchronous 16-byte copies (cp.async / LDGSTS, L2 evict_last): a warp pass covers 8
//     queries x 4 chunks of ONE window row (8 distinct 128 B lines per request instead of 32); out-of-image
//     chunks are zero-filled by the copy itself, so zero padding costs nothing in the inner loop; the tap
//     arithmetic runs while the copies are in flight
//   * the division by (size - 1) in the coordinate round trip is an exact reciprocal-multiply-correct sequence
//     (common.cuh: div_small, bit-identical to IEEE division, verified exhaustively on B200)
//   * warp w evaluates the window rows j = w*JPW .. of the (2r+1)^2 outputs from shared memory in ATen's
//     nw, ne, sw, se order.  When no lane of the warp sits within ~1e-5 of an integer coordinate (where the
//     round trip can flip a floor) the taps are provably the nominal ones and a register-window fast path
//     with static indexing runs; otherwise the general path with per-lane offsets
//   * outputs go straight into the final [B, L*K, H, W] layout: for a fixed channel the 32 lanes write 128
//     contiguous bytes (no cat / permute pass)
#include <cstdlib>

#include "common.cuh"
#include "sampling.cuh"

namespace cf {
namespace {

constexpr int kQT = 32;  // queries per CTA

#ifdef CF_BRINGUP   // per-CTA phase timeline (tools/lookup_ts.py); not in the shipped build
__device__ unsigned long long* g_lookup_ts = nullptr;
__device__ __forceinline__ void ts_stamp(int slot, float dep = 0.0f) {
    if (g_lookup_ts && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "f"(dep) : "memory");
        g_lookup_ts[(size_t)blockIdx.x * 8 + slot] = t;
        if (slot == 0) {
            unsigned sm;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
            g_lookup_ts[(size_t)blockIdx.x * 8 + 7] = sm;
        }
    }
}
#define CF_TS(slot, dep) ts_stamp(slot, dep)
#else
#define CF_TS(slot, dep)
#endif

// Window geometry.  The round trip moves a coordinate by far less than 0.5, so the floors of a query's taps can only
// flip DOWN when frac(c) < 0.5 and only UP otherwise: anchoring the window of a query that can flip at
// floor(c) - r - (frac < 0.5) makes 2r+3 rows / columns enough for every case; a nominal query (no flip possible) is
// anchored at floor(c) - r and uses the first 2r+2.
template <int R>
struct Geo {
    static constexpr int SIDE = 2 * R + 1;
    static constexpr int PR = 2 * R + 3;                 // window rows / logical columns
    static constexpr int PC = ((PR + 3 + 3) / 4) * 4;    // stored columns: PR plus the alignment shift (<= 3), in 16 B chunks
    static_assert(PC <= 16, "a window row is staged as at most 4 sixteen-byte chunks");
    static constexpr int QP = PR * PC + (((PR * PC / 4) % 2 == 0) ? 4 : 0);   // words per query: QP / 4 odd
    static_assert(QP % 4 == 0 && (QP / 4) % 2 == 1, "window pitch / 4 must be odd");
    static constexpr int WIN = kQT * QP + 16;            // + the 16 B skews of win_base()
    // 23.1 KB per CTA for r = 4: nine CTAs per SM.  (The x taps used to be computed once per CTA and shared through another
    // 3.4 KB of shared memory: 8 CTAs per SM, 3.14 waves of CTAs at cfg2 = a fourth, nearly empty round.  Every warp now
    // computes the x taps of its queries itself: ~130 more ALU instructions per warp, no shared-memory round trip.)
    static constexpr size_t SMEM = (size_t)WIN * sizeof(float);
};

struct LookupParams {
    const float* level[CF_MAX_LEVELS];
    int lh[CF_MAX_LEVELS];
    int lw[CF_MAX_LEVELS];
    float rcp_h[CF_MAX_LEVELS];   // 1 / (h_l - 1), 1 / (w_l - 1): IEEE-rounded on the host (div_small)
    float rcp_w[CF_MAX_LEVELS];
    const float* coords;
    float* out;
    float* taps;    // taps kernel only
    int32_t* idx;   // taps kernel only
    long N;         // B*H*W
    int HW;
    int L;
    int lshift;     // log2(L) when L is a power of two, else -1
    unsigned hw_magic;  // ceil(2^32 / HW)
    float l2_frac;  // fraction of pyramid lines fetched with L2 evict_last priority
    unsigned tiled; // bit l: level l is stored as 4x4 tiles (cardioflow.h: CF_LAYOUT_TILED4)
};

// flags a lane publishes about its query (shuffled to the lanes that stage its window)
constexpr unsigned kDead = 1u, kNomX = 2u, kNomY = 4u, kLoX = 8u, kLoY = 16u;

// The round trip moves a coordinate p = c + k (|k| <= r) by less than 2^-24 * (4|p| + S) (four roundings of
// magnitude <= |p| resp. S/2).  If frac(c) keeps twice that margin from both integers, floor(u_k) == floor(c) + k
// for every tap: the nominal window, no flipped floors.  `lo` = frac(c) < 0.5 (floors can only flip down).
template <int R>
__device__ __forceinline__ void classify(float c, int size, bool& nominal, bool& lo) {
    const float fr = __fsub_rn(c, floorf(c));
    const float margin = fmaf(4.0f, fabsf(c) + (float)(R + 1), (float)size) * 1.1920929e-7f;   // 2 * bound
    nominal = fr >= margin && fr <= 1.0f - margin;
    lo = fr < 0.5f;
}

// The pyramid is re-read by every lookup of a CorrBlock (12x per frame pair) while each lookup streams its output
// once: pyramid reads carry an L2 evict_last policy, output stores are streaming (st.global.cs).
__device__ __forceinline__ uint64_t l2_evict_last_policy(float frac = 1.0f) {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, %1;" : "=l"(pol) : "f"(frac));
    return pol;
}
// asynchronous global -> shared copies (LDGSTS): no staging registers, every copy of a warp is in flight at once;
// copies outside the image write zeros (zero padding) and read nothing
// 16-byte copy, zero-filled outside the image: the copy's own `ignore-src` predicate is formed inside the asm from the two unsigned
// bounds checks (row >= rows || col >= cols): 2 instructions per copy instead of the 5 the compiler needs to turn a bool into a byte count
__device__ __forceinline__ void cp_async_16_bounds(uint32_t dst, const void* src, unsigned row, unsigned rows, unsigned col, unsigned cols,
                                                   uint64_t pol) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ge.u32 p, %2, %3;\n\t"
        "setp.ge.or.u32 p, %4, %5, p;\n\t"
        "cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, p, %6;\n\t}"
        ::"r"(dst), "l"(src), "r"(row), "r"(rows), "r"(col), "r"(cols), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_4(uint32_t dst, const void* src, uint32_t src_bytes, uint64_t pol) {
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2, %3;"
                 ::"r"(dst), "l"(src), "r"(src_bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// word offset of query q's window: pitch QP (a multiple of 4 words: 16 B-aligned chunks) plus a 16 B skew every
// 8 queries, which makes the compute phase's scalar reads (column = per-lane shift + constant) conflict-free
// when consecutive queries have consecutive anchors
template <int R>
__device__ __forceinline__ int win_base(int q) { return q * Geo<R>::QP + 4 * (q >> 3); }

// Stage the windows of the CTA's 32 queries.  A pass = one window row of 8 queries x 4 sixteen-byte chunks; warp w
// takes rows w, w+NW, ...  VEC: level width is a multiple of 4, chunks are 16 B-aligned in global memory (window
// column a <-> image column (ax & ~3) + a); otherwise chunks start at the anchor column and are copied element-wise.
// TILED (implies VEC): the level image is stored as 4x4 tiles (64 B each, tile-row-major), so the 16 B chunk (row y, columns
// gx..gx+3) lives at ((y >> 2) * (wl / 4) + gx / 4) * 16 + (y & 3) * 4 words and the four rows of a tile share one DRAM atom.
template <int R, int NW, bool VEC, bool TILED = false>
__device__ __forceinline__ void stage_windows(uint32_t win_s, const float* __restrict__ img0, int hl, int wl, int w,
                                              int lane, int ax, int ay, unsigned flags, float l2_frac = 1.0f) {
    using G = Geo<R>;
    constexpr int PR = G::PR, PC = G::PC;
    constexpr int RPW = (PR + NW - 1) / NW;
    const int qq = lane >> 2, ch = lane & 3;
    const int img_words = hl * wl;
    const uint64_t pol = l2_evict_last_policy(l2_frac);
    // Owner side, once per lane: which of the 4 chunks its query's taps can read (nominal = the first 2r+2 columns,
    // else all PR) and the one row only a flipped floor can reach, packed for the staging lanes:
    // bit 0 dead, bits 1-4 chunk needed, bits 5-8 skipped window row (15 = none)
    unsigned info;
    {
        // a nominal query's window is anchored AT its first tap (columns / rows 0 .. 2r+1); only a query whose floors can flip
        // needs the extra column / row
        const int osh = VEC ? (ax & 3) : 0;
        const int c1 = (flags & kNomX) ? 2 * R + 1 : PR - 1;
        unsigned need = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if ((4 * c <= osh + c1) && (4 * c < PC)) need |= 1u << c;
        const unsigned oskip = (flags & kNomY) ? (unsigned)(PR - 1) : 15u;
        info = ((flags & kDead) ? 1u : 0u) | (need << 1) | (oskip << 5);
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const int src = 4 * g + (qq & 3) + 16 * (qq >> 2);   // query served by this lane in group g
        const int q_ax = __shfl_sync(0xffffffffu, ax, src);
        const int q_ay = __shfl_sync(0xffffffffu, ay, src);
        const unsigned q_in = __shfl_sync(0xffffffffu, info, src);
        const int gx = (VEC ? (q_ax & ~3) : q_ax) + 4 * ch;          // first image column of the chunk
        const bool ch_need = !(q_in & 1u) && ((q_in >> (1 + ch)) & 1u);
        const int skip = (int)((q_in >> 5) & 15u);
        const char* gq = TILED ? reinterpret_cast<const char*>(img0 + ((long)src * img_words + (long)gx * 4))
                               : reinterpret_cast<const char*>(img0 + ((long)src * img_words + (long)(q_ay + w) * wl + gx));
        const unsigned row_bytes = (unsigned)(NW * wl) * 4u;
        const int tile_row_bytes = wl * 16;                             // TILED: bytes of one row of tiles
        const uint32_t sq = win_s + (uint32_t)(win_base<R>(src) + w * PC + 4 * ch) * 4u;
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
            const int r = w + NW * k;
            if (ch_need && r < PR && r != skip) {
                const bool row_in = (unsigned)(q_ay + r) < (unsigned)hl;
                const int yrow = q_ay + r;
                const char* gp = TILED ? gq + ((yrow >> 2) * tile_row_bytes + (yrow & 3) * 16) : gq + (unsigned)k * row_bytes;
                const uint32_t d = sq + (uint32_t)(NW * k * PC) * 4u;
                if (VEC) {
                    // outside the image: 16 zero bytes are written and nothing is read.  (A/B-tested under ncu: redirecting these requests
                    // to one valid address changes neither the L2 sector counts nor the time -- profiles/r1_zfill_ab.log.)
                    cp_async_16_bounds(d, gp, (unsigned)yrow, (unsigned)hl, (unsigned)gx, (unsigned)wl, pol);
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const bool in = row_in && (unsigned)(gx + e) < (unsigned)wl;
                        cp_async_4(d + 4u * e, gp + 4 * e, in ? 4u : 0u, pol);
                    }
                }
            }
        }
    }
}

template <int R, int NW, bool BARREL>
__global__ void __launch_bounds__(32 * NW, NW == 3 ? 9 : 1) lookup_kernel(const __grid_constant__ LookupParams P) {
    using G = Geo<R>;
    constexpr int SIDE = G::SIDE, PR = G::PR, PC = G::PC;
    constexpr int JPW = (SIDE + NW - 1) / NW;   // output rows (and taps per axis) per warp
    extern __shared__ __align__(16) float smem[];
    float* const win = smem;                    // windows: win_base(q) + row * PC + column

    // Programmatic dependent launch: let the NEXT kernel of the stream (normally the next lookup) be scheduled while this
    // grid drains -- its CTAs become resident and run their index arithmetic, then block in griddepcontrol.wait below
    // until this grid has completed and flushed.  No effect on launches made without the PDL attribute.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    CF_TS(0, 0.0f);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    // levels interleave in launch order (read-heavy fine levels next to store-heavy coarse ones)
    const unsigned task = P.lshift >= 0 ? (blockIdx.x >> P.lshift) : blockIdx.x / (unsigned)P.L;
    const int l = (int)(blockIdx.x - task * (unsigned)P.L);
    const unsigned n0 = task * kQT;             // CTA-uniform
    const bool valid = n0 + lane < (unsigned)P.N;
    const int hl = P.lh[l], wl = P.lw[l];
    const bool vec = (wl & 3) == 0;

    // ---- per-lane query (every warp holds the same 32 queries); batch index from CTA-uniform arithmetic ----
    // n0 / HW by a host-computed reciprocal: ceil(2^32 / HW) over-estimates by at most one for n0 < 2^31
    unsigned b = __umulhi(n0, P.hw_magic);
    if (b * (unsigned)P.HW > n0) --b;
    unsigned p = n0 - b * (unsigned)P.HW + lane;
    while (p >= (unsigned)P.HW) { p -= (unsigned)P.HW; ++b; }   // one wrap at most when H*W >= 32; tiny maps (L <= 2) wrap more
    float cx = 0.0f, cy = 0.0f;
    // everything below reads memory the preceding kernels wrote (coords, pyramid) and writes `out`
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (valid) {
        const float inv = __int_as_float((127 - l) << 23);      // 2^-l: the reference's division by 2^l, exactly
        const float* cp = P.coords + (size_t)b * 2 * P.HW + p;
        cx = __fmul_rn(__ldg(cp), inv);
        cy = __fmul_rn(__ldg(cp + P.HW), inv);
    }
    // every tap of a query this far out is outside any image -> zeros (also catches NaN / inf)
    CF_TS(1, cx + cy);
    const bool dead = !valid || !(fabsf(cx) < kDeadCoord) || !(fabsf(cy) < kDeadCoord);
    bool nom_x, nom_y, lo_x, lo_y;
    classify<R>(cx, wl, nom_x, lo_x);
    classify<R>(cy, hl, nom_y, lo_y);
    // window anchor: the first nominal tap -- minus one only for a query whose floors can actually flip down.  Nominal queries
    // (all of them on the fast path) then read their taps at the same window rows / columns up to the alignment shift `sh`:
    // with the anchor always one lower for frac < 0.5 the per-lane read offset took 10 values instead of 4 and the shared-memory
    // reads of a warp collided 3.8-fold (ncu: 47 % of the wavefronts were bank-conflict replays; the kernel is LSU-bound)
    const int ax = dead ? 0 : __float2int_rd(cx) - R - ((!nom_x && lo_x) ? 1 : 0);
    const int ay = dead ? 0 : __float2int_rd(cy) - R - ((!nom_y && lo_y) ? 1 : 0);
    const int sh = vec ? (ax & 3) : 0;             // stored column of the anchor
    const bool fast = __all_sync(0xffffffffu, dead || (nom_x && nom_y));

    // ---- start the window copies, then compute the taps while they are in flight ----
    {
        const float* __restrict__ img0 = P.level[l] + (size_t)n0 * hl * wl;
        const unsigned flags = (dead ? kDead : 0u) | (nom_x ? kNomX : 0u) | (nom_y ? kNomY : 0u) | (lo_x ? kLoX : 0u) |
                               (lo_y ? kLoY : 0u);
        const uint32_t win_s = (uint32_t)__cvta_generic_to_shared(win);
        if ((P.tiled >> l) & 1u) stage_windows<R, NW, true, true>(win_s, img0, hl, wl, w, lane, ax, ay, flags, P.l2_frac);
        else if (vec) stage_windows<R, NW, true>(win_s, img0, hl, wl, w, lane, ax, ay, flags, P.l2_frac);
        else stage_windows<R, NW, false>(win_s, img0, hl, wl, w, lane, ax, ay, flags, P.l2_frac);
    }

    CF_TS(2, 0.0f);
    // taps (registers): every warp computes the x taps of its 32 queries and the y taps of its own output rows w*JPW..
    int yoff[JPW], xo[SIDE];
    float ywl[JPW], ywh[JPW], xwl[SIDE], xwh[SIDE];
    {
        const float s1x = (float)(wl - 1), s1y = (float)(hl - 1);
        const float rx = P.rcp_w[l], ry = P.rcp_h[l];
#pragma unroll
        for (int i = 0; i < SIDE; ++i) {
            int i0;
            axis_tap(cx, i - R, s1x, rx, i0, xwl[i], xwh[i]);
            xo[i] = min(max(i0 - ax, 0), PR - 2) + sh;
        }
#pragma unroll
        for (int k = 0; k < JPW; ++k) {
            const int t = w * JPW + k;
            yoff[k] = 0; ywl[k] = 0.0f; ywh[k] = 0.0f;
            if (t < SIDE) {
                int i0;
                axis_tap(cy, t - R, s1y, ry, i0, ywl[k], ywh[k]);
                yoff[k] = min(max(i0 - ay, 0), PR - 2) * PC;
            }
        }
    }
    CF_TS(3, ywl[0]);
    cp_async_wait_all();
    __syncthreads();
    CF_TS(4, 0.0f);

    // ---- (2r+1) x JPW outputs per lane, channel-major stores (128 B per warp per channel) ----
    if (!valid) return;
    constexpr int K = SIDE * SIDE;
    const unsigned HW = (unsigned)P.HW;
    const int j0 = w * JPW;
    // one store pointer per output row, advanced by SIDE channels per x tap; kept as byte addresses with a byte stride so
    // that a step is one 64-bit add (element indices cost an extra LEA pair per store)
    char* ok[JPW];
#pragma unroll
    for (int k = 0; k < JPW; ++k)
        ok[k] = reinterpret_cast<char*>(P.out + ((size_t)b * P.L * K + (size_t)l * K + j0 + k) * HW + p);
    const size_t ostep = (size_t)SIDE * HW * sizeof(float);
    const float* sq = win + win_base<R>(lane);
    if (dead) {
#pragma unroll
        for (int i = 0; i < SIDE; ++i)
#pragma unroll
            for (int k = 0; k < JPW; ++k)
                if (j0 + k < SIDE) { __stcs(reinterpret_cast<float*>(ok[k]), 0.0f); ok[k] += ostep; }
        return;
    }
    if (fast) {
        // nominal taps: x tap i reads stored columns sh+i, +1; output row j reads window rows j, +1.
        // The kernel issues on 63 % of the cycles (ncu), a fifth of them these blends: two output rows are blended at once with the
        // packed fp32 instructions of sm_100 (FMUL2 / FFMA2: IEEE round-to-nearest per component, so every output keeps the exact
        // sequence fmul, fma, fma, fma in ATen's nw, ne, sw, se order).  Pair kp = output rows (2kp, 2kp+1) from window rows 2kp .. 2kp+2;
        // an odd last row runs the scalar sequence with packed weight products.
        // Window rows come into registers one row at a time.  BARREL: four aligned LDS.128 per row (a quarter-warp's eight 16-byte
        // reads never collide with the window pitch used here) and the lane's alignment shift sh = 0..3 applied in registers by two
        // select stages (by 2, by 1); otherwise SIDE+1 scalar reads at column sh + c, which collide 2.6-fold when the shifts of the
        // 32 lanes are unrelated (noise coordinates) and not at all when consecutive queries have consecutive anchors (smooth flows).
        // Measured at cfg2 with noise coordinates: BARREL cuts the shared-memory wavefronts of a launch from 2.16 M to 1.63 M (conflicts 1.03 M
        // -> 0.38 M) and is SLOWER (17.2 vs 16.5 us): the launch is bound by the dependent phases of its resident CTAs, not by the LSU.
        // It stays behind CF_LOOKUP_BARREL=1.
        constexpr int NP = JPW / 2, NC = SIDE + 1;
        constexpr bool ODD = (JPW & 1) != 0;
        const float* const rows = sq + j0 * PC;
        const bool sh1 = (sh & 1) != 0, sh2 = (sh & 2) != 0;
        auto load_row = [&](int k, float* v) {           // v[c] = window row j0 + k, column sh + c
            if (j0 + k > SIDE) {
#pragma unroll
                for (int c = 0; c < NC; ++c) v[c] = 0.0f;
                return;
            }
            if (BARREL) {
                constexpr int NCH = (NC + 3 + 3) / 4;    // chunks covering columns 0 .. NC + 2
                float wv[NCH * 4], t[NC + 1];
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch) {
                    const float4 q4 = *reinterpret_cast<const float4*>(rows + k * PC + 4 * ch);
                    wv[4 * ch] = q4.x; wv[4 * ch + 1] = q4.y; wv[4 * ch + 2] = q4.z; wv[4 * ch + 3] = q4.w;
                }
#pragma unroll
                for (int c = 0; c <= NC; ++c) t[c] = sh2 ? wv[c + 2] : wv[c];
#pragma unroll
                for (int c = 0; c < NC; ++c) v[c] = sh1 ? t[c + 1] : t[c];
            } else {
#pragma unroll
                for (int c = 0; c < NC; ++c) v[c] = rows[k * PC + sh + c];
            }
        };
        float ra[NC], rb[NC], rc[NC];
        load_row(0, ra);
#pragma unroll
        for (int kp = 0; kp < NP; ++kp) {
            const int k = 2 * kp;
            load_row(k + 1, rb);
            load_row(k + 2, rc);
            const float2 yl2 = make_float2(ywl[k], ywl[k + 1]), yh2 = make_float2(ywh[k], ywh[k + 1]);
            if (j0 + k < SIDE) {
#pragma unroll
                for (int i = 0; i < SIDE; ++i) {
                    const float2 xll = make_float2(xwl[i], xwl[i]), xhh = make_float2(xwh[i], xwh[i]);
                    float2 acc = __fmul2_rn(make_float2(ra[i], rb[i]), __fmul2_rn(xll, yl2));
                    acc = __ffma2_rn(make_float2(ra[i + 1], rb[i + 1]), __fmul2_rn(xhh, yl2), acc);
                    acc = __ffma2_rn(make_float2(rb[i], rc[i]), __fmul2_rn(xll, yh2), acc);
                    acc = __ffma2_rn(make_float2(rb[i + 1], rc[i + 1]), __fmul2_rn(xhh, yh2), acc);
                    __stcs(reinterpret_cast<float*>(ok[k]), acc.x);
                    ok[k] += ostep;
                    if (j0 + k + 1 < SIDE) { __stcs(reinterpret_cast<float*>(ok[k + 1]), acc.y); ok[k + 1] += ostep; }
                }
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) ra[c] = rc[c];
        }
        if (ODD) {
            constexpr int k = JPW - 1;
            load_row(k + 1, rb);
            if (j0 + k < SIDE) {
                const float2 yl1 = make_float2(ywl[k], ywl[k]), yh1 = make_float2(ywh[k], ywh[k]);
#pragma unroll
                for (int i = 0; i < SIDE; ++i) {
                    const float2 x = make_float2(xwl[i], xwh[i]);         // (w_lo, w_hi) of x tap i
                    const float2 wl2 = __fmul2_rn(x, yl1), wh2 = __fmul2_rn(x, yh1);
                    float acc = __fmul_rn(ra[i], wl2.x);
                    acc = fmaf(ra[i + 1], wl2.y, acc);
                    acc = fmaf(rb[i], wh2.x, acc);
                    acc = fmaf(rb[i + 1], wh2.y, acc);
                    __stcs(reinterpret_cast<float*>(ok[k]), acc);
                    ok[k] += ostep;
                }
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < SIDE; ++i) {
            const float xl = xwl[i], xh = xwh[i];
#pragma unroll
            for (int k = 0; k < JPW; ++k) {
                if (j0 + k < SIDE) {
                    const float* t = sq + yoff[k] + xo[i];
                    float acc = __fmul_rn(t[0], __fmul_rn(xl, ywl[k]));
                    acc = fmaf(t[1], __fmul_rn(xh, ywl[k]), acc);
                    acc = fmaf(t[PC], __fmul_rn(xl, ywh[k]), acc);
                    acc = fmaf(t[PC + 1], __fmul_rn(xh, ywh[k]), acc);
                    __stcs(reinterpret_cast<float*>(ok[k]), acc);
                    ok[k] += ostep;
                }
            }
        }
    }
    CF_TS(5, 0.0f);
}

// Variant: ONE warp per CTA owns 32 queries x one level and evaluates all (2r+1)^2 outputs of its queries.  Nothing is
// shared between warps and every per-query quantity (anchor, classification, x taps, staging setup) is computed exactly
// once instead of once per warp of a 3-warp CTA; the output rows are processed in chunks of CH = 3 with a rolled loop
// (register window (CH+1) x (2r+2), compact code).  Fewer instructions, fewer resident warps (one window buffer each).
template <int R>
__global__ void __launch_bounds__(32) lookup_warp_kernel(const __grid_constant__ LookupParams P) {
    using G = Geo<R>;
    constexpr int SIDE = G::SIDE, PR = G::PR, PC = G::PC, K = SIDE * SIDE, CH = 3;
    extern __shared__ __align__(16) float win[];     // windows: win_base(q) + row * PC + column

    const int lane = threadIdx.x;
    const unsigned task = P.lshift >= 0 ? (blockIdx.x >> P.lshift) : blockIdx.x / (unsigned)P.L;
    const int l = (int)(blockIdx.x - task * (unsigned)P.L);
    const unsigned n0 = task * kQT;             // CTA-uniform
    const bool valid = n0 + lane < (unsigned)P.N;
    const int hl = P.lh[l], wl = P.lw[l];
    const bool vec = (wl & 3) == 0;
    const unsigned HW = (unsigned)P.HW;

    unsigned b = n0 / HW;
    unsigned p = n0 - b * HW + lane;
    while (p >= HW) { p -= HW; ++b; }               // one wrap at most when H*W >= 32; tiny maps (L <= 2) wrap more
    float cx = 0.0f, cy = 0.0f;
    if (valid) {
        const float inv = __int_as_float((127 - l) << 23);      // 2^-l: the reference's division by 2^l, exactly
        const float* cp = P.coords + (size_t)b * 2 * HW + p;
        cx = __fmul_rn(__ldg(cp), inv);
        cy = __fmul_rn(__ldg(cp + HW), inv);
    }
    const bool dead = !valid || !(fabsf(cx) < kDeadCoord) || !(fabsf(cy) < kDeadCoord);
    if (dead) cx = cy = 0.0f;                       // finite taps; the outputs are forced to zero below
    bool nom_x, nom_y, lo_x, lo_y;
    classify<R>(cx, wl, nom_x, lo_x);
    classify<R>(cy, hl, nom_y, lo_y);
    const int ax = dead ? 0 : __float2int_rd(cx) - R - ((!nom_x && lo_x) ? 1 : 0);
    const int ay = dead ? 0 : __float2int_rd(cy) - R - ((!nom_y && lo_y) ? 1 : 0);
    const int sh = vec ? (ax & 3) : 0;
    const bool fast = __all_sync(0xffffffffu, dead || (nom_x && nom_y));
    {
        const float* __restrict__ img0 = P.level[l] + (size_t)n0 * hl * wl;
        const unsigned flags = (dead ? kDead : 0u) | (nom_x ? kNomX : 0u) | (nom_y ? kNomY : 0u) | (lo_x ? kLoX : 0u) |
                               (lo_y ? kLoY : 0u);
        const uint32_t win_s = (uint32_t)__cvta_generic_to_shared(win);
        if ((P.tiled >> l) & 1u) stage_windows<R, 1, true, true>(win_s, img0, hl, wl, 0, lane, ax, ay, flags, P.l2_frac);
        else if (vec) stage_windows<R, 1, true>(win_s, img0, hl, wl, 0, lane, ax, ay, flags, P.l2_frac);
        else stage_windows<R, 1, false>(win_s, img0, hl, wl, 0, lane, ax, ay, flags, P.l2_frac);
    }
    // x taps (registers) while the copies are in flight
    float xl[SIDE], xh[SIDE];
    int xo[SIDE];
    {
        const float s1x = (float)(wl - 1), rx = P.rcp_w[l];
#pragma unroll
        for (int i = 0; i < SIDE; ++i) {
            int i0;
            axis_tap(cx, i - R, s1x, rx, i0, xl[i], xh[i]);
            xo[i] = min(max(i0 - ax, 0), PR - 2) + sh;
        }
    }
    const float s1y = (float)(hl - 1), ry = P.rcp_h[l];
    cp_async_wait_all();
    __syncwarp();
    if (!valid) return;

    const float* sq = win + win_base<R>(lane);
    // 32-bit element offsets from a per-lane base (L * K * HW < 2^30 is checked on the host)
    float* const ob = P.out + ((size_t)b * P.L * K + (size_t)l * K) * HW + p;
    const float* rp0 = sq + sh;
#pragma unroll 1
    for (int jc = 0; jc < SIDE; jc += CH) {
        int yo[CH];
        float yl[CH], yh[CH];
#pragma unroll
        for (int k = 0; k < CH; ++k) {
            int j0 = 0;
            yl[k] = yh[k] = 0.0f;
            if (jc + k < SIDE) axis_tap(cy, jc + k - R, s1y, ry, j0, yl[k], yh[k]);
            yo[k] = min(max(j0 - ay, 0), PR - 2) * PC;
        }
        if (fast) {
            // nominal taps: x tap i reads stored columns sh+i, +1; output row j reads window rows j, +1
            float r[CH + 1][SIDE + 1];
            const float* rp = rp0 + jc * PC;
#pragma unroll
            for (int k = 0; k <= CH; ++k)
#pragma unroll
                for (int c = 0; c <= SIDE; ++c) r[k][c] = (!dead && jc + k <= SIDE) ? rp[k * PC + c] : 0.0f;
#pragma unroll
            for (int k = 0; k < CH; ++k) {
                if (jc + k < SIDE) {
                    unsigned off = (unsigned)(jc + k) * HW;
#pragma unroll
                    for (int i = 0; i < SIDE; ++i, off += SIDE * HW) {
                        float acc = __fmul_rn(r[k][i], __fmul_rn(xl[i], yl[k]));
                        acc = fmaf(r[k][i + 1], __fmul_rn(xh[i], yl[k]), acc);
                        acc = fmaf(r[k + 1][i], __fmul_rn(xl[i], yh[k]), acc);
                        acc = fmaf(r[k + 1][i + 1], __fmul_rn(xh[i], yh[k]), acc);
                        __stcs(ob + off, acc);
                    }
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < CH; ++k) {
                if (jc + k < SIDE) {
                    unsigned off = (unsigned)(jc + k) * HW;
#pragma unroll
                    for (int i = 0; i < SIDE; ++i, off += SIDE * HW) {
                        const float* t = sq + yo[k] + xo[i];
                        float acc = __fmul_rn(t[0], __fmul_rn(xl[i], yl[k]));
                        acc = fmaf(t[1], __fmul_rn(xh[i], yl[k]), acc);
                        acc = fmaf(t[PC], __fmul_rn(xl[i], yh[k]), acc);
                        acc = fmaf(t[PC + 1], __fmul_rn(xh[i], yh[k]), acc);
                        __stcs(ob + off, dead ? 0.0f : acc);
                    }
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
    const float rx = P.rcp_w[l], ry = P.rcp_h[l];
    const bool far = !(fabsf(cx) < kDeadCoord) || !(fabsf(cy) < kDeadCoord);   // outside div_small's checked range
    const size_t base = (((size_t)l * P.N + n) * 2) * G::SIDE;
#pragma unroll
    for (int i = 0; i < G::SIDE; ++i) {
        const float px = __fadd_rn(cx, (float)(i - R)), py = __fadd_rn(cy, (float)(i - R));
        const float ux = far ? lookup_coord_ieee(px, s1x) : lookup_coord(px, s1x, rx);
        const float uy = far ? lookup_coord_ieee(py, s1y) : lookup_coord(py, s1y, ry);
        P.taps[base + i] = ux;
        P.taps[base + G::SIDE + i] = uy;
        P.idx[base + i] = __float2int_rd(ux);
        P.idx[base + G::SIDE + i] = __float2int_rd(uy);
    }
}

template <int R, int NW>
int launch_lookup_rw(const LookupParams& P, long ngroups, cudaStream_t s) {
    static const bool pdl = [] { const char* e = getenv("CF_PDL"); return e ? atoi(e) != 0 : true; }();   // tuning knob
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(ngroups * P.L));
    cfg.blockDim = dim3(32 * NW);
    cfg.dynamicSmemBytes = Geo<R>::SMEM;
    cfg.stream = s;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = pdl ? 1 : 0;
    static const bool barrel = [] { const char* e = getenv("CF_LOOKUP_BARREL"); return e ? atoi(e) != 0 : false; }();   // A/B knob: measured slower, see the fast path
    if (barrel) CF_CHECK_CUDA(cudaLaunchKernelEx(&cfg, lookup_kernel<R, NW, true>, P));
    else CF_CHECK_CUDA(cudaLaunchKernelEx(&cfg, lookup_kernel<R, NW, false>, P));
    return CF_OK;
}

template <int R>
int launch_lookup_r(const LookupParams& P, cudaStream_t s) {
    const long ngroups = (P.N + kQT - 1) / kQT;
    if (P.N > 0x7fffffffL || ngroups * P.L > 0x7fffffffL) return fail(CF_ERR_UNSUPPORTED, "lookup grid too large (%ld queries)", P.N);
    static const int nw = [] { const char* e = getenv("CF_LOOKUP_NW"); return e ? atoi(e) : 3; }();   // tuning knob
    if ((long)P.L * Geo<R>::SIDE * Geo<R>::SIDE * P.HW < (1L << 30) && nw == 1) {
        constexpr size_t smem = (size_t)Geo<R>::WIN * sizeof(float);
        lookup_warp_kernel<R><<<(unsigned)(ngroups * P.L), 32, smem, s>>>(P);
        CF_CHECK_LAUNCH("lookup_warp_kernel");
        return CF_OK;
    }
    switch (nw) {
        case 2: return launch_lookup_rw<R, 2>(P, ngroups, s);
        case 4: return launch_lookup_rw<R, 4>(P, ngroups, s);
        default: return launch_lookup_rw<R, 3>(P, ngroups, s);
    }
}

template <int R>
int launch_taps_r(const LookupParams& P, cudaStream_t s) {
    dim3 grid((unsigned)((P.N + 127) / 128), (unsigned)P.L);
    lookup_taps_kernel<R><<<grid, 128, 0, s>>>(P);
    CF_CHECK_LAUNCH("lookup_taps_kernel");
    return CF_OK;
}

}  // namespace
#ifdef CF_BRINGUP
void lookup_set_timestamps(unsigned long long* p) { cudaMemcpyToSymbol(g_lookup_ts, &p, sizeof(p)); }
#endif
namespace {

LookupParams make_params(const float* pyramid, const LevelDims& d, const float* coords, int B, int H, int W, int L) {
    LookupParams P{};
    for (int l = 0; l < L; ++l) {
        P.level[l] = pyramid ? reinterpret_cast<const float*>(reinterpret_cast<const char*>(pyramid) + d.off[l]) : nullptr;
        P.lh[l] = d.h[l];
        P.lw[l] = d.w[l];
        P.rcp_h[l] = 1.0f / (float)(d.h[l] - 1);
        P.rcp_w[l] = 1.0f / (float)(d.w[l] - 1);
    }
    P.coords = coords;
    P.N = (long)B * H * W;
    P.HW = H * W;
    P.L = L;
    P.hw_magic = (unsigned)(((1ull << 32) + (unsigned long long)(H * W) - 1) / (unsigned long long)(H * W));
    P.lshift = -1;
    for (int k = 0; k < 4; ++k)
        if ((1 << k) == L) P.lshift = k;
    return P;
}

}  // namespace

int launch_lookup(const float* pyramid, const LevelDims& d, const float* coords, int B, int H, int W, int L, int r,
                  float* out, unsigned tiled_mask, cudaStream_t s) {
    LookupParams P = make_params(pyramid, d, coords, B, H, W, L);
    P.out = out;
    for (int l = 0; l < L; ++l)
        if (((tiled_mask >> l) & 1u) && ((d.h[l] | d.w[l]) & 3))
            return fail(CF_ERR_BAD_ARG, "lookup: level %d (%dx%d) cannot be 4x4-tiled", l, d.h[l], d.w[l]);
    P.tiled = tiled_mask;
    static const float frac = [] { const char* e = getenv("CF_L2_FRAC"); return e ? (float)atof(e) : 1.0f; }();   // tuning knob
    P.l2_frac = frac;

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
