// Local normalised cross-correlation (the registration loss taken on the warped frame): replaces `NCC.forward` of
// /root/reference/nnunet/lib/loss.py:15-76 (five zero-padded 9x9 box sums of I, J, I*I, J*J, I*J through conv2d with a
// ones kernel, then cc = clamp(cross^2 / (I_var * J_var + eps), 0.001, 0.999)) and its autograd.
//
// One skeleton, three instantiations (box5_walk<Load, Epi>).  A warp owns a strip of 128 input columns (a lane = 4
// consecutive columns, one 128-bit load per array and row) and walks down a band of rows:
//   vertical   : running 9-row sums of the five quantities in registers (add the entering row, subtract the leaving one,
//                which is re-read through L1; both rows are fetched one iteration ahead of their use),
//   horizontal : a lane's four 9-sums need its own 4 columns, the total of the next lane and three prefixes + the total of
//                the lane after that: 5 shuffles per quantity per 4 pixels (a lane-per-pixel log-step chain costs 16),
//   epilogue   : per-pixel arithmetic on the five box sums, 128-bit stores.  120 of 128 columns produce outputs.
// Nothing is staged in shared memory and no intermediate box sum ever reaches HBM.
//   forward : Load = (I-cI, J-cJ and their products), Epi = cc -> optional map of 1 - cc + per-task partial sum of cc
//   bwd A   : same Load, Epi = the five coefficient maps a, b, c, m1, m2 of d(loss)/d(box sums) (workspace [5,N,H,W])
//   bwd B   : Load = the five maps (zero outside the image: the adjoint of a zero-padded box filter is the same box
//             filter), Epi = gI = J*A + 2*I*B - M1, gJ = I*A + 2*J*C - M2
// Pivot: the variance terms cancel catastrophically in fp32 (I2_sum ~ 81*mean^2 against I_var ~ 81*var).  cross, I_var and
// J_var are invariant under I -> I - cI, J -> J - cJ when the zero padding is shifted too, so every task subtracts the
// pixel at its centre (padding becomes -c) and evaluates the algebraically identical cross = IJ_sum - I_sum*J_sum/n,
// I_var = I2_sum - I_sum^2/n on the shifted sums: closer to the float64 value than the reference's own fp32 evaluation
// (tests bound both by their distance from float64).
#include <initializer_list>

#include "common.cuh"

namespace cf {
namespace {

constexpr int kWin = 9, kHalo = 4;
constexpr int kInCols = 128, kOutCols = kInCols - 2 * kHalo;   // 120 output columns per warp strip (lanes 0..29)
#ifndef NCC_ROWS
#define NCC_ROWS 16
#endif
#ifndef NCC_MINB
#define NCC_MINB 1
#endif
#ifndef NCC_PF
#define NCC_PF 1   // rows of look-ahead in the walk (1 or 2)
#endif
constexpr int kRows = NCC_ROWS;                                // output rows per task (+9 priming rows of adds only)
constexpr int kWarps = 1;   // one warp per CTA: the task is a function of blockIdx alone, so the compiler sees the row loop as
                            // warp-uniform and emits plain SHFLs (with 4 warps per CTA every shuffle carried a divergence check)

struct Task {
    int n, y0, x0, rows;   // image, first output row / column, output rows of this task
};

__device__ __forceinline__ Task decode_task(long t, int H, int W) {
    const int strips = (W + kOutCols - 1) / kOutCols;
    const int bands = (H + kRows - 1) / kRows;
    Task k;
    const int s = (int)(t % strips);
    const long u = t / strips;
    const int b = (int)(u % bands);
    k.n = (int)(u / bands);
    k.x0 = s * kOutCols;
    k.y0 = b * kRows;
    k.rows = min(kRows, H - k.y0);
    return k;
}

template <int NQ>
struct QN {
    float v[NQ][4];   // quantity, column
};
using Q5 = QN<5>;

// sm_100a packed fp32: one instruction for two IEEE round-to-nearest operations on an aligned register pair
// (FADD2 / FMUL2 / FFMA2 in SASS).  The kernels are issue-bound, so halving the instruction count of the element-wise
// stretches (pivot subtraction, vertical accumulation, per-pixel statistics) is what pays.
struct F2 {
    unsigned long long u;
};
__device__ __forceinline__ F2 pk(float lo, float hi) {
    F2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.u) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk(F2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.u)); }
__device__ __forceinline__ F2 add2(F2 a, F2 b) {
    F2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u));
    return r;
}
__device__ __forceinline__ F2 sub2(F2 a, F2 b) {
    F2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u));
    return r;
}
__device__ __forceinline__ F2 mul2(F2 a, F2 b) {
    F2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u));
    return r;
}
__device__ __forceinline__ F2 fma2(F2 a, F2 b, F2 c) {
    F2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.u) : "l"(a.u), "l"(b.u), "l"(c.u));
    return r;
}

// 4 consecutive floats of row y starting at column x (x % 4 == 0 when vec), `pad` outside the image
// Branch-free: the address is clamped into the image and the value selected afterwards, so the warp stays converged
// around the shuffles and the loads can be hoisted a whole iteration ahead.
template <bool VEC>
__device__ __forceinline__ float4 load4(const float* img, int y, int x, int H, int W, float pad) {
    const bool row_in = (unsigned)y < (unsigned)H;
    const float* row = img + (long)min(max(y, 0), H - 1) * W;
    float4 r;
    if (VEC) {
        r = __ldg(reinterpret_cast<const float4*>(row + min(max(x, 0), W - 4)));
        if (!(row_in && (unsigned)x < (unsigned)W)) r = make_float4(pad, pad, pad, pad);
    } else {
        const float e0 = __ldg(row + min(max(x + 0, 0), W - 1)), e1 = __ldg(row + min(max(x + 1, 0), W - 1));
        const float e2 = __ldg(row + min(max(x + 2, 0), W - 1)), e3 = __ldg(row + min(max(x + 3, 0), W - 1));
        r.x = (row_in && (unsigned)(x + 0) < (unsigned)W) ? e0 : pad;
        r.y = (row_in && (unsigned)(x + 1) < (unsigned)W) ? e1 : pad;
        r.z = (row_in && (unsigned)(x + 2) < (unsigned)W) ? e2 : pad;
        r.w = (row_in && (unsigned)(x + 3) < (unsigned)W) ? e3 : pad;
    }
    return r;
}
template <bool VEC>
__device__ __forceinline__ void store4(float* img, int y, int x, int W, const float (&o)[4]) {
    float* row = img + (long)y * W;
    if (VEC) {
        *reinterpret_cast<float4*>(row + x) = make_float4(o[0], o[1], o[2], o[3]);
    } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (x + c < W) row[x + c] = o[c];
    }
}

// horizontal 9-sums of a lane's four windows (window c starts at the lane's column c); valid on lanes 0..29
__device__ __forceinline__ void hsum9x4(const float (&v)[4], float (&o)[4]) {
    const float p2 = v[0] + v[1], s23 = v[2] + v[3];
    const float p3 = p2 + v[2], s123 = v[1] + s23, t = p2 + s23;
    const float T1 = __shfl_down_sync(0xffffffffu, t, 1);
    const float T2 = __shfl_down_sync(0xffffffffu, t, 2);
    const float P1 = __shfl_down_sync(0xffffffffu, v[0], 2);
    const float P2 = __shfl_down_sync(0xffffffffu, p2, 2);
    const float P3 = __shfl_down_sync(0xffffffffu, p3, 2);
    o[0] = (t + T1) + P1;
    o[1] = (s123 + T1) + P2;
    o[2] = (s23 + T1) + P3;
    o[3] = (v[3] + T1) + T2;
}

// Load: Raw fetch(y) (this lane's 4 columns of row y, any y) and accumulate(V, raw, sign).
// Epi(y, x, sums) for the 4 output pixels (y, x..x+3) of a lane; x < W is guaranteed, x + c < W is the functor's business.
template <class Load, class Epi>
__device__ __forceinline__ void box5_walk(const Task& k, int W, const Load& load, Epi& epi) {
    const int lane = threadIdx.x & 31;
    const int xout = k.x0 + 4 * lane;             // first output column of this lane (its windows start at xout - 4)
    const bool live = lane < kOutCols / 4 && xout < W;
    constexpr int NQ = Load::kNQ;
    QN<NQ> V;
#pragma unroll
    for (int q = 0; q < NQ; ++q)
#pragma unroll
        for (int c = 0; c < 4; ++c) V.v[q][c] = 0.f;
    // rows y0-5 .. y0+3: the window of output row y0 - 1 (row y0-5 leaves in the first iteration)
#pragma unroll 3
    for (int j = -kHalo - 1; j < kHalo; ++j) load.accumulate(V, load.fetch(k.y0 + j), 1.0f);
#if NCC_PF == 2
    // rows fetched two iterations ahead of their use (Load::kDeep loaders only: the pair loader's rows are 8 registers each)
    typename Load::Raw enter = load.fetch(k.y0 + kHalo), leave = load.fetch(k.y0 - kHalo - 1);
    typename Load::Raw enter2 = load.fetch(k.y0 + kHalo + 1), leave2 = load.fetch(k.y0 - kHalo);
    for (int y = k.y0; y < k.y0 + k.rows; ++y) {
        const typename Load::Raw e = enter, l = leave;
        enter = enter2;
        leave = leave2;
        enter2 = load.fetch(y + 2 + kHalo);
        leave2 = load.fetch(y + 1 - kHalo);
#else
    typename Load::Raw enter = load.fetch(k.y0 + kHalo), leave = load.fetch(k.y0 - kHalo - 1);
    for (int y = k.y0; y < k.y0 + k.rows; ++y) {
        const typename Load::Raw e = enter, l = leave;
        enter = load.fetch(y + 1 + kHalo);        // next iteration's rows: in flight while this row is computed
        leave = load.fetch(y - kHalo);
#endif
        load.accumulate(V, e, 1.0f);
        load.accumulate(V, l, -1.0f);
        QN<NQ> S;
#pragma unroll
        for (int q = 0; q < NQ; ++q) hsum9x4(V.v[q], S.v[q]);
        if (live) epi(y, xout, S);
    }
}

template <bool VEC>
struct PairLoad {   // (I - cI, J - cJ, squares, product); padding = (0 - c)
    static constexpr int kNQ = 5;
    struct Raw {
        float4 a, b;
    };
    const float *I, *J;
    int H, W, x;     // x = this lane's first input column
    float cI, cJ;
    __device__ __forceinline__ Raw fetch(int y) const {
        return Raw{load4<VEC>(I, y, x, H, W, 0.f), load4<VEC>(J, y, x, H, W, 0.f)};
    }
    // sign = +1: the row enters the window; -1: it leaves (products are negated through the (c - x) form, no extra multiply)
    __device__ __forceinline__ void accumulate(Q5& V, const Raw& r, float sign) const {
        const F2 c_i = pk(cI, cI), c_j = pk(cJ, cJ);
        const F2 ra[2] = {pk(r.a.x, r.a.y), pk(r.a.z, r.a.w)}, rb[2] = {pk(r.b.x, r.b.y), pk(r.b.z, r.b.w)};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const F2 a = sub2(ra[h], c_i), b = sub2(rb[h], c_j);
            const F2 sa = sign > 0.f ? a : sub2(c_i, ra[h]), sb = sign > 0.f ? b : sub2(c_j, rb[h]);
            F2 v0 = pk(V.v[0][2 * h], V.v[0][2 * h + 1]), v1 = pk(V.v[1][2 * h], V.v[1][2 * h + 1]);
            F2 v2 = pk(V.v[2][2 * h], V.v[2][2 * h + 1]), v3 = pk(V.v[3][2 * h], V.v[3][2 * h + 1]);
            F2 v4 = pk(V.v[4][2 * h], V.v[4][2 * h + 1]);
            v0 = add2(v0, sa);
            v1 = add2(v1, sb);
            v2 = fma2(sa, a, v2);
            v3 = fma2(sb, b, v3);
            v4 = fma2(sa, b, v4);
            upk(v0, V.v[0][2 * h], V.v[0][2 * h + 1]);
            upk(v1, V.v[1][2 * h], V.v[1][2 * h + 1]);
            upk(v2, V.v[2][2 * h], V.v[2][2 * h + 1]);
            upk(v3, V.v[3][2 * h], V.v[3][2 * h + 1]);
            upk(v4, V.v[4][2 * h], V.v[4][2 * h + 1]);
        }
    }
};

struct Stats {   // loss.py:59-66 on shifted sums: cross = IJ - Is*Js/n, I_var = I2 - Is^2/n (same algebra, fewer roundings)
    float nuI, nuJ, cross, Iv, Jv, D, cc0;   // nuI = -I_sum/n (the negated mean: feeds the FFMA2s without a negation)
};
// statistics of the column pair (2h, 2h+1) with packed arithmetic
__device__ __forceinline__ void stats_pair(const Q5& s, int h, float eps, Stats (&t)[2]) {
    const float ninv = -1.0f / (float)(kWin * kWin);
    const F2 S0 = pk(s.v[0][2 * h], s.v[0][2 * h + 1]), S1 = pk(s.v[1][2 * h], s.v[1][2 * h + 1]);
    const F2 S2 = pk(s.v[2][2 * h], s.v[2][2 * h + 1]), S3 = pk(s.v[3][2 * h], s.v[3][2 * h + 1]);
    const F2 S4 = pk(s.v[4][2 * h], s.v[4][2 * h + 1]);
    const F2 nuI = mul2(S0, pk(ninv, ninv)), nuJ = mul2(S1, pk(ninv, ninv));
    const F2 cross = fma2(nuI, S1, S4), Iv = fma2(nuI, S0, S2), Jv = fma2(nuJ, S1, S3);
    const F2 D = fma2(Iv, Jv, pk(eps, eps)), num = mul2(cross, cross);
    float n0, n1;
    upk(nuI, t[0].nuI, t[1].nuI);
    upk(nuJ, t[0].nuJ, t[1].nuJ);
    upk(cross, t[0].cross, t[1].cross);
    upk(Iv, t[0].Iv, t[1].Iv);
    upk(Jv, t[0].Jv, t[1].Jv);
    upk(D, t[0].D, t[1].D);
    upk(num, n0, n1);
    t[0].cc0 = __fdividef(n0, t[0].D);
    t[1].cc0 = __fdividef(n1, t[1].D);
}

// Pivot = the task's centre pixel scaled by the in-image fraction of the task's input footprint: ~ the mean of the windows
// INCLUDING their zero padding (an interior task subtracts the pixel itself, a 3x3 image whose windows are 8/9 padding
// subtracts almost nothing).  Any pivot gives the same result in exact arithmetic; this one keeps the shifted sums small.
__device__ __forceinline__ void task_pivot(const float* I, const float* J, const Task& k, int H, int W, float* cI, float* cJ) {
    const int cols = min(kOutCols, W - k.x0);
    const long p = (long)min(k.y0 + k.rows / 2, H - 1) * W + min(k.x0 + cols / 2, W - 1);
    const int in_rows = min(H, k.y0 + k.rows + kHalo) - max(0, k.y0 - kHalo);
    const int in_cols = min(W, k.x0 + cols + kHalo) - max(0, k.x0 - kHalo);
    const float f = (float)(in_rows * in_cols) / (float)((k.rows + 2 * kHalo) * (cols + 2 * kHalo));
    *cI = f * __ldg(I + p);
    *cJ = f * __ldg(J + p);
}

template <bool VEC>
struct FwdEpi {
    float* map;   // image base or nullptr
    int W;
    float eps, acc;
    __device__ __forceinline__ void operator()(int y, int x, const Q5& s) {
        float o[4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            Stats t[2];
            stats_pair(s, h, eps, t);
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int c = 2 * h + e;
                const float cc = fminf(fmaxf(t[e].cc0, 0.001f), 0.999f);
                o[c] = 1.0f - cc;                  // what reduction=None returns (loss.py:76)
                if (VEC || x + c < W) acc += cc;
            }
        }
        if (map) store4<VEC>(map, y, x, W, o);
    }
};

template <bool VEC>
__global__ void __launch_bounds__(kWarps * 32, NCC_MINB) ncc_fwd_kernel(const float* __restrict__ I, const float* __restrict__ J, int H, int W,
                                                             long n_tasks, float eps, float* __restrict__ loss_map,
                                                             float* __restrict__ partial) {
    const long t = blockIdx.x;
    if (t >= n_tasks) return;
    const Task k = decode_task(t, H, W);
    const long img = (long)k.n * H * W;
    PairLoad<VEC> load{I + img, J + img, H, W, k.x0 - kHalo + 4 * (int)(threadIdx.x & 31), 0.f, 0.f};
    task_pivot(load.I, load.J, k, H, W, &load.cI, &load.cJ);
    FwdEpi<VEC> epi{loss_map ? loss_map + img : nullptr, W, eps, 0.f};
    box5_walk(k, W, load, epi);
    float a = epi.acc;
#pragma unroll
    for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (partial && (threadIdx.x & 31) == 0) partial[t] = a;
}

// fixed-order sum of the per-task partials in double: the loss does not depend on scheduling
__global__ void __launch_bounds__(1024) ncc_reduce_kernel(const float* __restrict__ partial, long n, float* __restrict__ cc_sum) {
    __shared__ double sh[1024];
    double a = 0.0;
#pragma unroll 4
    for (long i = threadIdx.x; i < n; i += 1024) a += (double)partial[i];
    sh[threadIdx.x] = a;
    __syncthreads();
    for (int o = 512; o; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *cc_sum = (float)sh[0];
}

// MODE 0: both gradients (maps a, b, c, m1, m2); 1: grad_I only (a, b, m1); 2: grad_J only (a, c, m2) -- the usual case in
// training, where only the warped frame carries a gradient: 3 maps instead of 5 through HBM and through the second box filter.
template <int MODE>
struct BwdMaps {
    static constexpr int kNQ = MODE == 0 ? 5 : 3;
};

template <bool VEC, int MODE>
struct CoefEpi {   // d(loss)/d(box sums) folded into kNQ maps
    float* ws;     // [kNQ, N, H, W] base + image offset
    const float* gpix;   // per-pixel upstream (image base) or nullptr
    long plane;    // N*H*W
    int H, W;
    float eps, gscale, cI, cJ;
    __device__ __forceinline__ void operator()(int y, int x, const Q5& s) {
        constexpr int NQ = BwdMaps<MODE>::kNQ;
        float g[4] = {1.f, 1.f, 1.f, 1.f};
        if (gpix) {
            const float4 u = load4<VEC>(gpix, y, x, H, W, 0.f);
            g[0] = u.x; g[1] = u.y; g[2] = u.z; g[3] = u.w;
        }
        float m[NQ][4];
        Stats st[4];
        {
            Stats t2[2];
            stats_pair(s, 0, eps, t2);
            st[0] = t2[0]; st[1] = t2[1];
            stats_pair(s, 1, eps, t2);
            st[2] = t2[0]; st[3] = t2[1];
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const Stats t = st[c];
            float w = gscale * g[c];
            if (!(t.cc0 >= 0.001f && t.cc0 <= 0.999f)) w = 0.f;      // clamp passes gradient on the closed interval
            const float invD = __fdividef(1.0f, t.D);
            const float a = w * 2.0f * t.cross * invD;
            const float q = -w * t.cross * t.cross * invD * invD;
            const float b = q * t.Jv, cc = q * t.Iv;
            const float uI = cI - t.nuI, uJ = cJ - t.nuJ;             // absolute means: pass B multiplies absolute pixels
            m[0][c] = a;
            if (MODE == 0) {
                m[1][c] = b;
                m[2][c] = cc;
                m[3][c] = fmaf(a, uJ, 2.0f * b * uI);
                m[4][c] = fmaf(a, uI, 2.0f * cc * uJ);
            } else if (MODE == 1) {
                m[1][c] = b;
                m[2][c] = fmaf(a, uJ, 2.0f * b * uI);
            } else {
                m[1][c] = cc;
                m[2][c] = fmaf(a, uI, 2.0f * cc * uJ);
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q) store4<VEC>(ws + q * plane, y, x, W, m[q]);
    }
};

template <bool VEC, int MODE>
__global__ void __launch_bounds__(kWarps * 32, NCC_MINB) ncc_bwd_coef_kernel(const float* __restrict__ I, const float* __restrict__ J, int N, int H,
                                                                  int W, long n_tasks, float eps,
                                                                  const float* __restrict__ grad, int per_pixel, float mult,
                                                                  float* __restrict__ ws) {
    const long t = blockIdx.x;
    if (t >= n_tasks) return;
    const Task k = decode_task(t, H, W);
    const long img = (long)k.n * H * W;
    PairLoad<VEC> load{I + img, J + img, H, W, k.x0 - kHalo + 4 * (int)(threadIdx.x & 31), 0.f, 0.f};
    task_pivot(load.I, load.J, k, H, W, &load.cI, &load.cJ);
    CoefEpi<VEC, MODE> epi{ws + img, per_pixel ? grad + img : nullptr, (long)N * H * W, H, W, eps,
                per_pixel ? mult : mult * __ldg(grad), load.cI, load.cJ};
    box5_walk(k, W, load, epi);
}

template <bool VEC, int MODE>
struct MapLoad {
    static constexpr int kNQ = BwdMaps<MODE>::kNQ;
    struct Raw {
        float4 m[kNQ];
    };
    const float* ws;   // image base of map 0
    long plane;
    int H, W, x;
    __device__ __forceinline__ Raw fetch(int y) const {
        Raw r;
#pragma unroll
        for (int q = 0; q < kNQ; ++q) r.m[q] = load4<VEC>(ws + q * plane, y, x, H, W, 0.f);
        return r;
    }
    __device__ __forceinline__ void accumulate(QN<kNQ>& V, const Raw& r, float sign) const {
#pragma unroll
        for (int q = 0; q < kNQ; ++q) {
            const F2 lo = pk(r.m[q].x, r.m[q].y), hi = pk(r.m[q].z, r.m[q].w);
            const F2 v01 = pk(V.v[q][0], V.v[q][1]), v23 = pk(V.v[q][2], V.v[q][3]);
            upk(sign > 0.f ? add2(v01, lo) : sub2(v01, lo), V.v[q][0], V.v[q][1]);
            upk(sign > 0.f ? add2(v23, hi) : sub2(v23, hi), V.v[q][2], V.v[q][3]);
        }
    }
};

template <bool VEC, int MODE>
struct GradEpi {
    const float *I, *J;
    float *gI, *gJ;   // image bases (MODE 1: gJ unused, MODE 2: gI unused)
    int H, W;
    __device__ __forceinline__ void operator()(int y, int x, const QN<BwdMaps<MODE>::kNQ>& s) {
        const float4 i4 = load4<VEC>(I, y, x, H, W, 0.f), j4 = load4<VEC>(J, y, x, H, W, 0.f);
        const float i[4] = {i4.x, i4.y, i4.z, i4.w}, j[4] = {j4.x, j4.y, j4.z, j4.w};
        float oi[4], oj[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (MODE == 0) {
                oi[c] = fmaf(j[c], s.v[0][c], fmaf(2.0f * i[c], s.v[1][c], -s.v[3][c]));
                oj[c] = fmaf(i[c], s.v[0][c], fmaf(2.0f * j[c], s.v[2][c], -s.v[4][c]));
            } else if (MODE == 1) {
                oi[c] = fmaf(j[c], s.v[0][c], fmaf(2.0f * i[c], s.v[1][c], -s.v[2][c]));
            } else {
                oj[c] = fmaf(i[c], s.v[0][c], fmaf(2.0f * j[c], s.v[1][c], -s.v[2][c]));
            }
        }
        if (MODE != 2) store4<VEC>(gI, y, x, W, oi);
        if (MODE != 1) store4<VEC>(gJ, y, x, W, oj);
    }
};

template <bool VEC, int MODE>
__global__ void __launch_bounds__(kWarps * 32, NCC_MINB) ncc_bwd_gather_kernel(const float* __restrict__ I, const float* __restrict__ J, int N, int H,
                                                                    int W, long n_tasks, const float* __restrict__ ws,
                                                                    float* __restrict__ gI, float* __restrict__ gJ) {
    const long t = blockIdx.x;
    if (t >= n_tasks) return;
    const Task k = decode_task(t, H, W);
    const long img = (long)k.n * H * W;
    MapLoad<VEC, MODE> load{ws + img, (long)N * H * W, H, W, k.x0 - kHalo + 4 * (int)(threadIdx.x & 31)};
    GradEpi<VEC, MODE> epi{I + img, J + img, gI ? gI + img : nullptr, gJ ? gJ + img : nullptr, H, W};
    box5_walk(k, W, load, epi);
}

long task_count(int N, int H, int W) {
    return (long)N * ((H + kRows - 1) / kRows) * ((W + kOutCols - 1) / kOutCols);
}

// 128-bit loads / stores need rows that start on 16-byte boundaries
bool vec_ok(int W, std::initializer_list<const void*> ptrs) {
    if (W % 4) return false;
    for (const void* p : ptrs)
        if (p && !aligned16(p)) return false;
    return true;
}

}  // namespace

size_t ncc_workspace_bytes(int N, int H, int W, int backward) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    return backward ? (size_t)5 * N * H * W * sizeof(float) : round_up((size_t)task_count(N, H, W) * sizeof(float), 256);
}

int launch_ncc(const float* I, const float* J, int N, int H, int W, float eps, float* loss_map, float* cc_sum, float* partial,
               cudaStream_t s) {
    const long nt = task_count(N, H, W);
    const int vec = vec_ok(W, {I, J, loss_map});
    const dim3 grid(cdiv(nt, kWarps)), block(kWarps * 32);
    if (vec) ncc_fwd_kernel<true><<<grid, block, 0, s>>>(I, J, H, W, nt, eps, loss_map, cc_sum ? partial : nullptr);
    else ncc_fwd_kernel<false><<<grid, block, 0, s>>>(I, J, H, W, nt, eps, loss_map, cc_sum ? partial : nullptr);
    CF_CHECK_LAUNCH("ncc_fwd_kernel");
    if (cc_sum) {
        ncc_reduce_kernel<<<1, 1024, 0, s>>>(partial, nt, cc_sum);
        CF_CHECK_LAUNCH("ncc_reduce_kernel");
    }
    return CF_OK;
}

template <bool VEC, int MODE>
int launch_ncc_bwd_t(const float* I, const float* J, int N, int H, int W, float eps, const float* grad, int per_pixel, float mult,
                     float* gI, float* gJ, float* ws, cudaStream_t s) {
    const long nt = task_count(N, H, W);
    const dim3 grid(cdiv(nt, kWarps)), block(kWarps * 32);
    ncc_bwd_coef_kernel<VEC, MODE><<<grid, block, 0, s>>>(I, J, N, H, W, nt, eps, grad, per_pixel, mult, ws);
    CF_CHECK_LAUNCH("ncc_bwd_coef_kernel");
    ncc_bwd_gather_kernel<VEC, MODE><<<grid, block, 0, s>>>(I, J, N, H, W, nt, ws, gI, gJ);
    CF_CHECK_LAUNCH("ncc_bwd_gather_kernel");
    return CF_OK;
}

int launch_ncc_bwd(const float* I, const float* J, int N, int H, int W, float eps, const float* grad, int per_pixel, float mult,
                   float* gI, float* gJ, float* ws, cudaStream_t s) {
    const bool vec = vec_ok(W, {I, J, gI, gJ, ws, per_pixel ? grad : nullptr});
    const int mode = (gI && gJ) ? 0 : (gI ? 1 : 2);
#define CF_NCC_BWD(V, M) launch_ncc_bwd_t<V, M>(I, J, N, H, W, eps, grad, per_pixel, mult, gI, gJ, ws, s)
    if (vec) return mode == 0 ? CF_NCC_BWD(true, 0) : (mode == 1 ? CF_NCC_BWD(true, 1) : CF_NCC_BWD(true, 2));
    return mode == 0 ? CF_NCC_BWD(false, 0) : (mode == 1 ? CF_NCC_BWD(false, 1) : CF_NCC_BWD(false, 2));
#undef CF_NCC_BWD
}

}  // namespace cf
