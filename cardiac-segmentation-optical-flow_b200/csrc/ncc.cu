// Local normalised cross-correlation (the registration loss taken on the warped frame): replaces `NCC.forward` of
// /root/reference/nnunet/lib/loss.py:15-76 (five zero-padded 9x9 box sums of I, J, I*I, J*J, I*J through conv2d with a
// ones kernel, then cc = clamp(cross^2 / (I_var * J_var + eps), 0.001, 0.999)) and its autograd.
//
// One skeleton, three instantiations (box5_kernel<Load, Epi>): a warp owns a strip of 32 input columns and walks down the
// rows; a lane loads one pixel per row (coalesced 128 B per warp and array), the horizontal 9-sum of the five quantities is
// a log-step shuffle chain (v+1, +2, +4, +8: 4 shuffles per quantity, valid on lanes 0..23), the vertical 9-sum runs over a
// register ring of the last nine row sums (statically indexed: the row loop is unrolled by 9).  Nothing is staged in shared
// memory and no intermediate box sum ever reaches HBM.
//   forward : Load = (I-cI, J-cJ and their products), Epi = cc -> optional map of 1 - cc + per-task partial sum of cc
//   bwd A   : same Load, Epi = the five coefficient maps a, b, c, m1, m2 of d(loss)/d(box sums) (workspace [5,N,H,W])
//   bwd B   : Load = the five maps (zero outside the image: the adjoint of a zero-padded box filter is the same box
//             filter), Epi = gI = J*A + 2*I*B - M1, gJ = I*A + 2*J*C - M2
// Pivot: the variance terms cancel catastrophically in fp32 (I2_sum ~ 81*mean^2 against I_var ~ 81*var).  cross, I_var and
// J_var are invariant under I -> I - cI, J -> J - cJ when the zero padding is shifted too, so every task subtracts the
// pixel at its centre (padding becomes -c): the same formula, evaluated closer to the float64 value than the reference's
// own fp32 evaluation (tests bound both by their distance from float64).
#include "common.cuh"

namespace cf {
namespace {

constexpr int kWin = 9, kHalo = 4, kOutCols = 32 - 2 * kHalo;   // 24 output columns per warp strip
constexpr int kRows = 56;                                       // output rows per task (8 priming rows: 12.5 % overhead)
constexpr int kWarps = 4;

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

struct V5 {
    float v[5];
};

__device__ __forceinline__ V5 hsum9(const V5& a) {
    V5 r;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const float s2 = a.v[i] + __shfl_down_sync(0xffffffffu, a.v[i], 1);
        const float s4 = s2 + __shfl_down_sync(0xffffffffu, s2, 2);
        const float s8 = s4 + __shfl_down_sync(0xffffffffu, s4, 4);
        r.v[i] = s8 + __shfl_down_sync(0xffffffffu, a.v[i], 8);
    }
    return r;
}

// Load(y, x) -> five values of input pixel (y, x) (any y, x: the functor handles the padding);
// Epi(y, x, sums) for every in-image output pixel of the task.
template <class Load, class Epi>
__device__ __forceinline__ void box5_walk(const Task& k, int W, Load& load, Epi& epi) {
    const int lane = threadIdx.x & 31;
    const int xin = k.x0 - kHalo + lane;          // input column of this lane
    const int xout = k.x0 + lane;                 // output column (window start = xin, centre = xin + 4)
    const bool live = lane < kOutCols && xout < W;
    V5 ring[kWin];
#pragma unroll
    for (int j = 0; j < kWin - 1; ++j) ring[j] = hsum9(load(k.y0 - kHalo + j, xin));
    for (int y = 0; y < k.rows; y += kWin) {
#pragma unroll
        for (int j = 0; j < kWin; ++j) {
            if (y + j < k.rows) {                 // warp-uniform
                ring[(kWin - 1 + j) % kWin] = hsum9(load(k.y0 + y + j + kHalo, xin));
                V5 s;
#pragma unroll
                for (int i = 0; i < 5; ++i) {
                    float acc = ring[0].v[i];
#pragma unroll
                    for (int q = 1; q < kWin; ++q) acc += ring[q].v[i];
                    s.v[i] = acc;
                }
                if (live) epi(k.y0 + y + j, xout, s);
            }
        }
    }
}

struct PairLoad {   // (I - cI, J - cJ, squares, product); padding = (0 - c)
    const float *I, *J;
    int H, W;
    float cI, cJ;
    __device__ __forceinline__ V5 operator()(int y, int x) const {
        float a = -cI, b = -cJ;
        if ((unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W) {
            a = __ldg(I + (long)y * W + x) - cI;
            b = __ldg(J + (long)y * W + x) - cJ;
        }
        V5 r;
        r.v[0] = a; r.v[1] = b; r.v[2] = a * a; r.v[3] = b * b; r.v[4] = a * b;
        return r;
    }
};

struct Stats {   // loss.py:59-66 on (shifted) sums
    float uI, uJ, cross, Iv, Jv, D, cc0;
};
__device__ __forceinline__ Stats stats_of(const V5& s, float eps) {
    const float n = (float)(kWin * kWin);
    Stats t;
    t.uI = __fdiv_rn(s.v[0], n);
    t.uJ = __fdiv_rn(s.v[1], n);
    t.cross = s.v[4] - t.uJ * s.v[0] - t.uI * s.v[1] + t.uI * t.uJ * n;
    t.Iv = s.v[2] - 2.0f * t.uI * s.v[0] + t.uI * t.uI * n;
    t.Jv = s.v[3] - 2.0f * t.uJ * s.v[1] + t.uJ * t.uJ * n;
    t.D = t.Iv * t.Jv + eps;
    t.cc0 = __fdiv_rn(t.cross * t.cross, t.D);
    return t;
}

__device__ __forceinline__ void task_pivot(const float* I, const float* J, const Task& k, int H, int W, float* cI, float* cJ) {
    const long p = (long)min(k.y0 + k.rows / 2, H - 1) * W + min(k.x0 + kOutCols / 2, W - 1);
    *cI = __ldg(I + p);
    *cJ = __ldg(J + p);
}

struct FwdEpi {
    float* map;   // image base or nullptr
    int W;
    float eps, acc;
    __device__ __forceinline__ void operator()(int y, int x, const V5& s) {
        const float cc = fminf(fmaxf(stats_of(s, eps).cc0, 0.001f), 0.999f);
        if (map) map[(long)y * W + x] = 1.0f - cc;             // what reduction=None returns (loss.py:76)
        acc += cc;
    }
};

__global__ void __launch_bounds__(kWarps * 32) ncc_fwd_kernel(const float* __restrict__ I, const float* __restrict__ J, int H, int W,
                                                             long n_tasks, float eps, float* __restrict__ loss_map,
                                                             float* __restrict__ partial) {
    const long t = (long)blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (t >= n_tasks) return;
    const Task k = decode_task(t, H, W);
    const long img = (long)k.n * H * W;
    PairLoad load{I + img, J + img, H, W, 0.f, 0.f};
    task_pivot(load.I, load.J, k, H, W, &load.cI, &load.cJ);
    FwdEpi epi{loss_map ? loss_map + img : nullptr, W, eps, 0.f};
    box5_walk(k, W, load, epi);
    float a = epi.acc;
#pragma unroll
    for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (partial && (threadIdx.x & 31) == 0) partial[t] = a;
}

// fixed-order sum of the per-task partials in double: the loss does not depend on scheduling
__global__ void __launch_bounds__(256) ncc_reduce_kernel(const float* __restrict__ partial, long n, float* __restrict__ cc_sum) {
    __shared__ double sh[256];
    double a = 0.0;
    for (long i = threadIdx.x; i < n; i += 256) a += (double)partial[i];
    sh[threadIdx.x] = a;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *cc_sum = (float)sh[0];
}

struct CoefEpi {   // d(loss)/d(box sums) folded into five maps
    float* ws;     // [5, N, H, W] base + image offset
    const float* gpix;   // per-pixel upstream (image base) or nullptr
    long plane;    // N*H*W
    int W;
    float eps, gscale, cI, cJ;
    __device__ __forceinline__ void operator()(int y, int x, const V5& s) {
        const Stats t = stats_of(s, eps);
        const long p = (long)y * W + x;
        float w = gscale;
        if (gpix) w *= gpix[p];
        if (!(t.cc0 >= 0.001f && t.cc0 <= 0.999f)) w = 0.f;          // clamp passes gradient on the closed interval
        const float invD = __fdiv_rn(1.0f, t.D);
        const float a = w * 2.0f * t.cross * invD;
        const float q = -w * t.cross * t.cross * invD * invD;
        const float b = q * t.Jv, c = q * t.Iv;
        const float uI = t.uI + cI, uJ = t.uJ + cJ;                   // absolute means: pass B multiplies absolute pixels
        ws[p] = a;
        ws[plane + p] = b;
        ws[2 * plane + p] = c;
        ws[3 * plane + p] = a * uJ + 2.0f * b * uI;
        ws[4 * plane + p] = a * uI + 2.0f * c * uJ;
    }
};

__global__ void __launch_bounds__(kWarps * 32) ncc_bwd_coef_kernel(const float* __restrict__ I, const float* __restrict__ J, int N, int H,
                                                                  int W, long n_tasks, float eps, const float* __restrict__ grad,
                                                                  int per_pixel, float mult, float* __restrict__ ws) {
    const long t = (long)blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (t >= n_tasks) return;
    const Task k = decode_task(t, H, W);
    const long img = (long)k.n * H * W;
    PairLoad load{I + img, J + img, H, W, 0.f, 0.f};
    task_pivot(load.I, load.J, k, H, W, &load.cI, &load.cJ);
    CoefEpi epi{ws + img, per_pixel ? grad + img : nullptr, (long)N * H * W, W, eps,
                per_pixel ? mult : mult * __ldg(grad), load.cI, load.cJ};
    box5_walk(k, W, load, epi);
}

struct MapLoad {
    const float* ws;   // image base of map 0
    long plane;
    int H, W;
    __device__ __forceinline__ V5 operator()(int y, int x) const {
        V5 r;
        if ((unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W) {
            const long p = (long)y * W + x;
#pragma unroll
            for (int i = 0; i < 5; ++i) r.v[i] = __ldg(ws + i * plane + p);
        } else {
#pragma unroll
            for (int i = 0; i < 5; ++i) r.v[i] = 0.f;
        }
        return r;
    }
};

struct GradEpi {
    const float *I, *J;
    float *gI, *gJ;   // image bases or nullptr
    int W;
    __device__ __forceinline__ void operator()(int y, int x, const V5& s) {
        const long p = (long)y * W + x;
        const float i = __ldg(I + p), j = __ldg(J + p);
        if (gI) gI[p] = j * s.v[0] + 2.0f * i * s.v[1] - s.v[3];
        if (gJ) gJ[p] = i * s.v[0] + 2.0f * j * s.v[2] - s.v[4];
    }
};

__global__ void __launch_bounds__(kWarps * 32) ncc_bwd_gather_kernel(const float* __restrict__ I, const float* __restrict__ J, int N, int H,
                                                                    int W, long n_tasks, const float* __restrict__ ws,
                                                                    float* __restrict__ gI, float* __restrict__ gJ) {
    const long t = (long)blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (t >= n_tasks) return;
    const Task k = decode_task(t, H, W);
    const long img = (long)k.n * H * W;
    MapLoad load{ws + img, (long)N * H * W, H, W};
    GradEpi epi{I + img, J + img, gI ? gI + img : nullptr, gJ ? gJ + img : nullptr, W};
    box5_walk(k, W, load, epi);
}

long task_count(int N, int H, int W) {
    return (long)N * ((H + kRows - 1) / kRows) * ((W + kOutCols - 1) / kOutCols);
}

}  // namespace

size_t ncc_workspace_bytes(int N, int H, int W, int backward) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    return backward ? (size_t)5 * N * H * W * sizeof(float) : round_up((size_t)task_count(N, H, W) * sizeof(float), 256);
}

int launch_ncc(const float* I, const float* J, int N, int H, int W, float eps, float* loss_map, float* cc_sum, float* partial,
               cudaStream_t s) {
    const long nt = task_count(N, H, W);
    ncc_fwd_kernel<<<cdiv(nt, kWarps), kWarps * 32, 0, s>>>(I, J, H, W, nt, eps, loss_map, cc_sum ? partial : nullptr);
    CF_CHECK_LAUNCH("ncc_fwd_kernel");
    if (cc_sum) {
        ncc_reduce_kernel<<<1, 256, 0, s>>>(partial, nt, cc_sum);
        CF_CHECK_LAUNCH("ncc_reduce_kernel");
    }
    return CF_OK;
}

int launch_ncc_bwd(const float* I, const float* J, int N, int H, int W, float eps, const float* grad, int per_pixel, float mult,
                   float* gI, float* gJ, float* ws, cudaStream_t s) {
    const long nt = task_count(N, H, W);
    ncc_bwd_coef_kernel<<<cdiv(nt, kWarps), kWarps * 32, 0, s>>>(I, J, N, H, W, nt, eps, grad, per_pixel, mult, ws);
    CF_CHECK_LAUNCH("ncc_bwd_coef_kernel");
    ncc_bwd_gather_kernel<<<cdiv(nt, kWarps), kWarps * 32, 0, s>>>(I, J, N, H, W, nt, ws, gI, gJ);
    CF_CHECK_LAUNCH("ncc_bwd_gather_kernel");
    return CF_OK;
}

}  // namespace cf
