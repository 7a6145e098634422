// K2: pyramid pooling chain  level l -> l+1  =  F.avg_pool2d(x, 2, stride=2) (floor mode).
//
// Replaces torchvision raft.py:390-392 (3 pooling launches per CorrBlock; call site
// /root/reference/nnunet/network_architecture/SegFlowGaussian.py:929).  Bit-exact with ATen: the window is
// summed in (row, column) order and divided by 4.  HBM-bound: reads level l once, writes level l+1 once.
// When the correlation GEMM fuses level 1 into its epilogue this only produces levels 2.. (from_level = 1).
#include "common.cuh"

namespace cf {
namespace {

__device__ __forceinline__ float pool4(float a, float b, float c, float d) {
    return __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(a, b), c), d), 4.0f);
}

// two adjacent outputs per thread: 2 x float4 in, 1 x float2 out (needs w % 4 == 0)
__global__ void __launch_bounds__(256) pool2x2_vec_kernel(const float* __restrict__ in, float* __restrict__ out, int h,
                                                          int w, long total /* n_img * h2 * (w2/2) */) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int h2 = h >> 1, w2 = w >> 1, wq = w2 >> 1;
    const int xq = (int)(t % wq);
    const long r = t / wq;
    const int y = (int)(r % h2);
    const long n = r / h2;
    const float* src = in + ((size_t)n * h + 2 * y) * w + 4 * xq;
    const float4 a = ldg_f4(src), b = ldg_f4(src + w);
    float2 o;
    o.x = pool4(a.x, a.y, b.x, b.y);
    o.y = pool4(a.z, a.w, b.z, b.w);
    *reinterpret_cast<float2*>(out + ((size_t)n * h2 + y) * w2 + 2 * xq) = o;
}

__global__ void __launch_bounds__(256) pool2x2_kernel(const float* __restrict__ in, float* __restrict__ out, int h,
                                                      int w, long total /* n_img * h2 * w2 */) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int h2 = h >> 1, w2 = w >> 1;
    const int x = (int)(t % w2);
    const long r = t / w2;
    const int y = (int)(r % h2);
    const long n = r / h2;
    const float* src = in + ((size_t)n * h + 2 * y) * w + 2 * x;
    out[t] = pool4(__ldg(src), __ldg(src + 1), __ldg(src + w), __ldg(src + w + 1));
}

// input stored as 4x4 tiles (CF_LAYOUT_TILED4), output row-major: one thread per tile = 64 contiguous bytes in, a 2x2
// block out.  Same summation order as above.
__global__ void __launch_bounds__(256) pool2x2_tiled_in_kernel(const float* __restrict__ in, float* __restrict__ out, int h,
                                                               int w, long total /* n_img * (h/4) * (w/4) */) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int tw = w >> 2, th = h >> 2, w2 = w >> 1, h2 = h >> 1;
    const int tx = (int)(t % tw);
    const long r = t / tw;
    const int ty = (int)(r % th);
    const long n = r / th;
    const float* src = in + (size_t)t * 16;
    const float4 r0 = ldg_f4(src), r1 = ldg_f4(src + 4), r2 = ldg_f4(src + 8), r3 = ldg_f4(src + 12);
    float* o = out + ((size_t)n * h2 + 2 * ty) * w2 + 2 * tx;
    *reinterpret_cast<float2*>(o) = make_float2(pool4(r0.x, r0.y, r1.x, r1.y), pool4(r0.z, r0.w, r1.z, r1.w));
    *reinterpret_cast<float2*>(o + w2) = make_float2(pool4(r2.x, r2.y, r3.x, r3.y), pool4(r2.z, r2.w, r3.z, r3.w));
}

}  // namespace

int launch_pool2x2_tiled_in(const float* in, float* out, long n_img, int h, int w, cudaStream_t s) {
    if ((h & 3) || (w & 3) || !aligned16(in) || (reinterpret_cast<uintptr_t>(out) & 7u))
        return fail(CF_ERR_BAD_ARG, "pool2x2 (tiled input): %dx%d level cannot be 4x4-tiled", h, w);
    const long total = n_img * (h / 4) * (w / 4);
    pool2x2_tiled_in_kernel<<<cdiv(total, 256), 256, 0, s>>>(in, out, h, w, total);
    CF_CHECK_LAUNCH("pool2x2_tiled_in_kernel");
    return CF_OK;
}

int launch_pool2x2(const float* in, float* out, long n_img, int h, int w, cudaStream_t s) {
    if (w % 4 == 0 && aligned16(in) && ((reinterpret_cast<uintptr_t>(out) & 7u) == 0)) {
        const long total = n_img * (h / 2) * (w / 4);
        pool2x2_vec_kernel<<<cdiv(total, 256), 256, 0, s>>>(in, out, h, w, total);
    } else {
        const long total = n_img * (h / 2) * (w / 2);
        pool2x2_kernel<<<cdiv(total, 256), 256, 0, s>>>(in, out, h, w, total);
    }
    CF_CHECK_LAUNCH("pool2x2_kernel");
    return CF_OK;
}

int launch_pool_chain(float* pyramid, const LevelDims& d, long n_img, int L, int from_level, cudaStream_t s,
                      unsigned tiled_mask) {
    char* base = reinterpret_cast<char*>(pyramid);
    for (int l = from_level; l + 1 < L; ++l) {
        const float* in = reinterpret_cast<const float*>(base + d.off[l]);
        float* out = reinterpret_cast<float*>(base + d.off[l + 1]);
        const int rc = ((tiled_mask >> l) & 1u) ? launch_pool2x2_tiled_in(in, out, n_img, d.h[l], d.w[l], s)
                                                : launch_pool2x2(in, out, n_img, d.h[l], d.w[l], s);
        if (rc) return rc;
    }
    return CF_OK;
}

}  // namespace cf
