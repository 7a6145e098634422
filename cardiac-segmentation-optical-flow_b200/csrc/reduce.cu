// Deterministic float64 sum of an fp32 device array: the per-unit result rows of the cohort driver (checksum of the last
// lookup, sum of the warped frames; cardioflow_b200/shard.py: sum_f64, bench.py --workload cfg5) are compared bit for bit
// between the rank that computed a unit and a recomputation elsewhere, so the summation order is a function of n alone.
// Replaces `x.sum(dtype=torch.float64)`, which ATen runs as a cast copy (4n read + 8n written) plus a reduction over the
// copy (8n read): here the array is read once (4n) with 128-bit loads and converted in registers.
#include "common.cuh"

namespace cf {
namespace {

constexpr int kSumCtas = 148 * 4, kSumThreads = 256;

__device__ __forceinline__ double block_sum(double a, double* sh) {
    sh[threadIdx.x] = a;
    __syncthreads();
#pragma unroll
    for (int o = kSumThreads / 2; o; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    return sh[0];
}

__global__ void __launch_bounds__(kSumThreads) sum_f64_partial_kernel(const float* __restrict__ x, size_t n, size_t chunk,
                                                                      double* __restrict__ partial) {
    __shared__ double sh[kSumThreads];
    const size_t begin = (size_t)blockIdx.x * chunk, end = begin + chunk < n ? begin + chunk : n;   // chunk % 4 == 0
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    if (begin < end) {
        const size_t n4 = (end - begin) / 4;
        const float4* v = reinterpret_cast<const float4*>(x + begin);
        for (size_t i = threadIdx.x; i < n4; i += kSumThreads) {
            const float4 f = __ldcs(v + i);
            a0 += (double)f.x; a1 += (double)f.y; a2 += (double)f.z; a3 += (double)f.w;
        }
        for (size_t i = begin + 4 * n4 + threadIdx.x; i < end; i += kSumThreads) a0 += (double)x[i];
    }
    const double s = block_sum((a0 + a1) + (a2 + a3), sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

__global__ void __launch_bounds__(kSumThreads) sum_f64_final_kernel(const double* __restrict__ partial, double* __restrict__ out) {
    __shared__ double sh[kSumThreads];
    double a = 0.0;
    for (int i = threadIdx.x; i < kSumCtas; i += kSumThreads) a += partial[i];
    const double s = block_sum(a, sh);
    if (threadIdx.x == 0) *out = s;
}

}  // namespace

size_t sum_f64_workspace_bytes() { return kSumCtas * sizeof(double); }

int launch_sum_f64(const float* x, size_t n, double* out, double* partial, cudaStream_t s) {
    const size_t chunk = round_up((n + kSumCtas - 1) / kSumCtas, 4);
    sum_f64_partial_kernel<<<kSumCtas, kSumThreads, 0, s>>>(x, n, chunk, partial);
    CF_CHECK_LAUNCH("sum_f64_partial_kernel");
    sum_f64_final_kernel<<<1, kSumThreads, 0, s>>>(partial, out);
    CF_CHECK_LAUNCH("sum_f64_final_kernel");
    return CF_OK;
}

}  // namespace cf
