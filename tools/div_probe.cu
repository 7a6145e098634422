// Exhaustive check (run on the GPU box): is a reciprocal-multiply + FMA-correction division bit-identical to the
// IEEE division __fdiv_rn(a, s) for the divisors the coordinate normalisation uses (s = size - 1, small integers)?
//   v1: q0 = a*r; q = fma(fma(-q0, s, a), r, q0)                       (one correction)
//   v2: v1 followed by a second correction step
// r = 1/s correctly rounded (host division).  All fp32 a with exponent in [-24, 14) are tested, both variants,
// every s in [1, smax].   nvcc -arch=sm_100a -O3 -o tools/div_probe tools/div_probe.cu && tools/div_probe 255
#include <cstdio>
#include <cstdlib>
#include <cstdint>

__global__ void probe(int s_int, unsigned long long* bad1, unsigned long long* bad2, uint32_t lo_bits, uint32_t hi_bits) {
    const float s = (float)s_int;
    const float r = __frcp_rn(s);
    unsigned long long b1 = 0, b2 = 0;
    const uint64_t n = (uint64_t)hi_bits - lo_bits;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const float a = __uint_as_float(lo_bits + (uint32_t)i);
        const float ref = __fdiv_rn(a, s);
        const float q0 = __fmul_rn(a, r);
        const float q1 = __fmaf_rn(__fmaf_rn(-q0, s, a), r, q0);
        const float q2 = __fmaf_rn(__fmaf_rn(-q1, s, a), r, q1);
        b1 += (__float_as_uint(q1) != __float_as_uint(ref));
        b2 += (__float_as_uint(q2) != __float_as_uint(ref));
    }
    if (b1) atomicAdd(bad1, b1);
    if (b2) atomicAdd(bad2, b2);
}

int main(int argc, char** argv) {
    const int smax = argc > 1 ? atoi(argv[1]) : 255;
    unsigned long long *d1, *d2, h1, h2, t1 = 0, t2 = 0;
    cudaMalloc(&d1, 8);
    cudaMalloc(&d2, 8);
    // positive floats with biased exponent 103 (2^-24) .. 141 (2^14): bit patterns [103 << 23, 141 << 23)
    const uint32_t lo = 103u << 23, hi = 141u << 23;
    for (int s = 1; s <= smax; ++s) {
        cudaMemset(d1, 0, 8);
        cudaMemset(d2, 0, 8);
        probe<<<148 * 8, 256>>>(s, d1, d2, lo, hi);
        cudaMemcpy(&h1, d1, 8, cudaMemcpyDeviceToHost);
        cudaMemcpy(&h2, d2, 8, cudaMemcpyDeviceToHost);
        if (h1 || h2) printf("s=%d mismatches: one-step %llu  two-step %llu\n", s, h1, h2);
        t1 += h1;
        t2 += h2;
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("divisors 1..%d, %u values each: one-step mismatches %llu, two-step mismatches %llu (%s)\n", smax, hi - lo, t1, t2,
           cudaGetErrorString(e));
    return 0;
}
