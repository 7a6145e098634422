// K1 (exact-precision mode): all-pairs correlation on the CUDA cores, fp32 FMA.
//
//   corr[b, p, q] = <f1[b, :, p], f2[b, :, q]> / sqrt(C)
//
// Replaces torchvision raft.py:424-431 / RAFT corr.py `CorrBlock.corr` (call site
// /root/reference/nnunet/network_architecture/SegFlowGaussian.py:929), where the division is a second
// elementwise pass over the whole volume.  This kernel is the CF_PREC_FP32 mode (the reference never runs
// this path in reduced precision) and serves the shapes the TMA/tcgen05 kernel rejects
// (H*W not a multiple of 4, odd sizes).  The tensor-core path is corr_tcgen05.cu.
//
// Both operands are "MN-major" in memory ([C, HW] with HW contiguous), so a k-slab of either is read with
// unit stride.  64x64 output tile, 16-deep k-slabs, 4x4 register micro-tile, 256 threads.
#include "common.cuh"

namespace cf {
namespace {

constexpr int TM = 64, TN = 64, TK = 16;

__global__ void __launch_bounds__(256) corr_simt_kernel(const float* __restrict__ f1, const float* __restrict__ f2,
                                                        float* __restrict__ out, int C, int HW, float div) {
    __shared__ float As[TK][TM];
    __shared__ float Bs[TK][TN];
    const int b = blockIdx.z;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    const float* A = f1 + (size_t)b * C * HW;
    const float* Bm = f2 + (size_t)b * C * HW;
    const int tid = threadIdx.x;
    const int tx = tid % 16, ty = tid / 16;  // micro-tile: rows ty*4.., cols tx*4..
    const int lk = tid / 64, lm = tid % 64;  // loader: 4 k-rows per pass x 64 columns

    float acc[4][4] = {};
    for (int k0 = 0; k0 < C; k0 += TK) {
#pragma unroll
        for (int kk = 0; kk < TK; kk += 4) {
            const int k = k0 + kk + lk;
            const bool kin = k < C;
            As[kk + lk][lm] = (kin && m0 + lm < HW) ? __ldg(A + (size_t)k * HW + m0 + lm) : 0.0f;
            Bs[kk + lk][lm] = (kin && n0 + lm < HW) ? __ldg(Bm + (size_t)k * HW + n0 + lm) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < TK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 c = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* O = out + (size_t)b * HW * HW;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= HW) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n < HW) O[(size_t)m * HW + n] = __fdiv_rn(acc[i][j], div);
        }
    }
}

}  // namespace

int launch_corr_simt(const float* f1, const float* f2, int B, int C, int HW, float* level0, cudaStream_t s) {
    dim3 grid(cdiv(HW, TN), cdiv(HW, TM), B);
    if (B > 65535) return fail(CF_ERR_UNSUPPORTED, "corr_simt: batch %d > 65535", B);
    corr_simt_kernel<<<grid, 256, 0, s>>>(f1, f2, level0, C, HW, sqrtf((float)C));
    CF_CHECK_LAUNCH("corr_simt_kernel");
    return CF_OK;
}

}  // namespace cf
