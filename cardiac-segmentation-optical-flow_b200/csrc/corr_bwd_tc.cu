// Backward of the all-pairs correlation w.r.t. both feature maps on the tensor cores (the reference trains
// through CorrBlock: nnunet/training/network_training/SegFlowGaussian.py:2225 `l.mean().backward()`, the block is built at
// nnunet/network_architecture/SegFlowGaussian.py:929-935).  With G = d loss / d corr level 0, fp32 [B, HW_p, HW_q] (the pooling
// chain's backward has already folded the coarser levels into it, backward.cu):
//     gf1[b, c, p] = 1/sqrt(C) * sum_q G[b, p, q] * f2[b, c, q]          "gemm 0":  M = p, K = q
//     gf2[b, c, q] = 1/sqrt(C) * sum_p G[b, p, q] * f1[b, c, p]          "gemm 1":  M = q, K = p
// Round 1 ran these as fp32 CUDA-core GEMMs (1.2 ms at cfg2).  Here: tcgen05, D[128 spatial x 128 channels] in TMEM, fp32-grade
// through the bf16x3 split (x = hi + lo, hi*hi + hi*lo + lo*hi into one fp32 accumulator; the dropped lo*lo term is 2^-16 relative).
//   * operand B = feature map [c][k] (K-major as stored, NCHW): split once into bf16 hi / lo planes by a pre-pass, fetched by TMA;
//   * operand A = G: fp32 in HBM and by far the larger stream (HW^2 per pair against C*HW).  Eight loader warps read it with
//     coalesced loads, split it in registers and store the bf16 hi / lo tiles into the 128-byte-swizzled K-major layout the UMMA
//     descriptors expect -- for gemm 1 the same store transposes the tile (a thread owns one q = one operand row and writes 8
//     consecutive p = one 16-byte chunk), so G is never transposed in memory;
//   * warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-9 = loaders, then the epilogue (TMEM -> registers -> 1/sqrt(C) -> NCHW
//     stores, 32 consecutive positions per warp store).  The loader warps alternate K blocks in two groups of four so that two
//     register buffers per thread keep four K blocks (128 KB per SM) in flight.
// One CTA = one 128 x 128 output tile over the whole K = HW; the two channel halves of a tile sit in adjacent CTAs and share G in L2.
// Per K block (64): 3 x 4 MMAs of 128 x 128 x 16 = 0.68 us of tensor time against 32 KB of G = 0.73 us of HBM time per SM:
// the split costs about what streaming G costs.
#include "tc_common.cuh"

namespace cf {
namespace {
using namespace tc;

constexpr int BM = 128, BN = 128, BK = 64, UK = 16;
constexpr int NSTAGE = 3;
constexpr uint32_t TILE_BYTES = 128u * 128u;             // one operand tile: 128 rows x 64 bf16
constexpr uint32_t STAGE_BYTES = 4u * TILE_BYTES;        // A hi, A lo, B hi, B lo
constexpr int LOADER_WARPS = 16, GROUP_WARPS = 8;   // two groups of 8 warps (a thread = 4 chunks of a K block); with 2 x 4 warps there was one
                                                    // loader warp per scheduler and the kernel ran at its dependent-issue latency (1.7 us per K block)
constexpr int CPT = 128 * 8 / (32 * GROUP_WARPS);   // 16-byte chunks (8 K values of one operand row) per loader thread and K block
constexpr int NTHREADS = 64 + 32 * LOADER_WARPS;
constexpr int TMEM_COLS = BN;
constexpr size_t kSmemBytes = 1024 + (size_t)NSTAGE * STAGE_BYTES + 256;

struct BwdParams {
    const float* g;          // [B, HW, HW]
    float* out[2];           // gf1, gf2 (NCHW); gemm i writes out[i]
    int gemm_of[2];          // grid slot -> gemm index
    int B, C, HW, k_blocks, m_tiles, n_tiles;
    float div;
};
struct BwdMaps {
    CUtensorMap f[2][2];     // [gemm][hi / lo]: gemm 0 reads fmap2, gemm 1 reads fmap1; dims (HW, C, B) bf16
};

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 8 floats -> 8 bf16 hi (x rounded) + 8 bf16 lo (x - hi rounded), packed in K order.  Two values per step: one packed convert
// for hi, a shift / mask to get hi back as floats, two subtractions, one packed convert for lo.
__device__ __forceinline__ void split8(const float* v, uint4& hi, uint4& lo) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 hp = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);      // .x (low half) = the lower K index
        h[i] = *reinterpret_cast<const uint32_t*>(&hp);
        const float h0 = __uint_as_float(h[i] << 16), h1 = __uint_as_float(h[i] & 0xffff0000u);
        const __nv_bfloat162 lp = __floats2bfloat162_rn(__fsub_rn(v[2 * i], h0), __fsub_rn(v[2 * i + 1], h1));
        l[i] = *reinterpret_cast<const uint32_t*>(&lp);
    }
    hi = make_uint4(h[0], h[1], h[2], h[3]);
    lo = make_uint4(l[0], l[1], l[2], l[3]);
}

template <int GEMM>      // 0: gf1 (rows p, K = q); 1: gf2 (rows q, K = p): one specialised body per product, chosen per CTA
__device__ __forceinline__ void corr_bwd_tc_body(const BwdMaps& M, const BwdParams& P, uint8_t* smem_raw, uint32_t& tmem_base_slot,
                                                 int tile) {
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* const gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t s_bars = base + (uint32_t)NSTAGE * STAGE_BYTES;
    constexpr int kFull = 0, kEmpty = NSTAGE, kAcc = 2 * NSTAGE;
    auto bar = [&](int i) { return s_bars + 8u * (uint32_t)i; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mt = tile / P.n_tiles, nt = tile - mt * P.n_tiles;
    constexpr int gemm = GEMM;
    const int b = (int)blockIdx.y;
    const int m0 = mt * BM, n0 = nt * BN;
    const int HW = P.HW, KB = P.k_blocks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(bar(kFull + s), 1 + GROUP_WARPS); mbar_init(bar(kEmpty + s), 1); }
        mbar_init(bar(kAcc), 1);
        mbar_fence_init();
        tma_prefetch_desc(&M.f[gemm][0]);
        tma_prefetch_desc(&M.f[gemm][1]);
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(smem_u32(&tmem_base_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer: the feature-map operand =====================
        if (lane == 0) {
            for (int kb = 0; kb < KB; ++kb) {
                const int s = kb % NSTAGE;
                mbar_wait(bar(kEmpty + s), (uint32_t)(((kb / NSTAGE) & 1) ^ 1));
                mbar_expect_tx(bar(kFull + s), 2u * TILE_BYTES);
                const uint32_t dst = base + (uint32_t)s * STAGE_BYTES + 2u * TILE_BYTES;
                tma_load_3d(dst, &M.f[gemm][0], bar(kFull + s), kb * BK, n0, b);
                tma_load_3d(dst + TILE_BYTES, &M.f[gemm][1], bar(kFull + s), kb * BK, n0, b);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = umma_idesc(BM, BN);
            for (int kb = 0; kb < KB; ++kb) {
                const int s = kb % NSTAGE;
                mbar_wait(bar(kFull + s), (uint32_t)((kb / NSTAGE) & 1));
                tc_fence_after();
                const uint32_t sa = base + (uint32_t)s * STAGE_BYTES, sb = sa + 2u * TILE_BYTES;
#pragma unroll
                for (int prod = 0; prod < 3; ++prod) {        // hi*hi, hi*lo, lo*hi
                    const uint32_t a = sa + (prod == 2 ? TILE_BYTES : 0u), bb = sb + (prod == 1 ? TILE_BYTES : 0u);
#pragma unroll
                    for (int ks = 0; ks < BK / UK; ++ks)
                        umma_bf16(tmem_base, umma_desc(a + (uint32_t)ks * 32u, 1024u), umma_desc(bb + (uint32_t)ks * 32u, 1024u), idesc,
                                  (uint32_t)((kb | prod | ks) != 0));
                }
                umma_commit(bar(kEmpty + s));
            }
            umma_commit(bar(kAcc));
        }
    } else {
#include "corr_bwd_tc_loader.inl"
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// One launch covers both products (grid.x = slots x tiles): 928 CTAs = 6.3 waves at cfg2 instead of two launches of 3.1 waves each.
__global__ void __launch_bounds__(NTHREADS, 1) corr_bwd_tc_kernel(const __grid_constant__ BwdMaps M, const BwdParams P) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint32_t tmem_base_slot;
    const int tiles = P.m_tiles * P.n_tiles;
    const int slot = (int)blockIdx.x / tiles, tile = (int)blockIdx.x - slot * tiles;
    if (P.gemm_of[slot] == 0) corr_bwd_tc_body<0>(M, P, smem_raw, tmem_base_slot, tile);
    else corr_bwd_tc_body<1>(M, P, smem_raw, tmem_base_slot, tile);
}

// fp32 -> bf16 hi / lo planes, same [B, C, HW] layout
__global__ void __launch_bounds__(256) split_planes_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ hi,
                                                           __nv_bfloat16* __restrict__ lo, size_t n8) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n8; i += (size_t)gridDim.x * 256) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * i), c = __ldg(reinterpret_cast<const float4*>(src) + 2 * i + 1);
        const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
        uint4 h, l;
        split8(v, h, l);
        reinterpret_cast<uint4*>(hi)[i] = h;
        reinterpret_cast<uint4*>(lo)[i] = l;
    }
}

}  // namespace
}  // namespace cf

#include "corr_bwd_tc_host.inl"
#include "corr_bwd_tc_noalloc.inl"
