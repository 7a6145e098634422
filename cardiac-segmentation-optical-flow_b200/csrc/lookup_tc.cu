// K4 on the tensor cores: CorrBlock lookup WITHOUT a materialised volume (the optional path of north_star, BASELINE.json
// configs[3] "materialised vs on-the-fly").  Output contract: identical to corr_fn(coords1),
// /root/reference/nnunet/network_architecture/SegFlowGaussian.py:929-935 (arithmetic torchvision raft.py:394-431).
//
// avg_pool over q of <f1[p], f2[q]> == <f1[p], avg_pool(f2)[q]>, so level l of the correlation pyramid restricted to the
// windows the queries touch is a set of dot products between f1 and the l-times pooled fmap2.  Memory: B*C*HW*(1 + 4/3)
// bf16 values instead of B*HW^2*(4/3) floats (cfg4: 0.18 GB instead of 5.70 GB); the window correlations are recomputed by
// every lookup -- on tcgen05 instead of the fp32 CUDA cores of lookup_fused.cu.
//
// One CTA (persistent over query blocks) = 128 queries = an 8 x 16 patch of the fmap, all levels:
//   * the queries' f1 rows (128 x C bf16, K-major, SWIZZLE_128B, one TMA box per 64 channels) stay resident = UMMA operand A;
//   * per level the epilogue threads (thread = query = TMEM lane) reduce the bounding box of the block's windows and publish
//     it; the box is tiled by 8 x 16-position chunks, each chunk = one TMA box of the pooled fmap2 (zero-filled outside the
//     image = zero padding) = UMMA operand B, N = 128;
//   * D[query, position] accumulates over C in TMEM (double-buffered); the epilogue threads of a query gather the cells of
//     ITS (2r+4)^2 window out of its TMEM lane into shared memory ([cell][query]: bank = query, conflict-free for any
//     per-lane cell), and once the level's chunks are through evaluate the 81 bilinear outputs from the window -- dense,
//     divergence-free, with the same exact tap arithmetic as the materialised lookup (sampling.cuh: axis_tap);
//   * warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-17 = epilogue (four threads per query); bounding boxes of block n+1 are computed and
//     published while block n is being evaluated, so the producer and the tensor pipe never wait for the epilogue.
// A GEMM over a bounding box computes ~5x the window products a per-query kernel needs; at >= 1 PFLOP/s that is still
// an order of magnitude faster than the 30 TFLOP/s of CUDA-core FMAs.
#include "sampling.cuh"
#include "tc_common.cuh"

namespace cf {
namespace {
using namespace tc;

constexpr int QB_H = 8, QB_W = 16;         // query block: 8 rows x 16 columns of the fmap
constexpr int CH_H = 8, CH_W = 16;         // position chunk: 8 rows x 16 columns of a pooled fmap2 level
constexpr int NQ = 128, NPOS = 128;
constexpr int BK = 64, UK = 16;
constexpr int NSTAGE = 5;
constexpr int EPT = 4;                                           // epilogue threads per query
constexpr int EPI_WARPS = 4 * EPT, NTHREADS = 64 + 32 * EPI_WARPS;
constexpr int TMEM_COLS = 2 * NPOS;
constexpr int MAX_KB = 4;
constexpr uint32_t OP_BYTES = 128u * 128u;

struct LevelInfo { int bx0, by0, ncx, ncy; };

struct OtfParams {
    const float* coords;
    float* out;
    int B, C, H, W, HW, L;
    int k_blocks;
    int qb_x, qb_per_img;      // query blocks per fmap row of blocks / per image
    long n_blocks;
    int lh[CF_MAX_LEVELS], lw[CF_MAX_LEVELS];
    float rcp_h[CF_MAX_LEVELS], rcp_w[CF_MAX_LEVELS];
    float div, mul;
};

struct OtfMaps {
    CUtensorMap f1;
    CUtensorMap f2[CF_MAX_LEVELS];
};

constexpr int WIN_CELLS = (2 * CF_MAX_RADIUS + 4) * (2 * CF_MAX_RADIUS + 4);       // window of a query: (2r+4)^2 cells
constexpr int MISC_BYTES = 2 * CF_MAX_LEVELS * 16 + 2 * EPI_WARPS * CF_MAX_LEVELS * 16 + 32 * 8;   // tables, boxes, mbarriers
constexpr size_t smem_bytes(int k_blocks) {
    return 1024 + (size_t)k_blocks * OP_BYTES + (size_t)NSTAGE * OP_BYTES + (size_t)WIN_CELLS * NQ * 4 + MISC_BYTES;
}

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory"); }

template <int R>
__global__ void __launch_bounds__(NTHREADS, 1) lookup_tc_kernel(const __grid_constant__ OtfMaps M, const OtfParams P) {
    constexpr int SIDE = 2 * R + 1, PR = 2 * R + 4, K = SIDE * SIDE;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint32_t tmem_base_slot;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t s_a = base;                                              // [kb][128 x 128 B]   queries
    const uint32_t s_ring = s_a + (uint32_t)P.k_blocks * OP_BYTES;          // [stage][128 x 128 B] positions
    const uint32_t s_cs = s_ring + (uint32_t)NSTAGE * OP_BYTES;             // [window cell][query] fp32
    const uint32_t s_misc = s_cs + (uint32_t)WIN_CELLS * NQ * 4;
    uint8_t* const gen_base = smem_raw + (base - smem_u32(smem_raw));
    float* const Cs = reinterpret_cast<float*>(gen_base + (s_cs - base));
    LevelInfo* const info = reinterpret_cast<LevelInfo*>(gen_base + (s_misc - base));            // [2][CF_MAX_LEVELS]
    int* const red = reinterpret_cast<int*>(gen_base + (s_misc - base) + 2 * CF_MAX_LEVELS * 16);  // [2][4 warps][L][4]
    const uint32_t s_bars = s_misc + 2 * CF_MAX_LEVELS * 16 + 2 * EPI_WARPS * CF_MAX_LEVELS * 16;
    constexpr int kFull = 0, kEmpty = NSTAGE, kAFull = 2 * NSTAGE, kAEmpty = kAFull + 1, kAccFull = kAFull + 2,
                  kAccEmpty = kAFull + 4, kInfoFull = kAFull + 6, kInfoEmpty = kAFull + 8;
    auto bar = [&](int i) { return s_bars + 8u * (uint32_t)i; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long nblk = (P.n_blocks - (long)blockIdx.x + (long)gridDim.x - 1) / (long)gridDim.x;   // blocks of this CTA

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(bar(kFull + s), 1); mbar_init(bar(kEmpty + s), 1); }
        mbar_init(bar(kAFull), 1);
        mbar_init(bar(kAEmpty), 1);
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar(kAccFull + a), 1);
            mbar_init(bar(kAccEmpty + a), EPI_WARPS);
            mbar_init(bar(kInfoFull + a), 1);
            mbar_init(bar(kInfoEmpty + a), 2);
        }
        mbar_fence_init();
        tma_prefetch_desc(&M.f1);
        for (int l = 0; l < P.L; ++l) tma_prefetch_desc(&M.f2[l]);
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(smem_u32(&tmem_base_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (long n = 0; n < nblk; ++n) {
                const long blk = (long)blockIdx.x + n * (long)gridDim.x;
                const int b = (int)(blk / P.qb_per_img), rem = (int)(blk - (long)b * P.qb_per_img);
                const int qy0 = (rem / P.qb_x) * QB_H, qx0 = (rem % P.qb_x) * QB_W;
                const int slot = (int)(n & 1);
                mbar_wait(bar(kInfoFull + slot), (uint32_t)((n >> 1) & 1));
                LevelInfo li[CF_MAX_LEVELS];
                for (int l = 0; l < P.L; ++l) li[l] = info[slot * CF_MAX_LEVELS + l];
                mbar_arrive(bar(kInfoEmpty + slot));
                if (n > 0) mbar_wait(bar(kAEmpty), (uint32_t)((n - 1) & 1));   // the previous block's MMAs have retired
                mbar_expect_tx(bar(kAFull), (uint32_t)P.k_blocks * OP_BYTES);
                for (int kb = 0; kb < P.k_blocks; ++kb)
                    tma_load_4d(s_a + (uint32_t)kb * OP_BYTES, &M.f1, bar(kAFull), kb * BK, qx0, qy0, b);
                for (int l = 0; l < P.L; ++l)
                    for (int cy = 0; cy < li[l].ncy; ++cy)
                        for (int cx = 0; cx < li[l].ncx; ++cx)
                            for (int kb = 0; kb < P.k_blocks; ++kb) {
                                mbar_wait(bar(kEmpty + stage), (uint32_t)(phase ^ 1));
                                mbar_expect_tx(bar(kFull + stage), OP_BYTES);
                                tma_load_4d(s_ring + (uint32_t)stage * OP_BYTES, &M.f2[l], bar(kFull + stage), kb * BK,
                                            li[l].bx0 + cx * CH_W, li[l].by0 + cy * CH_H, b);
                                if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
                            }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = umma_idesc(NQ, NPOS);
            int stage = 0, phase = 0;
            long t = 0;
            for (long n = 0; n < nblk; ++n) {
                const int slot = (int)(n & 1);
                mbar_wait(bar(kInfoFull + slot), (uint32_t)((n >> 1) & 1));
                int chunks = 0;
                for (int l = 0; l < P.L; ++l) chunks += info[slot * CF_MAX_LEVELS + l].ncx * info[slot * CF_MAX_LEVELS + l].ncy;
                mbar_arrive(bar(kInfoEmpty + slot));
                mbar_wait(bar(kAFull), (uint32_t)(n & 1));
                tc_fence_after();
                for (int ch = 0; ch < chunks; ++ch, ++t) {
                    const int acc = (int)(t & 1);
                    mbar_wait(bar(kAccEmpty + acc), (uint32_t)(((t >> 1) & 1) ^ 1));
                    tc_fence_after();
                    const uint32_t d = tmem_base + (uint32_t)(acc * NPOS);
                    for (int kb = 0; kb < P.k_blocks; ++kb) {
                        mbar_wait(bar(kFull + stage), (uint32_t)phase);
                        tc_fence_after();
#pragma unroll
                        for (int ks = 0; ks < BK / UK; ++ks) {
                            const uint64_t ad = umma_desc(s_a + (uint32_t)kb * OP_BYTES + (uint32_t)ks * 32u, 1024u);
                            const uint64_t bd = umma_desc(s_ring + (uint32_t)stage * OP_BYTES + (uint32_t)ks * 32u, 1024u);
                            umma_bf16(d, ad, bd, idesc, (uint32_t)((kb | ks) != 0));
                        }
                        umma_commit(bar(kEmpty + stage));
                        if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
                    }
                    umma_commit(bar(kAccFull + acc));
                }
                umma_commit(bar(kAEmpty));     // every MMA that read this block's queries has retired when this arrives
            }
        }
    } else {
#include "lookup_tc_epilogue.inl"
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

}  // namespace
}  // namespace cf

#include "lookup_tc_host.inl"
