// K1: all-pairs correlation as a TMA-fed tcgen05 (5th-gen tensor core) GEMM with a fused epilogue.
//
//   corr[b, p, q] = <f1[b, :, p], f2[b, :, q]> / sqrt(C)          level 0  [B*HW, 1, H, W]
//   level1[b, p]  = avg_pool2d(corr[b, p], 2, 2)                   level 1  [B*HW, 1, H/2, W/2]  (fused)
//
// Replaces torchvision raft.py:424-431 + the first avg_pool2d of raft.py:390-392 (call site
// /root/reference/nnunet/network_architecture/SegFlowGaussian.py:929): there a cuBLAS batched SGEMM, a
// separate elementwise division over the whole volume and a pooling launch that re-reads it.
//
// Mapping to the hardware
//   (Version 1 of the build, round 1.  Since round 2 cabi.cu uses it for plain bf16 at W = 32 and for shapes version 2,
//   corr_tc2.cu, rejects; everything else -- the fp32-grade bf16x3 default, other widths, bf16 / channels-last inputs -- is v2.)
//   * The caller's maps are [C, HW] with HW contiguous, i.e. "MN-major" operands.  tcgen05 accepts them for kind::f16
//     (kind::tf32 returns zeros, tools/umma_probe.cu), but measured on B200 an MN-major 128x128x16 MMA takes ~195 cycles
//     instead of ~64-96 (per-role timeline, tools/gemm_ts.py with -DCF_BRINGUP).  So the bf16 rounding pre-pass TRANSPOSES
//     both maps to [B, HW, C] (one tiled pass) and the MMA reads canonical K-major SWIZZLE_128B operands: TMA boxes of 64
//     channels (128 B) x rows, 8-row groups 1024 B apart (SBO), K advanced by 32 B inside the swizzle atom.  fp32 accumulate
//     in TMEM.
//   * one CTA owns a 128-row block of queries p for one frame pair: the A block (128 x C bf16 = 64 KB at
//     C = 256) is loaded ONCE and stays resident; B streams through a 5-stage ring (NSTAGE) of 128(n) x 64(k) tiles;
//     accumulators (128 x BN fp32) are double-buffered in TMEM so the epilogue of tile t overlaps the MMAs of
//     tile t+1.  Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warps 2-5 =
//     epilogue (tcgen05.ld -> scale -> swizzled smem -> TMA store of level 0; second TMEM pass pools 2x2 over
//     the (h2, w2) axes and stores level 1 directly).
//   * BN = 2*W*k so that a tile covers whole pairs of image rows of fmap2 (the pooling window never crosses a
//     tile); the divide and the ((a+b)+c)+d)/4 order match ATen bit-for-bit given the fp32 accumulators.
// The kernel is output-write bound (K = C = 256 only): 64 KB of level 0 + 16 KB of level 1 per 128x128 tile.
#include <cuda.h>
#include <cuda_bf16.h>

#include <cstdlib>
#include <mutex>

#include "tc_common.cuh"

namespace cf {
namespace {

constexpr int BM = 128;        // query rows per CTA (UMMA M)
constexpr int BK = 64;         // k rows per B stage (4 MMAs of K = 16)
constexpr int UK = 16;         // K of one kind::f16 MMA
constexpr int NSTAGE = 5;
constexpr int NEPI = 5;         // level-0 staging buffers: TMA stores in flight per CTA (the kernel is store-latency bound)
constexpr int EPI_WARPS = 4;
constexpr int NTHREADS = 64 + 32 * EPI_WARPS;
constexpr int MAX_C = 256;     // A block resident in smem: C * 256 B
constexpr int MAX_BN = 128;
constexpr int TMEM_COLS = 256; // 2 accumulators x MAX_BN columns

struct GemmParams {
    float* level1;       // null -> no fused pooling
    float* level2;       // W == 32 fast epilogue only: levels 2 and 3 fused as well (null -> not fused)
    float* level3;
    int C, HW, H, W;
    int BN;              // tile width (multiple of 64, <= 128)
    int n_tiles;         // ceil(HW / BN)
    int n_split;         // CTAs per (b, m-tile) along n
    int m_tiles;
    float div;           // sqrt(C)
    float mul;           // 1/div when that is exact (power of two), else 0
    unsigned long long* ts;  // bring-up only: per-role timestamps of CTA `ts_cta` (null in production)
    int ts_cta;
    float* dbg;          // bring-up only: raw accumulators + smem operand dump of CTA 0 (null in production)
    int tile_l0;         // W == 32 epilogue: write level 0 as 4x4 tiles (CF_LAYOUT_TILED4)
    int tile_l1;         // generic epilogue: write the fused level 1 as 4x4 tiles
};

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#ifdef CF_BRINGUP   // bring-up builds only (tools/gemm_ts.py, tools/gemm_debug.py): the shipped library carries neither
#define CF_TS(slot) do { if (P.ts != nullptr && (int)blockIdx.x == P.ts_cta) P.ts[(slot)] = gtime(); } while (0)
#else
#define CF_TS(slot) do { } while (0)
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Bounded spin: a protocol bug traps (launch error on the host) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    uint32_t spins = 0;
    do {
        if (++spins == (1u << 28)) __trap();
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory"); }

// UMMA shared-memory descriptor, K-major operand, SWIZZLE_128B (cute/arch/mma_sm100_desc.hpp bit layout): rows of 128 B
// (64 bf16 along K), 8-row groups `sbo_bytes` apart; the leading-byte-offset field is unused for swizzled K-major (1)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) |
           (1ull << 46) /* sm_100 descriptor version */ | (2ull << 61) /* SWIZZLE_128B */;
}
// instruction descriptor: D = f32, A = B = bf16, both K-major, M x N
__device__ __forceinline__ uint32_t umma_idesc(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float pool4(float a, float b, float c, float d) {
    return __fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(a, b), c), d), 0.25f);  // /4 is exact as *0.25
}

// ---------------------------------------------------------------- shared memory carve-up (dynamic, 1024-aligned)
struct SmemLayout {
    uint32_t a;                 // [ceil(C/64) k-chunks][128 rows][128 B]
    uint32_t b0;                // NSTAGE x [BN rows][128 B] (one 64-channel k-block)
    uint32_t epi0;              // NEPI x [128 rows][128 B] swizzled staging for the level-0 TMA store
    uint32_t bars;              // mbarriers
    __device__ uint32_t b(int stage) const { return b0 + (uint32_t)stage * (MAX_BN * 128u); }
    __device__ uint32_t epi(int i) const { return epi0 + (uint32_t)i * (128u * 128u); }
};
constexpr int kBarAFull = 0, kBarBFull = 1, kBarBEmpty = 1 + NSTAGE, kBarAccFull = 1 + 2 * NSTAGE,
              kBarAccEmpty = 3 + 2 * NSTAGE, kNumBars = 5 + 2 * NSTAGE;

// W32: the feature maps are 32 wide and BN = 128, so a tile is exactly 4 image rows of fmap2: the epilogue keeps the
// whole accumulator row (128 fp32) in registers, reads TMEM once, and produces levels 0, 1, 2 and (over two consecutive
// tiles) 3 of the pyramid from them.
template <bool W32>
__global__ void __launch_bounds__(NTHREADS, 1)
corr_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmL1, const GemmParams P) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint32_t tmem_base_slot;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    SmemLayout S;
    S.a = base;
    const int k_blocks = (P.C + BK - 1) / BK;
    S.b0 = base + (uint32_t)k_blocks * (BM * 128u);
    S.epi0 = S.b0 + NSTAGE * MAX_BN * 128u;
    S.bars = S.epi0 + (uint32_t)NEPI * 128u * 128u;
    auto bar = [&](int i) { return S.bars + 8u * (uint32_t)i; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // work item: (b, m-tile, n-split)
    const int split = blockIdx.x % P.n_split;
    const int mt = (blockIdx.x / P.n_split) % P.m_tiles;
    const int b = blockIdx.x / (P.n_split * P.m_tiles);
    const int m0 = mt * BM;
    const int tiles_per = (P.n_tiles + P.n_split - 1) / P.n_split;
    const int nt_begin = split * tiles_per;
    const int nt_end = min(P.n_tiles, nt_begin + tiles_per);
    const int k_steps = P.C / UK;   // MMAs along K

    if (threadIdx.x == 0) {
        mbar_init(bar(kBarAFull), 1);
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(bar(kBarBFull + s), 1); mbar_init(bar(kBarBEmpty + s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar(kBarAccFull + a), 1); mbar_init(bar(kBarAccEmpty + a), EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // TMEM allocation (one warp), address lands in shared memory
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_base_slot)), "r"((uint32_t)TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            // the operands are written by the pre-pass right before this launch (launched with the PDL attribute: no effect otherwise);
            // this thread is the only one that reads them, and every store of this grid comes after MMAs that need these loads
            asm volatile("griddepcontrol.wait;" ::: "memory");
            mbar_expect_tx(bar(kBarAFull), (uint32_t)k_blocks * (BM * 128u));   // out-of-range channels / rows are zero-filled
            for (int kc = 0; kc < k_blocks; ++kc)
                tma_load_3d(S.a + (uint32_t)kc * (BM * 128u), &tmA, bar(kBarAFull), kc * BK, m0, b);
            int stage = 0, phase = 0;
            CF_TS(0);
            for (int nt = nt_begin; nt < nt_end; ++nt) {
                const int n0 = nt * P.BN;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(bar(kBarBEmpty + stage), phase ^ 1);
                    if (kb == 0) CF_TS(16 + (nt - nt_begin));
                    mbar_expect_tx(bar(kBarBFull + stage), (uint32_t)P.BN * 128u);
                    tma_load_3d(S.b(stage), &tmB, bar(kBarBFull + stage), kb * BK, n0, b);
                    if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = umma_idesc(BM, P.BN);
            mbar_wait(bar(kBarAFull), 0);
            tc_fence_after();
            CF_TS(1);
            int stage = 0, phase = 0;
            int t = 0;
            for (int nt = nt_begin; nt < nt_end; ++nt, ++t) {
                const int acc = t & 1;
                mbar_wait(bar(kBarAccEmpty + acc), ((t >> 1) & 1) ^ 1);
                tc_fence_after();
                CF_TS(32 + t);
                const uint32_t d = tmem_base + (uint32_t)(acc * MAX_BN);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(bar(kBarBFull + stage), phase);
                    tc_fence_after();
                    const int nks = min(BK / UK, k_steps - kb * (BK / UK));
                    for (int ks = 0; ks < nks; ++ks) {   // K = 16 per MMA = 32 B inside the 128 B swizzle row
                        const uint64_t ad = umma_desc(S.a + (uint32_t)kb * (BM * 128u) + (uint32_t)ks * 32u, 1024u);
                        const uint64_t bd = umma_desc(S.b(stage) + (uint32_t)ks * 32u, 1024u);
                        umma_bf16(d, ad, bd, idesc, (kb | ks) != 0);
                    }
                    umma_commit(bar(kBarBEmpty + stage));  // frees the stage when these MMAs retire
                    if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
                }
                umma_commit(bar(kBarAccFull + acc));
                CF_TS(48 + t);
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int quad = warp & 3;                       // TMEM lane quadrant this warp may read
        const int row = quad * 32 + lane;                // accumulator row == query p - m0
        const int ep_tid = (warp - 2) * 32 + lane;
        const bool row_ok = (m0 + row) < P.HW;
        const int W = P.W, W2 = P.W >> 1;
        int t = 0, sb = 0;
        float l2_keep[8] = {};                               // W32: level-2 row of the even tile, for level 3
        for (int nt = nt_begin; nt < nt_end; ++nt, ++t) {
            const int acc = t & 1;
            const int n0 = nt * P.BN;
            if (ep_tid == 0) CF_TS(64 + 4 * t);
            mbar_wait(bar(kBarAccFull + acc), (t >> 1) & 1);
            tc_fence_after();
            if (ep_tid == 0) CF_TS(65 + 4 * t);
            const uint32_t tacc = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * MAX_BN);
#ifdef CF_BRINGUP
            if (P.dbg != nullptr && blockIdx.x == 0 && t == 0) {
                for (int c = 0; c < P.BN / 32; ++c) {
                    float v[32];
                    tmem_ld32(tacc + (uint32_t)(c * 32), v);
                    tmem_wait_ld();
                    for (int i = 0; i < 32; ++i) P.dbg[row * 128 + c * 32 + i] = v[i];
                }
                const float* sa = reinterpret_cast<const float*>(smem_raw + (S.a - smem_u32(smem_raw)));
                const float* sbp = reinterpret_cast<const float*>(smem_raw + (S.b0 - smem_u32(smem_raw)));
                for (int i = ep_tid; i < P.C * 64; i += 128) P.dbg[16384 + i] = sa[i];
                for (int i = ep_tid; i < MAX_BN * 32; i += 128) P.dbg[16384 + P.C * 64 + i] = sbp[i];
                if (ep_tid == 0) { P.dbg[16384 + P.C * 64 + MAX_BN * 32] = __uint_as_float(tmem_base); }
            }
#endif
            if (W32) {
                // ---- one TMEM pass: the lane's 128 accumulators (4 image rows of 32) ----
                float v[128];
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld32(tacc + (uint32_t)(c * 32), v + 32 * c);
                tmem_wait_ld();
                // the accumulator is in registers: hand the TMEM buffer back to the MMA warp right away
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(kBarAccEmpty + acc));
                if (P.mul != 0.0f) {
#pragma unroll
                    for (int i = 0; i < 128; ++i) v[i] = __fmul_rn(v[i], P.mul);
                } else {
#pragma unroll
                    for (int i = 0; i < 128; ++i) v[i] = __fdiv_rn(v[i], P.div);
                }
                // levels 1..3 from registers, in avg_pool2d's summation order (bit-exact with the pooling kernels)
                float l1[32];                                 // 2 rows x 16
                if (P.level1 != nullptr) {
#pragma unroll
                    for (int rr = 0; rr < 2; ++rr)
#pragma unroll
                        for (int x = 0; x < 16; ++x)
                            l1[16 * rr + x] = pool4(v[64 * rr + 2 * x], v[64 * rr + 2 * x + 1], v[64 * rr + 32 + 2 * x],
                                                    v[64 * rr + 32 + 2 * x + 1]);
                }
                // level 0: four swizzled 128 x 32 boxes, level 1: a fifth one (128 queries x 32 floats) -> five TMA stores
                // (direct per-lane stores of level 1 are 1 KB apart across the warp and stall the epilogue)
                if (ep_tid == 0) tma_store_wait_read<0>();   // the previous tile's stores have read their buffers
                epi_bar_sync();
                if (ep_tid == 0) CF_TS(66 + 4 * t);
                // Tiled level 0: the tile's 128 accumulators are image rows 4k..4k+3 = one row of eight 4x4 tiles, i.e. the same
                // 128-float range of the query's image in both layouts; only the order inside it changes.  16 B piece j of
                // chunk c holds row (j & 3) of tile 2c + (j >> 2): a static permutation of register indices, no extra work.
                if (P.tile_l0) {
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const uint32_t dst = S.epi(c) + (uint32_t)row * 128u;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float* src = v + (j & 3) * 32 + (2 * c + (j >> 2)) * 4;
                            const uint32_t addr = dst + (uint32_t)((j ^ (row & 7)) << 4);
                            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(src[0]), "f"(src[1]),
                                         "f"(src[2]), "f"(src[3]) : "memory");
                        }
                    }
                }
#pragma unroll
                for (int c = 0; c < 5; ++c) {
                    if (c == 4 && P.level1 == nullptr) break;
                    if (c < 4 && P.tile_l0) continue;
                    const float* src = c < 4 ? v + 32 * c : l1;
                    const uint32_t dst = S.epi(c) + (uint32_t)row * 128u;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint32_t addr = dst + (uint32_t)((j ^ (row & 7)) << 4);
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(src[4 * j]), "f"(src[4 * j + 1]),
                                     "f"(src[4 * j + 2]), "f"(src[4 * j + 3]) : "memory");
                    }
                }
                fence_async_smem();
                epi_bar_sync();
                if (ep_tid == 0) {
#pragma unroll
                    for (int c = 0; c < 4; ++c) tma_store_3d(&tmC, S.epi(c), n0 + c * 32, m0, b);
                    if (P.level1 != nullptr) tma_store_3d(&tmL1, S.epi(4), (n0 >> 7) * 32, m0, b);
                    tma_store_commit();
                    CF_TS(67 + 4 * t);
                }
                if (P.level1 != nullptr) {
                    const size_t q = (size_t)b * P.HW + (size_t)(m0 + row);
                    if (P.level2 != nullptr) {
                        float l2[8];
#pragma unroll
                        for (int x = 0; x < 8; ++x) l2[x] = pool4(l1[2 * x], l1[2 * x + 1], l1[16 + 2 * x], l1[16 + 2 * x + 1]);
                        if (row_ok) {
                            float4* d2 = reinterpret_cast<float4*>(P.level2 + q * (size_t)((P.H >> 2) * 8) + (size_t)(n0 >> 7) * 8);
                            d2[0] = make_float4(l2[0], l2[1], l2[2], l2[3]);
                            d2[1] = make_float4(l2[4], l2[5], l2[6], l2[7]);
                        }
                        if (P.level3 == nullptr) {
                        } else if ((nt & 1) == 0) {
#pragma unroll
                            for (int x = 0; x < 8; ++x) l2_keep[x] = l2[x];
                        } else if (row_ok) {                      // tiles 2k, 2k+1 = level-2 rows 2k, 2k+1 = level-3 row k
                            float4 o3;
                            o3.x = pool4(l2_keep[0], l2_keep[1], l2[0], l2[1]);
                            o3.y = pool4(l2_keep[2], l2_keep[3], l2[2], l2[3]);
                            o3.z = pool4(l2_keep[4], l2_keep[5], l2[4], l2[5]);
                            o3.w = pool4(l2_keep[6], l2_keep[7], l2[6], l2[7]);
                            *reinterpret_cast<float4*>(P.level3 + q * (size_t)((P.H >> 3) * 4) + (size_t)(n0 >> 8) * 4) = o3;
                        }
                    }
                }
                continue;
            }
            // ---- pass 1: level 0, 32 columns at a time -> swizzled smem -> TMA store ----
            for (int c = 0; c < P.BN / 32; ++c) {
                float v[32];
                tmem_ld32(tacc + (uint32_t)(c * 32), v);
                tmem_wait_ld();
                if (P.mul != 0.0f) {          // the test stays outside the unrolled loops (inside, it is re-evaluated per element)
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __fmul_rn(v[i], P.mul);
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __fdiv_rn(v[i], P.div);
                }
                if (ep_tid == 0) tma_store_wait_read<NEPI - 1>();   // the store that used buffer `sb` NEPI chunks ago is done
                epi_bar_sync();
                const uint32_t dst = S.epi(sb) + (uint32_t)row * 128u;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t addr = dst + (uint32_t)((j ^ (row & 7)) << 4);
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v[4 * j]), "f"(v[4 * j + 1]),
                                 "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
                }
                fence_async_smem();
                epi_bar_sync();
                if (ep_tid == 0) {
                    tma_store_3d(&tmC, S.epi(sb), n0 + c * 32, m0, b);
                    tma_store_commit();
                }
                sb = (sb + 1 == NEPI) ? 0 : sb + 1;
            }
            // ---- pass 2: fused level 1 = 2x2 average over the (h2, w2) axes of the scaled values ----
            if (P.level1 != nullptr) {
                const int pairs = P.BN / (2 * W);            // image-row pairs covered by this tile
                const int pair0 = n0 / (2 * W);
                float* l1row = P.level1 + ((size_t)b * P.HW + (size_t)(m0 + row)) * (size_t)((P.H >> 1) * W2);
                for (int j = 0; j < pairs; ++j) {
                    for (int m = 0; m < W / 8; ++m) {
                        float top[8], bot[8];
                        tmem_ld8(tacc + (uint32_t)((2 * j) * W + 8 * m), top);
                        tmem_ld8(tacc + (uint32_t)((2 * j + 1) * W + 8 * m), bot);
                        tmem_wait_ld();
                        if (P.mul != 0.0f) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) { top[i] = __fmul_rn(top[i], P.mul); bot[i] = __fmul_rn(bot[i], P.mul); }
                        } else {
#pragma unroll
                            for (int i = 0; i < 8; ++i) { top[i] = __fdiv_rn(top[i], P.div); bot[i] = __fdiv_rn(bot[i], P.div); }
                        }
                        float4 o;
                        o.x = pool4(top[0], top[1], bot[0], bot[1]);
                        o.y = pool4(top[2], top[3], bot[2], bot[3]);
                        o.z = pool4(top[4], top[5], bot[4], bot[5]);
                        o.w = pool4(top[6], top[7], bot[6], bot[7]);
                        if (row_ok) {
                            const int y1 = pair0 + j;
                            float* dst = P.tile_l1 ? l1row + ((size_t)((y1 >> 2) * (W2 >> 2) + m) * 16 + (y1 & 3) * 4)
                                                   : l1row + (size_t)y1 * W2 + 4 * m;
                            *reinterpret_cast<float4*>(dst) = o;
                        }
                    }
                }
            }
            // accumulator drained -> hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kBarAccEmpty + acc));
        }
        if (ep_tid == 0) tma_store_wait_all();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------- host side
using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn get_encode_fn() {
    static EncodeFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeFn>(p);
    });
    return fn;
}

// 3-D tensor map: dims (d0 innermost), strides in bytes for d1, d2; 128-byte swizzle
int make_map(CUtensorMap* m, CUtensorMapDataType dtype, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1,
             uint64_t s2, uint32_t b0, uint32_t b1) {
    EncodeFn enc = get_encode_fn();
    if (!enc) return fail(CF_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    const cuuint64_t dims[3] = {d0, d1, d2};
    const cuuint64_t strides[2] = {s1, s2};
    const cuuint32_t box[3] = {b0, b1, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(m, dtype, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(CF_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return CF_OK;
}

// tile width for the fused-pool epilogue: BN = 2*W*k, multiple of 32, <= MAX_BN, k | H/2
int fused_bn(int H, int W) {
    if ((H & 1) || (W & 7)) return 0;
    int best = 0;
    for (int k = 1; 2 * W * k <= MAX_BN; ++k)
        if ((2 * W * k) % 32 == 0 && (H / 2) % k == 0) best = 2 * W * k;   // epilogue works in 32-column chunks
    return best;
}

// fp32 [B, C, HW] -> bf16 [B, HW, C] (round to nearest even): 64 x 64 tiles through shared memory, 128-bit loads along
// HW and 128-bit stores along C.  blockIdx.z = 2 * b + (0: fmap1, 1: fmap2).
__global__ void __launch_bounds__(256) to_bf16_t_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                        __nv_bfloat16* __restrict__ oa, __nv_bfloat16* __restrict__ ob, int C,
                                                        int HW) {
    __shared__ float tile[64][65];
    // programmatic dependent launch: the GEMM that follows in the stream may become resident (barrier set-up, TMEM allocation, tensor-map
    // prefetch) while this grid drains; its producer thread waits for this grid's completion before its first TMA load
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int which = blockIdx.z & 1, bi = blockIdx.z >> 1;
    const float* src = (which ? b : a) + (size_t)bi * C * HW;
    __nv_bfloat16* dst = (which ? ob : oa) + (size_t)bi * HW * C;
    const int p0 = blockIdx.x * 64, c0 = blockIdx.y * 64;
    {   // load: 16 threads x float4 cover 64 positions of one channel; 16 channels per pass
        const int tp = (threadIdx.x & 15) * 4, tc = threadIdx.x >> 4;
#pragma unroll
        for (int r = 0; r < 64; r += 16) {
            const int c = c0 + tc + r, pp = p0 + tp;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c < C && pp < HW) v = ldg_f4(src + (size_t)c * HW + pp);   // HW % 4 == 0
            tile[tc + r][tp] = v.x; tile[tc + r][tp + 1] = v.y; tile[tc + r][tp + 2] = v.z; tile[tc + r][tp + 3] = v.w;
        }
    }
    __syncthreads();
    {   // store: 8 threads x 8 channels (16 B) cover 64 channels of one position; 32 positions per pass
        const int tc = (threadIdx.x & 7) * 8, tp = threadIdx.x >> 3;
#pragma unroll
        for (int r = 0; r < 64; r += 32) {
            const int pp = p0 + tp + r, c = c0 + tc;
            if (pp < HW && c < C) {   // C % 8 == 0
                __nv_bfloat162 q0 = __floats2bfloat162_rn(tile[tc][tp + r], tile[tc + 1][tp + r]);
                __nv_bfloat162 q1 = __floats2bfloat162_rn(tile[tc + 2][tp + r], tile[tc + 3][tp + r]);
                __nv_bfloat162 q2 = __floats2bfloat162_rn(tile[tc + 4][tp + r], tile[tc + 5][tp + r]);
                __nv_bfloat162 q3 = __floats2bfloat162_rn(tile[tc + 6][tp + r], tile[tc + 7][tp + r]);
                uint4 o;
                o.x = *reinterpret_cast<uint32_t*>(&q0); o.y = *reinterpret_cast<uint32_t*>(&q1);
                o.z = *reinterpret_cast<uint32_t*>(&q2); o.w = *reinterpret_cast<uint32_t*>(&q3);
                *reinterpret_cast<uint4*>(dst + (size_t)pp * C + c) = o;
            }
        }
    }
}

}  // namespace

bool corr_tcgen05_supported(int B, int C, int H, int W, bool fuse_pool) {
    const long HW = (long)H * W;
    if (B < 1 || C % UK != 0 || C > MAX_C || C % 8 != 0 || HW % 8 != 0 || HW < 64) return false;
    if (fuse_pool && fused_bn(H, W) == 0) return false;
    return get_encode_fn() != nullptr;
}

#ifdef CF_BRINGUP
static float* g_dbg = nullptr;
void corr_tcgen05_set_debug(float* p) { g_dbg = p; }
static unsigned long long* g_ts = nullptr;
static int g_ts_cta = 0;
void corr_tcgen05_set_timestamps(unsigned long long* p, int cta) { g_ts = p; g_ts_cta = cta; }
#else
static float* const g_dbg = nullptr;            // no process-global mutable state in the shipped library
static unsigned long long* const g_ts = nullptr;
static const int g_ts_cta = 0;
#endif

size_t corr_tcgen05_workspace_bytes(int B, int C, int H, int W) {
    return 2 * round_up((size_t)B * C * H * W * sizeof(__nv_bfloat16), 256);
}

int launch_corr_tcgen05(const float* f1, const float* f2, int B, int C, int H, int W, float* const* levels, int L,
                        void* workspace, size_t ws_bytes, int* levels_done, bool allow_tiled, unsigned* tiled_mask,
                        cudaStream_t s) {
    if (tiled_mask) *tiled_mask = 0;
    const int HW = H * W;
    float* level0 = levels[0];
    float* level1 = (L > 1 && fused_bn(H, W) != 0) ? levels[1] : nullptr;
    if (!workspace || ws_bytes < corr_tcgen05_workspace_bytes(B, C, H, W) || ((uintptr_t)workspace & 255u))
        return fail(CF_ERR_BAD_ARG, "corr_tcgen05: workspace of %zu bytes (256-byte aligned) required, got %zu",
                    corr_tcgen05_workspace_bytes(B, C, H, W), ws_bytes);
    __nv_bfloat16* h1 = reinterpret_cast<__nv_bfloat16*>(workspace);
    __nv_bfloat16* h2 = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(workspace) +
                                                          round_up((size_t)B * C * HW * sizeof(__nv_bfloat16), 256));
    {
        if (2L * B > 65535) return fail(CF_ERR_UNSUPPORTED, "corr_tcgen05: batch %d > 32767", B);
        dim3 grid((unsigned)cdiv(HW, 64), (unsigned)cdiv(C, 64), (unsigned)(2 * B));
        to_bf16_t_kernel<<<grid, 256, 0, s>>>(f1, f2, h1, h2, C, HW);
        CF_CHECK_LAUNCH("to_bf16_t_kernel");
    }
    GemmParams P{};
    P.level1 = level1;
    P.C = C; P.HW = HW; P.H = H; P.W = W;
    P.dbg = g_dbg;
    P.ts = g_ts;
    P.ts_cta = g_ts_cta;
    P.BN = level1 ? fused_bn(H, W) : MAX_BN;
    P.n_tiles = (HW + P.BN - 1) / P.BN;
    P.m_tiles = (HW + BM - 1) / BM;
    P.div = sqrtf((float)C);
    int e;
    const float fr = frexpf(P.div, &e);
    P.mul = (fr == 0.5f) ? 1.0f / P.div : 0.0f;  // power of two: scaling by the reciprocal is exact
    // one CTA per SM at a time (smem): split the n range only until every SM has a CTA -- each CTA re-loads its 64 KB A
    // block before its first MMA, so fewer, longer CTAs win once the machine is full (measured at cfg2: 1 > 2 > 4)
    int split = 1;
    while ((long)B * P.m_tiles * split < sm_count() && P.n_tiles / (split * 2) >= 1) split *= 2;
    static const int force_split = [] { const char* e = getenv("CF_GEMM_SPLIT"); return e ? atoi(e) : 0; }();   // tuning knob
    if (force_split > 0 && P.n_tiles / force_split >= 1) split = force_split;
    P.n_split = split;
    // register-resident epilogue: a tile = 4 whole image rows; levels 2 / 3 need tiles in aligned pairs per CTA
    const bool w32 = level1 && W == 32 && P.BN == 128 && (H % 4) == 0 && g_dbg == nullptr;
    const int tiles_per = (P.n_tiles + split - 1) / split;
    if (w32 && L > 2) P.level2 = levels[2];
    if (w32 && L > 3 && (H % 8) == 0 && (tiles_per % 2) == 0 && P.n_tiles % split == 0) P.level3 = levels[3];
    *levels_done = P.level3 ? 4 : (P.level2 ? 3 : (level1 ? 2 : 1));
    // Levels as 4x4 tiles (fewer DRAM atoms per lookup window; worth it for rows >= 128 B).  W == 32: the register-resident
    // epilogue reorders level 0 for free.  Other widths: the fused level 1 is tiled by addressing (the pooling chain reads it
    // with its tiled-input kernel); level 0 stays row-major there -- a 5-D tensor-map store of 16-byte tile rows (boxes
    // {4,8,1,1,128} and {4 of 16,8,1,128,1}, SWIZZLE_128B) faulted on B200 in both forms and was dropped.
    // Measured at cfg4: tiling level 1 speeds the lookup up by 7 % (183 -> 171 us) but a 128-column accumulator tile holds one
    // level-1 row only, so a 64 B tile is completed by four different tiles' scattered 16 B stores and the build loses more
    // (3.85 -> 4.13 ms per step): bit 1 is off by default.
    const char* tile_env = getenv("CF_GEMM_TILED");                  // bit 0: level 0 (W == 32), bit 1: level 1 (other widths)
    const int tile_knob = tile_env ? atoi(tile_env) : 1;
    const bool can_tile = allow_tiled && tiled_mask != nullptr && g_dbg == nullptr;
    if (w32) {
        P.tile_l0 = (can_tile && (tile_knob & 1)) ? 1 : 0;
    } else if (can_tile) {
        P.tile_l1 = (level1 && (tile_knob & 2) && (W / 2) % 4 == 0 && (H / 2) % 4 == 0 && W / 2 >= 32) ? 1 : 0;
    }
    if (tiled_mask) *tiled_mask = (P.tile_l0 ? 1u : 0u) | (P.tile_l1 ? 2u : 0u);

    CUtensorMap tmA, tmB, tmC;
    int rc;
    // operands [B, HW, C] bf16: box = 64 channels (one 128 B swizzle row) x rows
    const uint64_t sR = (uint64_t)C * 2, sB = (uint64_t)C * HW * 2;
    if ((rc = make_map(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, h1, C, HW, B, sR, sB, BK, BM))) return rc;
    if ((rc = make_map(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, h2, C, HW, B, sR, sB, BK, (uint32_t)P.BN))) return rc;
    if ((rc = make_map(&tmC, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, level0, HW, HW, B, (uint64_t)HW * 4, (uint64_t)HW * HW * 4, 32,
                       BM)))
        return rc;

    CUtensorMap tmL1 = tmC;   // only read by the W32 epilogue
    if (w32 && (rc = make_map(&tmL1, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, level1, (uint64_t)(H / 2) * 16, HW, B, (uint64_t)(H / 2) * 64,
                              (uint64_t)HW * (H / 2) * 64, 32, BM)))
        return rc;
    const size_t smem = 1024 + (size_t)((C + BK - 1) / BK) * BM * 128 + (size_t)NSTAGE * MAX_BN * 128 + NEPI * 128 * 128 + kNumBars * 8 + 64;
    // opt in to the large dynamic shared memory (sized for MAX_C; a per-device attribute: once per device and kernel)
    {
        const size_t max_smem = 1024 + MAX_C * 256 + NSTAGE * MAX_BN * 128 + NEPI * 128 * 128 + kNumBars * 8 + 64;
        int rc2 = smem_optin(reinterpret_cast<const void*>(&corr_tcgen05_kernel<false>), max_smem, 16);
        if (!rc2) rc2 = smem_optin(reinterpret_cast<const void*>(&corr_tcgen05_kernel<true>), max_smem, 17);
        if (rc2) return rc2;
    }
    const long grid = (long)B * P.m_tiles * P.n_split;
    static const bool pdl = [] { const char* e = getenv("CF_GEMM_PDL"); return e ? atoi(e) != 0 : true; }();   // A/B knob
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = pdl ? 1 : 0;
    if (w32) CF_CHECK_CUDA(cudaLaunchKernelEx(&cfg, corr_tcgen05_kernel<true>, tmA, tmB, tmC, tmL1, P));
    else CF_CHECK_CUDA(cudaLaunchKernelEx(&cfg, corr_tcgen05_kernel<false>, tmA, tmB, tmC, tmL1, P));
    return CF_OK;
}

}  // namespace cf
