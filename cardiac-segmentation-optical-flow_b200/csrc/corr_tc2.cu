// K1 + K2, version 2: all-pairs correlation + the first two pooled levels as ONE persistent tcgen05 kernel.
//
//   corr[b, p, q] = <f1[b, :, p], f2[b, :, q]> / sqrt(C)      level 0  [B*HW, 1, H, W]
//   level 1, 2    = avg_pool2d chain over the q axes           fused in the epilogue (levels >= 3: pool.cu)
//
// Replaces torchvision raft.py:424-431 + raft.py:387-392 (call site /root/reference/nnunet/network_architecture/
// SegFlowGaussian.py:929): there a cuBLAS batched SGEMM, an elementwise division over the whole volume and three
// pooling launches that re-read it.
//
// Differences to version 1 (corr_tcgen05.cu, kept for shapes this kernel rejects):
//   * ROLES SWAPPED: the fmap2 positions q are the UMMA M dimension (TMEM lanes), the queries p the N dimension (TMEM
//     columns).  A tcgen05.ld then hands every epilogue thread ONE q and 32 consecutive p: for a fixed register the 32
//     lanes of a warp hold 32 positions of the SAME query's image, so level 0 goes out with plain coalesced 128-byte
//     st.global -- no shared-memory staging, no TMA store, and any layout is just address arithmetic: levels 0 and 1
//     can be written as 4x4 tiles (CF_LAYOUT_TILED4) for EVERY width (version 1 could only tile level 0 at W = 32).
//   * an M tile is a 4-row x 32-column patch of the fmap2 image loaded as four TMA boxes (channels, 8 x, 4 y), one per
//     TMEM lane quadrant: the quadrant of an epilogue warp is then a 4 x 8 block of q, i.e. two 4x4 tiles = 128 contiguous bytes of a
//     tiled level 0, two full rows of 4 level-1 values and one pair of level-2 values: levels 1 and 2 are pooled with
//     warp shuffles in avg_pool2d's summation order ((a+b)+c)+d (bit-identical to pooling the stored level 0).
//   * PERSISTENT: one CTA per SM walks a contiguous range of (pair, query block, q tile) units, so the 232 one-per-SM
//     CTAs of version 1 at cfg2 (1.57 waves: the second wave ran on 84 SMs) become 148 equal shares.
//   * precision CF_PREC_BF16X3: both maps are split hi + lo in bf16 by the pre-pass and every K step issues
//     hi*hi + hi*lo + lo*hi into the same fp32 TMEM accumulator -- an fp32-grade volume (~1e-6) on the tensor pipe.  The
//     freed staging memory is what makes the doubled operands fit.
//   * inputs may be fp32 or bf16, NCHW or channels-last; bf16 channels-last maps feed the TMA directly (no pre-pass).
//
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warps 2-17 = epilogue (four warps per TMEM
// lane quadrant, 32 queries each).  The 128-query block (K-major, SWIZZLE_128B) stays resident in shared memory while
// the q tiles stream through a ring; fp32 accumulators are double-buffered in TMEM.
#include <cstdlib>
#include <mutex>

#include "tc_common.cuh"

namespace cf {

// ---------------------------------------------------------------- host helpers declared in tc_common.cuh
TmapEncodeFn tmap_encode_fn() {
    static TmapEncodeFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<TmapEncodeFn>(p);
    });
    return fn;
}

int make_tmap_bf16(CUtensorMap* m, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box) {
    TmapEncodeFn enc = tmap_encode_fn();
    if (!enc) return fail(CF_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    cuuint64_t d[5], st[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(CF_ERR_CUDA, "cuTensorMapEncodeTiled (rank %d) failed with CUresult %d", rank, (int)r);
    return CF_OK;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute of a kernel: set once per (device, slot)
int smem_optin(const void* func, size_t bytes, int slot) {
    static std::mutex mu;
    static bool done[64][32] = {};
    int dev = 0;
    CF_CHECK_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && slot >= 0 && slot < 32 && done[dev][slot]) return CF_OK;
    CF_CHECK_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    if (dev >= 0 && dev < 64 && slot >= 0 && slot < 32) done[dev][slot] = true;
    return CF_OK;
}

int sm_count() {
    static std::mutex mu;
    static int cache[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    std::lock_guard<std::mutex> lock(mu);
    if (cache[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev] = n;
    }
    return cache[dev];
}

namespace {
using namespace tc;

constexpr int TQ = 128;          // fmap2 positions per tile (UMMA M): 4 rows x 32 columns
constexpr int NPS = 128;         // queries per sub-block; a resident block is NB sub-blocks = UMMA N of 128 or 256
constexpr int BK = 64, UK = 16;  // channels per ring stage / per MMA
constexpr int EPI_WARPS = 16;         // four per TMEM lane quadrant
constexpr int EPQ = 32;               // queries per epilogue warp and sub-block (one tcgen05.ld x32)
constexpr int NTHREADS = 64 + 32 * EPI_WARPS;
constexpr int MAX_KB = 4;           // C <= 256
constexpr uint32_t OP_BYTES = 128u * 128u;   // one 128-row x 64-channel bf16 operand block

// NB = 2 (plain bf16 only: the split operands would not fit): 256 queries resident, every q stage feeds ONE N = 256 MMA, so the
// operand re-reads from L2 are halved.  Measured slower than NB = 1 (see launch_corr_tc2): an experiment knob, not the default.
template <bool SPLIT, int NB> struct Cfg {
    static_assert(NB == 1 || (NB == 2 && !SPLIT), "two sub-blocks only without the lo parts");
    static constexpr int PARTS = SPLIT ? 2 : 1;          // hi (+ lo)
    static constexpr int NSTAGE = SPLIT ? 3 : (NB == 2 ? 5 : 6);   // (9 stages measured no faster than 5: L2-throughput bound)
    static constexpr int NP = NPS * NB;                  // UMMA N
    static constexpr int TMEM_COLS = 2 * NP;             // two accumulators
    static constexpr uint32_t STAGE_BYTES = OP_BYTES * PARTS;
    static constexpr size_t smem_bytes(int k_blocks) {
        return 1024 + (size_t)k_blocks * OP_BYTES * PARTS * NB + (size_t)NSTAGE * STAGE_BYTES + 256;
    }
};

struct G2Params {
    float* level0;
    float* level1;       // null: not produced
    float* level2;       // null: not produced
    int B, C, H, W, HW;
    int k_blocks;
    int qt_x, QT, PB;    // q tiles per image row of tiles / per image; query blocks per image
    long units;          // B * PB * QT
    float div, mul;      // sqrt(C); 1/sqrt(C) when exact (power of two), else 0
    int tile_l0, tile_l1;
    int sched;           // 0: contiguous shares of all units, 1: whole keys round-robin (see the kernel)
    int skip;            // bring-up timing knob (CF_TC2_SKIP): bit 0 no level-0 stores, bit 1 no level 1, bit 2 no level 2
};

__device__ __forceinline__ float pool4(float a, float b, float c, float d) {
    return __fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(a, b), c), d), 0.25f);   // avg_pool2d order; /4 is exact as *0.25
}

template <bool SPLIT, int NB>
__global__ void __launch_bounds__(NTHREADS, 1)
corr_tc2_kernel(const __grid_constant__ CUtensorMap tmQh, const __grid_constant__ CUtensorMap tmQl,
                const __grid_constant__ CUtensorMap tmPh, const __grid_constant__ CUtensorMap tmPl, const G2Params P) {
    using K = Cfg<SPLIT, NB>;
    constexpr int NSTAGE = K::NSTAGE, NP = K::NP, TMEM_COLS = K::TMEM_COLS;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint32_t tmem_base_slot;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t s_res = base;                                                        // [part][kb][128 x 128 B]
    const uint32_t s_ring = s_res + (uint32_t)P.k_blocks * OP_BYTES * K::PARTS * NB;     // [stage][part][128 x 128 B]
    const uint32_t s_bars = s_ring + (uint32_t)NSTAGE * K::STAGE_BYTES;
    constexpr int kFull = 0, kEmpty = NSTAGE, kPFull = 2 * NSTAGE, kPEmpty = 2 * NSTAGE + 1, kAccFull = 2 * NSTAGE + 2,
                  kAccEmpty = 2 * NSTAGE + 4;
    auto bar = [&](int i) { return s_bars + 8u * (uint32_t)i; };
    auto res = [&](int part, int kb) { return s_res + (uint32_t)((part * P.k_blocks + kb) * NB) * OP_BYTES; };   // NB x 128 rows, contiguous
    auto ring = [&](int stage, int part) { return s_ring + (uint32_t)stage * K::STAGE_BYTES + (uint32_t)part * OP_BYTES; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // Work of this CTA, as a sequence of (key = pair * PB + query block, q tile) units with the q tile fastest:
    //   sched 0: a contiguous 1/gridDim share of all units (equal shares; a CTA reloads its query block 2-3 times);
    //   sched 1: whole keys dealt round-robin -- at any time the CTAs work on neighbouring keys, i.e. on a few pairs, so the
    //            fmap2 operand they all re-read stays in L2 when the batch's operands exceed it (cfg4: 134 MB).
    // 32-bit arithmetic (the host checks units < 2^31) and incremental stepping: no division per tile.
    const int G = (int)gridDim.x, bid = (int)blockIdx.x;
    const int n_keys = P.B * P.PB;
    const int u_begin = (int)((long)P.units * bid / G);
    const int n_it = P.sched == 0 ? (int)((long)P.units * (bid + 1) / G) - u_begin : ((n_keys - bid + G - 1) / G) * P.QT;
    struct UnitIter {
        int key, qt, step, QT;
        __device__ __forceinline__ void next() { if (++qt == QT) { qt = 0; key += step; } }
        __device__ __forceinline__ int next_key() const { return qt + 1 == QT ? key + step : key; }
    };
    const UnitIter it0 = P.sched == 0 ? UnitIter{u_begin / P.QT, u_begin % P.QT, 1, P.QT} : UnitIter{bid, 0, G, P.QT};

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(bar(kFull + s), 1); mbar_init(bar(kEmpty + s), 1); }
        mbar_init(bar(kPFull), 1);
        mbar_init(bar(kPEmpty), 1);
        for (int a = 0; a < 2; ++a) { mbar_init(bar(kAccFull + a), 1); mbar_init(bar(kAccEmpty + a), EPI_WARPS); }
        mbar_fence_init();
        tma_prefetch_desc(&tmQh);
        tma_prefetch_desc(&tmPh);
        if (SPLIT) { tma_prefetch_desc(&tmQl); tma_prefetch_desc(&tmPl); }
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(smem_u32(&tmem_base_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0, phase = 0, n_res = 0;
            int cur_key = -1, pb = 0, b = 0;
            UnitIter it = it0;
            for (int n = 0; n < n_it; ++n, it.next()) {
                const int key = it.key, qt = it.qt;        // key = b * PB + pb
                if (key != cur_key) {
                    b = key / P.PB;
                    pb = key - b * P.PB;
                    // the MMAs of the previous query block must have retired before its operand is overwritten
                    if (n_res > 0) mbar_wait(bar(kPEmpty), (uint32_t)((n_res - 1) & 1));
                    mbar_expect_tx(bar(kPFull), (uint32_t)P.k_blocks * OP_BYTES * K::PARTS * NB);
                    for (int kb = 0; kb < P.k_blocks; ++kb)
#pragma unroll
                        for (int nb = 0; nb < NB; ++nb) {
                            tma_load_3d(res(0, kb) + (uint32_t)nb * OP_BYTES, &tmPh, bar(kPFull), kb * BK, pb * NP + nb * NPS, b);
                            if (SPLIT) tma_load_3d(res(1, kb) + (uint32_t)nb * OP_BYTES, &tmPl, bar(kPFull), kb * BK, pb * NP + nb * NPS, b);
                        }
                    cur_key = key;
                    ++n_res;
                }
                const int ty = qt / P.qt_x, tx = qt - ty * P.qt_x;
                for (int kb = 0; kb < P.k_blocks; ++kb) {
                    mbar_wait(bar(kEmpty + stage), (uint32_t)(phase ^ 1));
                    mbar_expect_tx(bar(kFull + stage), K::STAGE_BYTES);
                    // four boxes (64 channels, 8 x, 4 y) = 32 rows = 4 KB each: smem row = 32 * x-block + 8 * y + x (= TMEM lane);
                    // x-blocks beyond the map are zero-filled
#pragma unroll
                    for (int xb = 0; xb < 4; ++xb) {
                        tma_load_4d(ring(stage, 0) + (uint32_t)xb * 4096u, &tmQh, bar(kFull + stage), kb * BK, 32 * tx + 8 * xb, 4 * ty, b);
                        if (SPLIT)
                            tma_load_4d(ring(stage, 1) + (uint32_t)xb * 4096u, &tmQl, bar(kFull + stage), kb * BK, 32 * tx + 8 * xb, 4 * ty, b);
                    }
                    if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = umma_idesc(TQ, NP);
            int stage = 0, phase = 0, n_res = 0;
            int cur_key = -1;
            int t = 0;
            UnitIter it = it0;
            for (int n = 0; n < n_it; ++n, ++t, it.next()) {
                const int key = it.key;
                if (key != cur_key) {
                    mbar_wait(bar(kPFull), (uint32_t)(n_res & 1));
                    ++n_res;
                    cur_key = key;
                }
                const int acc = (int)(t & 1);
                mbar_wait(bar(kAccEmpty + acc), (uint32_t)(((t >> 1) & 1) ^ 1));
                tc_fence_after();
                const uint32_t d = tmem_base + (uint32_t)(acc * NP);
                for (int kb = 0; kb < P.k_blocks; ++kb) {
                    mbar_wait(bar(kFull + stage), (uint32_t)phase);
                    tc_fence_after();
#pragma unroll
                    for (int ks = 0; ks < BK / UK; ++ks) {   // K = 16 per MMA = 32 B inside the 128 B swizzle row
                        const uint64_t qh = umma_desc(ring(stage, 0) + (uint32_t)ks * 32u, 1024u);
                        const uint64_t ph = umma_desc(res(0, kb) + (uint32_t)ks * 32u, 1024u);
                        umma_bf16(d, qh, ph, idesc, (uint32_t)((kb | ks) != 0));
                        if (SPLIT) {
                            const uint64_t ql = umma_desc(ring(stage, 1) + (uint32_t)ks * 32u, 1024u);
                            const uint64_t pl = umma_desc(res(1, kb) + (uint32_t)ks * 32u, 1024u);
                            umma_bf16(d, qh, pl, idesc, 1u);
                            umma_bf16(d, ql, ph, idesc, 1u);
                        }
                    }
                    umma_commit(bar(kEmpty + stage));        // frees the stage when these MMAs retire
                    if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
                }
                umma_commit(bar(kAccFull + acc));
                if (n + 1 == n_it || it.next_key() != key) umma_commit(bar(kPEmpty));   // last use of this query block
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int quad = warp & 3;                 // TMEM lane quadrant this warp may read = x-block of the q tile
        const int part = (warp - 2) >> 2;          // which EPQ of the 128 queries
        const int r = lane >> 3, c = lane & 7;     // position of this lane's q inside the quadrant's 4 x 8 block
        const int W = P.W, HW = P.HW;
        const int w1 = W >> 1, w2 = W >> 2;
        const unsigned img1 = (unsigned)((P.H >> 1) * w1), img2 = (unsigned)((P.H >> 2) * w2);
        const bool K0 = (c & 1) != 0, K1 = (r & 1) != 0;       // position of this lane's q inside its 2x2 block
        const int kk = (K1 ? 2 : 0) + (K0 ? 1 : 0);
        const bool B0 = (c & 2) != 0, B1 = (r & 2) != 0;       // position of that 2x2 block inside its 4x4 block
        const bool do1 = P.level1 != nullptr && !(P.skip & 2), do2 = do1 && P.level2 != nullptr && !(P.skip & 4);
        int t = 0, cur_key = -1, pb = 0, b = 0;
        UnitIter it = it0;
        for (int n = 0; n < n_it; ++n, ++t, it.next()) {
            const int key = it.key, qt = it.qt;
            if (key != cur_key) {
                cur_key = key;
                b = key / P.PB;
                pb = key - b * P.PB;
            }
            const int ty = qt / P.qt_x, tx = qt - ty * P.qt_x;
            const int acc = (int)(t & 1);
            mbar_wait(bar(kAccFull + acc), (uint32_t)((t >> 1) & 1));
            tc_fence_after();
            const int x0 = 32 * tx + 8 * quad;
            const int y = 4 * ty + r, x = x0 + c;
            const size_t q0 = P.tile_l0 ? (size_t)(((y >> 2) * (W >> 2) + (x >> 2)) * 16 + (y & 3) * 4 + (x & 3))
                                        : (size_t)y * W + x;
            const int y1 = y >> 1, x1 = x >> 1;
            const size_t q1 = P.tile_l1 ? (size_t)(((y1 >> 2) * (w1 >> 2) + (x1 >> 2)) * 16 + (y1 & 3) * 4 + (x1 & 3))
                                        : (size_t)y1 * w1 + x1;
            const size_t q2 = (size_t)(y >> 2) * w2 + (x >> 2);
#pragma unroll 1
          for (int nb = 0; nb < NB; ++nb) {
            float v[EPQ];
            tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * NP + nb * NPS + part * EPQ), v);
            tmem_wait_ld();
            if (nb == NB - 1) {
                // the last of this warp's accumulator columns are in registers: hand the TMEM buffer back to the MMA warp
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(kAccEmpty + acc));
            }
            const int p_first = pb * NP + nb * NPS + part * EPQ;   // first query of this warp's share of the sub-block
            const int n_valid = min(EPQ, HW - p_first);         // warp-uniform; <= 0 when the block hangs over the map
            if (x0 >= W || n_valid <= 0) continue;              // warp-uniform (W is a multiple of 8)
            const size_t img = (size_t)b * HW + p_first;        // first image (query) index of this warp
            float* const o0 = P.level0 + img * HW + q0;
            float* const o1 = do1 ? P.level1 + img * img1 + q1 : nullptr;
            float* const o2 = do2 ? P.level2 + img * img2 + q2 : nullptr;
            // Straight-line code over the warp's queries (no branch inside: the 32 shuffle / add chains interleave);
            // stores are predicated, everything else is computed for all lanes.
            if (P.mul != 0.0f) {
#pragma unroll
                for (int j = 0; j < EPQ; ++j) v[j] = __fmul_rn(v[j], P.mul);
            } else {
#pragma unroll
                for (int j = 0; j < EPQ; ++j) v[j] = __fdiv_rn(v[j], P.div);
            }
            if (P.skip & 1) {
            } else if (n_valid == EPQ) {                      // the common case: no per-store predicate
#pragma unroll
                for (int j = 0; j < EPQ; ++j) o0[(unsigned)j * (unsigned)HW] = v[j];
            } else {
#pragma unroll
                for (int j = 0; j < EPQ; ++j)
                    if (j < n_valid) o0[(unsigned)j * (unsigned)HW] = v[j];
            }
            if (do1) {
                // 2x2 mean over (y, x) of a query's image, in avg_pool2d's order ((a+b)+c)+d.  The four lanes of a 2x2 block
                // exchange round-robin: for every group of four queries each lane ends up with all four values of ITS block
                // for ONE of the four queries (the one matching its position in the block), so three shuffles serve four
                // queries (a plain "+1, +8, +9" gather needs three per query and uses a quarter of the lanes).
#pragma unroll
                for (int g = 0; g < EPQ / 4; ++g) {
                    const float v0 = v[4 * g], v1 = v[4 * g + 1], v2 = v[4 * g + 2], v3 = v[4 * g + 3];
                    const float t0 = K0 ? v1 : v0, t1 = K0 ? v0 : v1, u0 = K0 ? v3 : v2, u1 = K0 ? v2 : v3;
                    const float own = K1 ? u0 : t0;                                        // my value of query 4g + kk
                    const float r1 = __shfl_xor_sync(0xffffffffu, K1 ? u1 : t1, 1);        // x neighbour's value of MY query
                    const float r2 = __shfl_xor_sync(0xffffffffu, K1 ? t0 : u0, 8);        // y neighbour's
                    const float r3 = __shfl_xor_sync(0xffffffffu, K1 ? t1 : u1, 9);        // diagonal neighbour's
                    const float x0 = K1 ? r2 : own, x1 = K1 ? r3 : r1;                     // top row: my column, other column
                    const float y0 = K1 ? own : r2, y1 = K1 ? r1 : r3;                     // bottom row
                    const float l1 = pool4(K0 ? x1 : x0, K0 ? x0 : x1, K0 ? y1 : y0, K0 ? y0 : y1);
                    const bool ok = 4 * g + kk < n_valid;
                    if (ok) o1[(unsigned)(4 * g + kk) * img1] = l1;
                    if (do2) {
                        // the same exchange between the four 2x2 blocks of a 4x4 block (lanes of equal block position hold
                        // level-1 values of the same query)
                        const float q1 = __shfl_xor_sync(0xffffffffu, l1, 2);
                        const float q2 = __shfl_xor_sync(0xffffffffu, l1, 16);
                        const float q3 = __shfl_xor_sync(0xffffffffu, l1, 18);
                        const float s0 = B1 ? q2 : l1, s1 = B1 ? q3 : q1;
                        const float z0 = B1 ? l1 : q2, z1 = B1 ? q1 : q3;
                        const float l2 = pool4(B0 ? s1 : s0, B0 ? s0 : s1, B0 ? z1 : z0, B0 ? z0 : z1);
                        if (ok && !B0 && !B1) o2[(unsigned)(4 * g + kk) * img2] = l2;
                    }
                }
            }
          }   // sub-blocks
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// ---------------------------------------------------------------- pre-pass: any input -> bf16 [B, HW, C] (+ lo part)
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }

// [B, C, HW] -> [B, HW, C]: 64 x 64 tiles through shared memory.  blockIdx.z = 2 * b + (0: fmap1, 1: fmap2).
// lo = bf16(x - float(hi)) when requested (CF_PREC_BF16X3).
template <typename TIn>
__global__ void __launch_bounds__(256) prep_transpose_kernel(const TIn* __restrict__ a, const TIn* __restrict__ b,
                                                             __nv_bfloat16* __restrict__ ha, __nv_bfloat16* __restrict__ hb,
                                                             __nv_bfloat16* __restrict__ la, __nv_bfloat16* __restrict__ lb,
                                                             int C, int HW, int nmaps) {
    __shared__ float tile[64][65];
    const int which = nmaps == 2 ? (blockIdx.z & 1) : 0, bi = nmaps == 2 ? (blockIdx.z >> 1) : blockIdx.z;
    const TIn* src = (which ? b : a) + (size_t)bi * C * HW;
    __nv_bfloat16* hi = (which ? hb : ha) + (size_t)bi * HW * C;
    __nv_bfloat16* lo = (which ? lb : la);
    if (lo) lo += (size_t)bi * HW * C;
    const int p0 = blockIdx.x * 64, c0 = blockIdx.y * 64;
    {   // load: lanes along HW (coalesced), 4 channels per pass and thread row
        const int tp = threadIdx.x & 63, tc = threadIdx.x >> 6;
#pragma unroll
        for (int rr = 0; rr < 64; rr += 4) {
            const int ch = c0 + tc + rr, pp = p0 + tp;
            tile[tc + rr][tp] = (ch < C && pp < HW) ? to_f32(src[(size_t)ch * HW + pp]) : 0.0f;
        }
    }
    __syncthreads();
    {   // store: 8 threads x 8 channels (16 B) cover 64 channels of one position; 32 positions per pass
        const int tc = (threadIdx.x & 7) * 8, tp = threadIdx.x >> 3;
#pragma unroll
        for (int rr = 0; rr < 64; rr += 32) {
            const int pp = p0 + tp + rr, ch = c0 + tc;
            if (pp < HW && ch < C) {   // C % 8 == 0
                uint32_t h[4], l[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float x0 = tile[tc + 2 * k][tp + rr], x1 = tile[tc + 2 * k + 1][tp + rr];
                    const __nv_bfloat162 q = __floats2bfloat162_rn(x0, x1);
                    h[k] = *reinterpret_cast<const uint32_t*>(&q);
                    const __nv_bfloat162 ql = __floats2bfloat162_rn(x0 - __low2float(q), x1 - __high2float(q));
                    l[k] = *reinterpret_cast<const uint32_t*>(&ql);
                }
                *reinterpret_cast<uint4*>(hi + (size_t)pp * C + ch) = make_uint4(h[0], h[1], h[2], h[3]);
                if (lo) *reinterpret_cast<uint4*>(lo + (size_t)pp * C + ch) = make_uint4(l[0], l[1], l[2], l[3]);
            }
        }
    }
}

// fp32 channels-last [B, HW, C] -> bf16 (+ lo), elementwise; n4 = number of float4 groups of ONE map
__global__ void __launch_bounds__(256) prep_convert_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                           __nv_bfloat16* __restrict__ ha, __nv_bfloat16* __restrict__ hb,
                                                           __nv_bfloat16* __restrict__ la, __nv_bfloat16* __restrict__ lb,
                                                           size_t n4) {
    const size_t stride = (size_t)gridDim.x * 256;
    const int which = blockIdx.y;
    const float* src = which ? b : a;
    __nv_bfloat16* hi = which ? hb : ha;
    __nv_bfloat16* lo = which ? lb : la;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += stride) {
        const float4 x = ldg_f4(src + 4 * i);
        const __nv_bfloat162 q0 = __floats2bfloat162_rn(x.x, x.y), q1 = __floats2bfloat162_rn(x.z, x.w);
        *reinterpret_cast<uint2*>(hi + 4 * i) =
            make_uint2(*reinterpret_cast<const uint32_t*>(&q0), *reinterpret_cast<const uint32_t*>(&q1));
        if (lo) {
            const __nv_bfloat162 l0 = __floats2bfloat162_rn(x.x - __low2float(q0), x.y - __high2float(q0));
            const __nv_bfloat162 l1 = __floats2bfloat162_rn(x.z - __low2float(q1), x.w - __high2float(q1));
            *reinterpret_cast<uint2*>(lo + 4 * i) =
                make_uint2(*reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1));
        }
    }
}

size_t part_bytes(int B, int C, int HW) { return round_up((size_t)B * C * HW * sizeof(__nv_bfloat16), 256); }

}  // namespace

bool corr_tc2_supported(int B, int C, int H, int W) {
    static const bool force_v1 = [] { const char* e = getenv("CF_FORCE_V1"); return e && atoi(e) != 0; }();   // A/B knob
    if (force_v1) return false;
    if (B < 1 || C < 8 || C % 8 != 0 || C > 64 * MAX_KB || H < 4 || H % 4 != 0 || W < 8 || W % 8 != 0) return false;
    if ((long)H * W > (1L << 24)) return false;
    return tmap_encode_fn() != nullptr;
}

size_t corr_tc2_workspace_bytes(int B, int C, int H, int W, bool split) {
    return (split ? 4 : 2) * part_bytes(B, C, H * W);
}

// Converts both maps to bf16 channels-last in `workspace` (hi parts first, then lo parts when split) unless they already
// are (zero-copy).  Returns the four operand pointers.
int corr_tc2_prepare(const void* f1, const void* f2, int in_dtype, int in_layout, int B, int C, int HW, bool split,
                     void* workspace, size_t ws_bytes, const __nv_bfloat16** h1, const __nv_bfloat16** h2,
                     const __nv_bfloat16** l1, const __nv_bfloat16** l2, cudaStream_t s) {
    *l1 = *l2 = nullptr;
    if (in_dtype == CF_DTYPE_BF16 && in_layout == CF_LAYOUT_NHWC) {
        *h1 = reinterpret_cast<const __nv_bfloat16*>(f1);
        *h2 = reinterpret_cast<const __nv_bfloat16*>(f2);
        return CF_OK;
    }
    const size_t pbytes = part_bytes(B, C, HW);
    const size_t need = (split ? 4 : 2) * pbytes;
    if (!workspace || ws_bytes < need || ((uintptr_t)workspace & 255u))
        return fail(CF_ERR_BAD_ARG, "corr build: workspace of %zu bytes (256-byte aligned) required, got %zu", need, ws_bytes);
    char* w = reinterpret_cast<char*>(workspace);
    __nv_bfloat16* o1 = reinterpret_cast<__nv_bfloat16*>(w);
    __nv_bfloat16* o2 = reinterpret_cast<__nv_bfloat16*>(w + pbytes);
    __nv_bfloat16* p1 = split ? reinterpret_cast<__nv_bfloat16*>(w + 2 * pbytes) : nullptr;
    __nv_bfloat16* p2 = split ? reinterpret_cast<__nv_bfloat16*>(w + 3 * pbytes) : nullptr;
    if (in_layout == CF_LAYOUT_NCHW) {
        if (2L * B > 65535) return fail(CF_ERR_UNSUPPORTED, "corr build: batch %d > 32767", B);
        const dim3 grid((unsigned)cdiv(HW, 64), (unsigned)cdiv(C, 64), (unsigned)(2 * B));
        if (in_dtype == CF_DTYPE_F32)
            prep_transpose_kernel<float><<<grid, 256, 0, s>>>(reinterpret_cast<const float*>(f1), reinterpret_cast<const float*>(f2),
                                                              o1, o2, p1, p2, C, HW, 2);
        else
            prep_transpose_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(f1),
                                                                      reinterpret_cast<const __nv_bfloat16*>(f2), o1, o2, p1, p2, C, HW, 2);
        CF_CHECK_LAUNCH("prep_transpose_kernel");
    } else {   // fp32 channels-last
        const size_t n4 = (size_t)B * C * HW / 4;   // C % 8 == 0
        const dim3 grid((unsigned)min((size_t)4 * sm_count(), (n4 + 255) / 256), 2u);
        prep_convert_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const float*>(f1), reinterpret_cast<const float*>(f2), o1, o2, p1,
                                                 p2, n4);
        CF_CHECK_LAUNCH("prep_convert_kernel");
    }
    *h1 = o1; *h2 = o2; *l1 = p1; *l2 = p2;
    return CF_OK;
}

// one fp32 [B, C, HW] map -> bf16 [B, HW, C] (operands of the on-the-fly lookup, lookup_tc.cu)
int launch_prep_transpose_one(const float* src, void* dst_bf16, int B, int C, int HW, cudaStream_t s) {
    if (B > 65535) return fail(CF_ERR_UNSUPPORTED, "prep: batch %d > 65535", B);
    const dim3 grid((unsigned)cdiv(HW, 64), (unsigned)cdiv(C, 64), (unsigned)B);
    prep_transpose_kernel<float><<<grid, 256, 0, s>>>(src, src, reinterpret_cast<__nv_bfloat16*>(dst_bf16),
                                                      reinterpret_cast<__nv_bfloat16*>(dst_bf16), nullptr, nullptr, C, HW, 1);
    CF_CHECK_LAUNCH("prep_transpose_kernel");
    return CF_OK;
}

int launch_corr_tc2(const void* f1, const void* f2, int in_dtype, int in_layout, int B, int C, int H, int W,
                    float* const* levels, int L, bool split, void* workspace, size_t ws_bytes, int* levels_done,
                    bool allow_tiled, unsigned* tiled_mask, cudaStream_t s) {
    if (tiled_mask) *tiled_mask = 0;
    const int HW = H * W;
    if (in_dtype == CF_DTYPE_BF16) split = false;   // a bf16 map has no low part
    const __nv_bfloat16 *h1, *h2, *l1, *l2;
    int rc = corr_tc2_prepare(f1, f2, in_dtype, in_layout, B, C, HW, split, workspace, ws_bytes, &h1, &h2, &l1, &l2, s);
    if (rc) return rc;

    G2Params P{};
    P.level0 = levels[0];
    P.level1 = L > 1 ? levels[1] : nullptr;
    P.level2 = L > 2 ? levels[2] : nullptr;
    P.B = B; P.C = C; P.H = H; P.W = W; P.HW = HW;
    P.k_blocks = (C + BK - 1) / BK;
    P.qt_x = (W + 31) / 32;
    P.QT = P.qt_x * (H / 4);
    // Two 128-query sub-blocks per resident block (N = 256, plain bf16 only) halve the operand re-reads -- and measured SLOWER:
    // 78.4 vs 74.5 us at cfg2, 1784 vs 1552 us at cfg4 (the accumulator is held until the second half is read and the epilogue,
    // not the operand stream, is what the tile time follows).  Kept behind CF_TC2_NB=2 as the evidence; the default is one.
    static const int nb_knob = [] { const char* e = getenv("CF_TC2_NB"); return e ? atoi(e) : 1; }();   // tuning knob
    const int NB = (nb_knob == 2 && !split) ? 2 : 1;
    const int NP = NPS * NB;
    P.PB = (HW + NP - 1) / NP;
    P.units = (long)B * P.PB * P.QT;
    P.div = sqrtf((float)C);
    int e;
    P.mul = (frexpf(P.div, &e) == 0.5f) ? 1.0f / P.div : 0.0f;   // power of two: scaling by the reciprocal is exact
    *levels_done = P.level2 ? 3 : (P.level1 ? 2 : 1);
    // Levels as 4x4 tiles (fewer DRAM atoms per lookup window) pay for rows of >= 128 B; inference builds only.
    static const int tile_knob = [] { const char* e = getenv("CF_GEMM_TILED"); return e ? atoi(e) : 3; }();   // tuning knob
    const bool can_tile = allow_tiled && tiled_mask != nullptr;
    P.tile_l0 = (can_tile && (tile_knob & 1) && W >= 32) ? 1 : 0;                                   // H % 4 == 0, W % 8 == 0
    P.tile_l1 = (can_tile && (tile_knob & 2) && P.level1 && W >= 64 && H % 8 == 0) ? 1 : 0;
    if (tiled_mask) *tiled_mask = (P.tile_l0 ? 1u : 0u) | (P.tile_l1 ? 2u : 0u);

    { const char* e = getenv("CF_TC2_SKIP"); P.skip = e ? atoi(e) : 0; }
    // operands of the whole batch beyond a third of L2 (re-read by every query block): keep concurrent CTAs on neighbouring pairs
    P.sched = ((size_t)B * HW * C * 2 * (split ? 2 : 1) > (size_t)40 << 20 && (long)B * P.PB >= 4L * sm_count()) ? 1 : 0;
    { const char* e = getenv("CF_TC2_SCHED"); if (e) P.sched = atoi(e) != 0; }
    CUtensorMap tmQh, tmQl, tmPh, tmPl;
    const uint64_t rb = (uint64_t)C * 2;     // bytes of one position's channels
    {   // fmap2 as (channels, x, y, batch); box = 8 x of 4 rows
        const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
        const uint64_t str[3] = {rb, rb * W, rb * HW};
        const uint32_t box[4] = {BK, 8, 4, 1};
        if ((rc = make_tmap_bf16(&tmQh, h2, 4, dims, str, box))) return rc;
        tmQl = tmQh;
        if (split && (rc = make_tmap_bf16(&tmQl, l2, 4, dims, str, box))) return rc;
    }
    {   // fmap1 as (channels, position, batch)
        const uint64_t dims[3] = {(uint64_t)C, (uint64_t)HW, (uint64_t)B};
        const uint64_t str[2] = {rb, rb * HW};
        const uint32_t box[3] = {BK, NPS, 1};
        if ((rc = make_tmap_bf16(&tmPh, h1, 3, dims, str, box))) return rc;
        tmPl = tmPh;
        if (split && (rc = make_tmap_bf16(&tmPl, l1, 3, dims, str, box))) return rc;
    }
    if (P.units >= (1L << 31)) return fail(CF_ERR_UNSUPPORTED, "corr build: %ld tiles in one launch", P.units);
    const int grid = (int)min((long)sm_count(), P.units);
    // the opt-in to large dynamic shared memory belongs to the device: once per (device, kernel), sized for C = 256
    if (split) {
        if ((rc = smem_optin(reinterpret_cast<const void*>(&corr_tc2_kernel<true, 1>), Cfg<true, 1>::smem_bytes(MAX_KB), 0))) return rc;
        corr_tc2_kernel<true, 1><<<grid, NTHREADS, Cfg<true, 1>::smem_bytes(P.k_blocks), s>>>(tmQh, tmQl, tmPh, tmPl, P);
    } else if (NB == 2) {
        if ((rc = smem_optin(reinterpret_cast<const void*>(&corr_tc2_kernel<false, 2>), Cfg<false, 2>::smem_bytes(MAX_KB), 2))) return rc;
        corr_tc2_kernel<false, 2><<<grid, NTHREADS, Cfg<false, 2>::smem_bytes(P.k_blocks), s>>>(tmQh, tmQl, tmPh, tmPl, P);
    } else {
        if ((rc = smem_optin(reinterpret_cast<const void*>(&corr_tc2_kernel<false, 1>), Cfg<false, 1>::smem_bytes(MAX_KB), 1))) return rc;
        corr_tc2_kernel<false, 1><<<grid, NTHREADS, Cfg<false, 1>::smem_bytes(P.k_blocks), s>>>(tmQh, tmQl, tmPh, tmPl, P);
    }
    CF_CHECK_LAUNCH("corr_tc2_kernel");
    return CF_OK;
}

}  // namespace cf
