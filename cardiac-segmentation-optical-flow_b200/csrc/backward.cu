// backward of the correlation block (the reference trains through it:
// `l.mean().backward()`, /root/reference/nnunet/training/network_training/SegFlowGaussian.py:2225, with the
// CorrBlock built on encoder features, nnunet/network_architecture/SegFlowGaussian.py:929-935).
//
//   cf_corr_lookup_bwd   d(lookup)/d(pyramid): every (query, level) owns a private h_l x w_l image, so the 81
//                        bilinear taps of a query scatter into a window that NO other thread touches -> the
//                        window is accumulated in shared memory and ADDED to the gradient pyramid with one reduction
//                        per touched cell (deterministic: one addition per cell and launch): the lookups of one
//                        CorrBlock share one gradient buffer.
//                        No gradient w.r.t. coords: the reference detaches them (SegFlowGaussian.py:933).
//   cf_corr_pool_bwd     backward of the avg_pool2d chain, coarse -> fine, in place: g_l += up2(g_{l+1}) / 4.
//   cf_corr_volume_bwd   d(corr)/d(fmaps): gf1 = f2 . G^T / sqrt(C), gf2 = f1 . G / sqrt(C) with G = grad of
//                        level 0 [HW_p, HW_q]; fp32 CUDA-core GEMMs (training in the reference is fp32).
#include "common.cuh"
#include "sampling.cuh"

namespace cf {
namespace {

constexpr int kQ = 32;   // queries per CTA (lane <-> query)

// *p += v unless v == 0 (cells no tap touched, and cells outside the image, which the caller passes as 0), as ONE predicated reduction
__device__ __forceinline__ void red_add_if_nonzero(float* p, float v) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "setp.neu.f32 q, %1, 0f00000000;\n\t"
        "@q red.global.add.f32 [%0], %1;\n\t}"
        ::"l"(p), "f"(v) : "memory");
}

struct LookupBwdParams {
    float* glevel[CF_MAX_LEVELS];
    int lh[CF_MAX_LEVELS];
    int lw[CF_MAX_LEVELS];
    const float* coords;
    const float* go;
    long N;
    int HW;
    int L;
};

// CTA = 32 consecutive queries x one level, one warp.  Window = (2r+4)^2 cells anchored at floor(c) - r - 1
// (nominal taps use cells 1..2r+2; a flipped floor can reach cell 0 or 2r+3).
template <int R>
__global__ void __launch_bounds__(32) lookup_bwd_kernel(const __grid_constant__ LookupBwdParams P) {
    constexpr int SIDE = 2 * R + 1, PR = 2 * R + 4, CELLS = PR * PR, K = SIDE * SIDE;
    constexpr int PITCH = CELLS + 1;   // odd pitch: lane-private windows, conflict-free accumulation
    __shared__ float win[kQ * PITCH];
    __shared__ int anchor[kQ][2];

    const int lane = threadIdx.x;
    const unsigned task = blockIdx.x / (unsigned)P.L;
    const int l = (int)(blockIdx.x - task * (unsigned)P.L);
    const long n0 = (long)task * kQ;
    const long n = n0 + lane;
    const bool valid = n < P.N;
    const int hl = P.lh[l], wl = P.lw[l];

    long b = 0;
    int p = 0;
    float cx = 0.0f, cy = 0.0f;
    if (valid) {
        b = n / P.HW;
        p = (int)(n - b * P.HW);
        const float inv = 1.0f / (float)(1 << l);
        cx = __fmul_rn(__ldg(P.coords + (b * 2 + 0) * P.HW + p), inv);
        cy = __fmul_rn(__ldg(P.coords + (b * 2 + 1) * P.HW + p), inv);
    }
    const bool dead = !valid || !(fabsf(cx) < kDeadCoord) || !(fabsf(cy) < kDeadCoord);
    const int ax = dead ? 0 : __float2int_rd(cx) - R - 1;
    const int ay = dead ? 0 : __float2int_rd(cy) - R - 1;
    anchor[lane][0] = dead ? (1 << 30) : ax;   // far outside any image: nothing is written back
    anchor[lane][1] = ay;
    float* my = win + lane * PITCH;
    // taps of both axes up front: they decide between the register path (every tap of every lane is the nominal one: x tap i at
    // window column i + 1, y tap j at window row j + 1 -- all but queries within ~1e-5 of an integer coordinate) and the general path
    int xo[SIDE], yo[SIDE];
    float xl[SIDE], xh[SIDE], yl[SIDE], yh[SIDE];
    bool nominal = true;
    {
        const float s1x = (float)(wl - 1), s1y = (float)(hl - 1);
        const float rcx = __frcp_rn(s1x), rcy = __frcp_rn(s1y);   // IEEE-rounded reciprocals for div_small
        const float ccx = dead ? 0.0f : cx, ccy = dead ? 0.0f : cy;
#pragma unroll
        for (int i = 0; i < SIDE; ++i) {
            int i0, j0;
            axis_tap(ccx, i - R, s1x, rcx, i0, xl[i], xh[i]);
            axis_tap(ccy, i - R, s1y, rcy, j0, yl[i], yh[i]);
            xo[i] = i0 - ax;   // in [0, PR-2] by the round-trip bound (lookup.cu: classify)
            yo[i] = j0 - ay;
            nominal = nominal && xo[i] == i + 1 && yo[i] == i + 1;
        }
    }
    const float* g = P.go + ((size_t)b * P.L * K + (size_t)l * K) * P.HW + p;
    const bool fastw = __all_sync(0xffffffffu, dead || nominal);
    if (fastw) {
        // register path: window row r receives output row j = r - 1 through its upper weights and j = r - 2 through its lower ones; a row
        // is accumulated in 12 registers with static indices (no shared-memory read-modify-write chain: the general path below spends
        // 324 dependent LDS / FFMA / STS round trips per query) and stored once.  Rows / columns 0 and PR - 1 stay untouched and unread.
        float gprev[SIDE], gcur[SIDE];
#pragma unroll
        for (int i = 0; i < SIDE; ++i) gprev[i] = gcur[i] = 0.0f;
#pragma unroll
        for (int r = 1; r <= SIDE + 1; ++r) {
            float acc[PR];
#pragma unroll
            for (int c = 0; c < PR; ++c) acc[c] = 0.0f;
            if (r - 1 < SIDE) {
#pragma unroll
                for (int i = 0; i < SIDE; ++i) gcur[i] = dead ? 0.0f : __ldg(g + (size_t)(i * SIDE + (r - 1)) * P.HW);   // coalesced across lanes
#pragma unroll
                for (int i = 0; i < SIDE; ++i) {
                    acc[i + 1] = fmaf(gcur[i], xl[i] * yl[r - 1], acc[i + 1]);
                    acc[i + 2] = fmaf(gcur[i], xh[i] * yl[r - 1], acc[i + 2]);
                }
            }
            if (r - 2 >= 0) {
#pragma unroll
                for (int i = 0; i < SIDE; ++i) {
                    acc[i + 1] = fmaf(gprev[i], xl[i] * yh[r - 2], acc[i + 1]);
                    acc[i + 2] = fmaf(gprev[i], xh[i] * yh[r - 2], acc[i + 2]);
                }
            }
#pragma unroll
            for (int i = 0; i < SIDE; ++i) gprev[i] = gcur[i];
#pragma unroll
            for (int c = 1; c <= SIDE + 1; ++c) my[r * PR + c] = acc[c];
        }
        __syncwarp();
        // cooperative write-back of the (2r+2)^2 touched cells: a lane's cells are the same for every query, so their window coordinates
        // are computed once
        constexpr int TS = SIDE + 1, TCELLS = TS * TS, NIT = (TCELLS + 31) / 32;
        int wyk[NIT], wxk[NIT], wofs[NIT], gofs[NIT];    // window row / column, offset inside the window, offset inside the level image
#pragma unroll
        for (int k = 0; k < NIT; ++k) {
            const int c = lane + 32 * k;
            wyk[k] = c < TCELLS ? 1 + c / TS : -(1 << 20);   // a lane without a cell fails the row test below
            wxk[k] = 1 + c % TS;
            wofs[k] = (1 + c / TS) * PR + wxk[k];
            gofs[k] = (1 + c / TS) * wl + wxk[k];
        }
        const int nq = (int)min((long)kQ, P.N - n0);
        for (int q = 0; q < nq; ++q) {
            const int qx = anchor[q][0], qy = anchor[q][1];
            if (qx == (1 << 30)) continue;
            float* img = P.glevel[l] + (size_t)(n0 + q) * hl * wl;
            const float* srcw = win + q * PITCH;
            const int qbase = qy * wl + qx;
#pragma unroll
            for (int k = 0; k < NIT; ++k) {
                // branch-free: a predicated shared load and a predicated reduction (as `if`s each cell cost a divergence region of
                // its own: BSSY / branch / BSYNC -- 5000 instructions per warp in a kernel that runs at its issue latency)
                const bool in = (unsigned)(qy + wyk[k]) < (unsigned)hl && (unsigned)(qx + wxk[k]) < (unsigned)wl;
                const float v = in ? srcw[wofs[k]] : 0.0f;
                red_add_if_nonzero(img + (in ? qbase + gofs[k] : 0), v);     // one addition per cell and launch: deterministic (see below)
            }
        }
        return;
    }
    for (int i = 0; i < CELLS; ++i) my[i] = 0.0f;

    if (!dead) {
#pragma unroll 1
        for (int j = 0; j < SIDE; ++j) {
            float* rowp = my + min(max(yo[j], 0), PR - 2) * PR;
#pragma unroll
            for (int i = 0; i < SIDE; ++i) {
                const float gv = __ldg(g + (size_t)(i * SIDE + j) * P.HW);   // coalesced across lanes
                float* c = rowp + min(max(xo[i], 0), PR - 2);
                c[0] += gv * (xl[i] * yl[j]);
                c[1] += gv * (xh[i] * yl[j]);
                c[PR] += gv * (xl[i] * yh[j]);
                c[PR + 1] += gv * (xh[i] * yh[j]);
            }
        }
    }
    __syncwarp();
    // cooperative write-back: for each query the 32 lanes sweep its window cells (rows of PR contiguous floats)
    const int nq = (int)min((long)kQ, P.N - n0);
    for (int q = 0; q < nq; ++q) {
        const int qx = anchor[q][0], qy = anchor[q][1];
        if (qx == (1 << 30)) continue;
        float* img = P.glevel[l] + (size_t)(n0 + q) * hl * wl;
        const float* src = win + q * PITCH;
        for (int c = lane; c < CELLS; c += 32) {
            const int wy = c / PR, wx = c - wy * PR;
            const int y = qy + wy, x = qx + wx;
            // accumulate: the 12 lookups of one CorrBlock add into ONE gradient pyramid.  Launches are stream-ordered and a
            // (query, level) image belongs to one warp, so every cell receives exactly one addition per launch: a reduction
            // (RED.ADD.F32, no return value) gives the same bits as load-add-store without the 160 dependent global round
            // trips per warp that bounded this kernel (220 us per cfg2 lookup at 5 % of HBM).  Cells no tap touched are skipped.
            const float v = src[c];
            if (v != 0.0f && (unsigned)y < (unsigned)hl && (unsigned)x < (unsigned)wl) atomicAdd(&img[(size_t)y * wl + x], v);
        }
    }
}

// g_fine[n, y, x] += 0.25 * g_coarse[n, y/2, x/2] for the cells an avg_pool2d(2, 2) window covers (floor mode)
__global__ void __launch_bounds__(256) pool_bwd_kernel(float* __restrict__ fine, const float* __restrict__ coarse,
                                                       int h, int w, long total) {
    const long t = (long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int h2 = h >> 1, w2 = w >> 1;
    const int x = (int)(t % w);
    const long r = t / w;
    const int y = (int)(r % h);
    const long n = r / h;
    const int y2 = y >> 1, x2 = x >> 1;
    if (y2 < h2 && x2 < w2) fine[t] = __fadd_rn(fine[t], __fdiv_rn(__ldg(coarse + ((size_t)n * h2 + y2) * w2 + x2), 4.0f));
}

// the same for level widths that are multiples of 4: one thread = 4 consecutive cells of a fine row (16-byte accesses, two coarse
// cells), 32-bit index arithmetic (the scalar kernel spends most of its instructions on three 64-bit divisions)
__global__ void __launch_bounds__(256) pool_bwd_vec4_kernel(float4* __restrict__ fine, const float2* __restrict__ coarse, int h,
                                                            int w4, unsigned total4) {
    const unsigned t = blockIdx.x * 256u + threadIdx.x;
    if (t >= total4) return;
    const unsigned x4 = t % (unsigned)w4, r = t / (unsigned)w4;
    const unsigned y = r % (unsigned)h, n = r / (unsigned)h;
    const unsigned h2 = (unsigned)h >> 1, y2 = y >> 1;
    if (y2 >= h2) return;                        // odd last row: outside every pooling window (floor mode)
    const float2 c = __ldg(coarse + ((size_t)n * h2 + y2) * (unsigned)w4 + x4);
    const float a = __fdiv_rn(c.x, 4.0f), b = __fdiv_rn(c.y, 4.0f);
    float4 f = fine[t];
    f.x = __fadd_rn(f.x, a);
    f.y = __fadd_rn(f.y, a);
    f.z = __fadd_rn(f.z, b);
    f.w = __fadd_rn(f.w, b);
    fine[t] = f;
}

// C[m, n] = alpha * sum_k A[m, k] * B(k, n);  A row-major [M, K] (k contiguous).
// BT = false: B row-major [K, N] (n contiguous);  BT = true: B stored as [N, K] (k contiguous).
constexpr int TM = 64, TN = 64, TK = 16;
template <bool BT>
__global__ void __launch_bounds__(256) sgemm_kernel(const float* __restrict__ A, const float* __restrict__ Bm,
                                                    float* __restrict__ Cm, int M, int N, int K, float div,
                                                    size_t strideA, size_t strideB, size_t strideC) {
    __shared__ float As[TK][TM + 1];
    __shared__ float Bs[TK][TN + (BT ? 1 : 0)];
    const int bz = blockIdx.z;
    A += bz * strideA;
    Bm += bz * strideB;
    Cm += bz * strideC;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    const int tid = threadIdx.x;
    const int tx = tid % 16, ty = tid / 16;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += TK) {
        {   // A tile: 64 rows x 16 k, k contiguous in memory -> lanes along k
            const int kk = tid % TK, mm = tid / TK;   // 16 rows per pass
#pragma unroll
            for (int r = 0; r < TM; r += 16) {
                const int m = m0 + mm + r, k = k0 + kk;
                As[kk][mm + r] = (m < M && k < K) ? __ldg(A + (size_t)m * K + k) : 0.0f;
            }
        }
        if (BT) {
            const int kk = tid % TK, nn = tid / TK;
#pragma unroll
            for (int r = 0; r < TN; r += 16) {
                const int n = n0 + nn + r, k = k0 + kk;
                Bs[kk][nn + r] = (n < N && k < K) ? __ldg(Bm + (size_t)n * K + k) : 0.0f;
            }
        } else {
            const int nn = tid % TN, kk = tid / TN;   // 4 k-rows per pass
#pragma unroll
            for (int r = 0; r < TK; r += 4) {
                const int n = n0 + nn, k = k0 + kk + r;
                Bs[kk + r][nn] = (n < N && k < K) ? __ldg(Bm + (size_t)k * N + n) : 0.0f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < TK; ++k) {
            float av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = As[k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = Bs[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n < N) Cm[(size_t)m * N + n] = __fdiv_rn(acc[i][j], div);
        }
    }
}

template <int R>
int launch_lookup_bwd_r(const LookupBwdParams& P, cudaStream_t s) {
    const long grid = (P.N + kQ - 1) / kQ * P.L;
    if (grid > 0x7fffffffL) return fail(CF_ERR_UNSUPPORTED, "lookup_bwd grid too large (%ld)", grid);
    lookup_bwd_kernel<R><<<(unsigned)grid, 32, 0, s>>>(P);
    CF_CHECK_LAUNCH("lookup_bwd_kernel");
    return CF_OK;
}

}  // namespace

int launch_lookup_bwd(const LevelDims& d, const float* coords, const float* go, int B, int H, int W, int L, int r,
                      float* gpyr, cudaStream_t s) {
    LookupBwdParams P{};
    for (int l = 0; l < L; ++l) {
        P.glevel[l] = reinterpret_cast<float*>(reinterpret_cast<char*>(gpyr) + d.off[l]);
        P.lh[l] = d.h[l];
        P.lw[l] = d.w[l];
    }
    P.coords = coords;
    P.go = go;
    P.N = (long)B * H * W;
    P.HW = H * W;
    P.L = L;
    switch (r) {
        case 1: return launch_lookup_bwd_r<1>(P, s);
        case 2: return launch_lookup_bwd_r<2>(P, s);
        case 3: return launch_lookup_bwd_r<3>(P, s);
        case 4: return launch_lookup_bwd_r<4>(P, s);
    }
    return fail(CF_ERR_UNSUPPORTED, "lookup radius %d not in [1,%d]", r, CF_MAX_RADIUS);
}

int launch_pool_chain_bwd(float* gpyr, const LevelDims& d, long n_img, int L, cudaStream_t s) {
    char* base = reinterpret_cast<char*>(gpyr);
    for (int l = L - 2; l >= 0; --l) {
        float* fine = reinterpret_cast<float*>(base + d.off[l]);
        const float* coarse = reinterpret_cast<const float*>(base + d.off[l + 1]);
        const long total = n_img * d.h[l] * d.w[l];
        if (d.w[l] % 4 == 0 && total / 4 < 0xffffffffL && aligned16(fine) && (reinterpret_cast<uintptr_t>(coarse) & 7u) == 0) {
            const unsigned total4 = (unsigned)(total / 4);
            pool_bwd_vec4_kernel<<<cdiv((long)total4, 256), 256, 0, s>>>(reinterpret_cast<float4*>(fine), reinterpret_cast<const float2*>(coarse),
                                                                         d.h[l], d.w[l] / 4, total4);
        } else {
            pool_bwd_kernel<<<cdiv(total, 256), 256, 0, s>>>(fine, coarse, d.h[l], d.w[l], total);
        }
        CF_CHECK_LAUNCH("pool_bwd_kernel");
    }
    return CF_OK;
}

int launch_corr_volume_bwd(const float* f1, const float* f2, const float* gcorr, int B, int C, int HW, float* gf1,
                           float* gf2, cudaStream_t s) {
    if (B > 65535) return fail(CF_ERR_UNSUPPORTED, "corr_volume_bwd: batch %d > 65535", B);
    const float div = sqrtf((float)C);
    const size_t sf = (size_t)C * HW, sg = (size_t)HW * HW;
    dim3 grid(cdiv(HW, TN), cdiv(C, TM), B);
    // gf1[c, p] = sum_q f2[c, q] * G[p, q]   (B operand stored [N = p, K = q])
    if (gf1) sgemm_kernel<true><<<grid, 256, 0, s>>>(f2, gcorr, gf1, C, HW, HW, div, sf, sg, sf);
    // gf2[c, q] = sum_p f1[c, p] * G[p, q]   (B operand row-major [K = p, N = q])
    if (gf2) sgemm_kernel<false><<<grid, 256, 0, s>>>(f1, gcorr, gf2, C, HW, HW, div, sf, sg, sf);
    CF_CHECK_LAUNCH("sgemm_kernel");
    return CF_OK;
}

}  // namespace cf
