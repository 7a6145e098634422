// Bring-up probe: one CTA, one tcgen05.mma kind::tf32 (M=128, N=128, K=8) over several smem layout variants.
// nvcc -gencode arch=compute_100a,code=sm_100a -o umma_probe umma_probe.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cmath>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct Variant { int a_major, b_major; int layout; uint32_t lbo, sbo; int sentinel; int kind_bf16; };

__global__ void probe(const uint8_t* a_img, const uint8_t* b_img, int a_bytes, int b_bytes, Variant v, float* d_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_slot;
    __shared__ __align__(8) uint64_t bar;
    uint8_t* sa = (uint8_t*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);
    uint8_t* sb = sa + 32768;
    for (int i = threadIdx.x; i < a_bytes / 4; i += blockDim.x) ((uint32_t*)sa)[i] = ((const uint32_t*)a_img)[i];
    for (int i = threadIdx.x; i < b_bytes / 4; i += blockDim.x) ((uint32_t*)sb)[i] = ((const uint32_t*)b_img)[i];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_u32(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the async proxy (UMMA)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    if (v.sentinel) {   // pre-fill the accumulator so "MMA wrote zeros" and "MMA wrote nothing" can be told apart
        for (int c = 0; c < 128; ++c) {
            uint32_t s = __float_as_uint(777.0f);
            asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr + c), "r"(s) : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    if (threadIdx.x == 0) {
        auto desc = [&](uint32_t addr) {
            return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((v.lbo >> 4) & 0x3FFFu) << 16) |
                   ((uint64_t)((v.sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | ((uint64_t)v.layout << 61);
        };
        const uint32_t fmt = v.kind_bf16 ? 1u : 2u;
        const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)v.a_major << 15) | ((uint32_t)v.b_major << 16) |
                               ((128u >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t ad = desc(smem_u32(sa)), bd = desc(smem_u32(sb));
        if (v.kind_bf16)
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                         ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(0u) : "memory");
        else
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                         ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(0u) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    uint32_t done = 0, spins = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        if (++spins > (1u << 26)) { if (threadIdx.x == 0) printf("TIMEOUT waiting for MMA\n"); break; }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int c = 0; c < 128; ++c) {
        uint32_t r;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr + c) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        d_out[(warp * 32 + lane) * 128 + c] = __uint_as_float(r);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem) : "memory");
}

static float tf32(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }

int main() {
    const int M = 128, N = 128, K = 8;
    std::vector<float> A(M * K), B(N * K);
    srand(1);
    for (auto& x : A) x = tf32((rand() % 2001 - 1000) / 500.0f);
    for (auto& x : B) x = tf32((rand() % 2001 - 1000) / 500.0f);
    std::vector<float> ref(M * N);
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { float s = 0; for (int k = 0; k < K; ++k) s += A[m * K + k] * B[n * K + k]; ref[m * N + n] = s; }

    struct Named { const char* name; Variant v; int kind; };
    // kind: 0 = MN-major SW128, 1 = MN-major no swizzle, 2 = K-major SW128, 3 = K-major no swizzle
    Named vs[] = {
        {"MN-major SW128   (LBO=1024 chunk stride, SBO=1024)", {1, 1, 2, 1024, 1024, 1, 0}, 0},
        {"MN-major SW128   swapped (LBO<->SBO roles)        ", {1, 1, 2, 4096, 1024, 1, 0}, 4},
        {"MN-major no-swz  (SBO=128 m-group, LBO=..)        ", {1, 1, 0, 4096, 128, 1, 0}, 1},
        {"K-major  SW128   (SBO=1024)                       ", {0, 0, 2, 0, 1024, 1, 0}, 2},
        {"K-major  no-swz  (LBO=128, SBO=256)               ", {0, 0, 0, 128, 256, 1, 0}, 3},
    };
    uint8_t *da, *db; float* dd;
    CK(cudaMalloc(&da, 32768)); CK(cudaMalloc(&db, 32768)); CK(cudaMalloc(&dd, M * N * 4));
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000));
    for (auto& nv : vs) {
        std::vector<uint8_t> ia(32768, 0), ib(32768, 0);
        auto place = [&](std::vector<uint8_t>& img, const std::vector<float>& X, int rows) {
            for (int m = 0; m < rows; ++m) for (int k = 0; k < K; ++k) {
                size_t off;
                if (nv.kind == 0 || nv.kind == 4) {            // chunks of 32 m; each chunk 8 k-rows x 128 B, 16B groups XOR (k%8)
                    int mc = m / 32, g = (m % 32) / 4, e = m % 4;
                    off = (size_t)mc * 1024 + k * 128 + ((g ^ (k % 8)) * 16) + e * 4;
                } else if (nv.kind == 1) {    // core matrix: 8 k x 16 B (4 m); m-groups stride SBO=128
                    off = (size_t)(m / 4) * 128 + k * 16 + (m % 4) * 4;
                } else if (nv.kind == 2) {    // rows m: 128 B each (32 k), groups of 8 rows = 1024 B; 16B group XOR (m%8)
                    int g = k / 4, e = k % 4;
                    off = (size_t)(m / 8) * 1024 + (m % 8) * 128 + ((g ^ (m % 8)) * 16) + e * 4;
                } else {                      // core matrix: 8 m x 16 B (4 k); k-groups stride LBO=128, m-groups stride SBO=256
                    off = (size_t)(m / 8) * 256 + (k / 4) * 128 + (m % 8) * 16 + (k % 4) * 4;
                }
                memcpy(&img[off], &X[m * K + k], 4);
            }
        };
        place(ia, A, M); place(ib, B, N);
        CK(cudaMemcpy(da, ia.data(), 32768, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(db, ib.data(), 32768, cudaMemcpyHostToDevice));
        CK(cudaMemset(dd, 0, M * N * 4));
        probe<<<1, 128, 70000>>>(da, db, 32768, 32768, nv.v, dd);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: CUDA error %s\n", nv.name, cudaGetErrorString(e)); return 1; }
        std::vector<float> out(M * N);
        CK(cudaMemcpy(out.data(), dd, M * N * 4, cudaMemcpyDeviceToHost));
        double maxerr = 0, maxabs = 0; int sent = 0, zeros = 0;
        for (int i = 0; i < M * N; ++i) { maxerr = fmax(maxerr, fabs(out[i] - ref[i])); maxabs = fmax(maxabs, fabs(out[i])); sent += out[i] == 777.0f; zeros += out[i] == 0.0f; }
        printf("%s: maxerr=%.4g out_absmax=%.4g sentinel_left=%d zeros=%d  out[0][0..3]=%g %g %g %g ref=%g %g %g %g\n", nv.name, maxerr, maxabs,
               sent, zeros, out[0], out[1], out[2], out[3], ref[0], ref[1], ref[2], ref[3]);
    }
    // ---- bf16 (kind::f16), K = 16 ----
    {
        const int K2 = 16;
        auto bf = [](float x) { uint32_t u; memcpy(&u, &x, 4); return (uint16_t)(u >> 16); };
        auto bff = [](uint16_t h) { uint32_t u = (uint32_t)h << 16; float x; memcpy(&x, &u, 4); return x; };
        std::vector<uint16_t> A2(M * K2), B2(N * K2);
        for (auto& x : A2) x = bf((rand() % 2001 - 1000) / 500.0f);
        for (auto& x : B2) x = bf((rand() % 2001 - 1000) / 500.0f);
        std::vector<float> ref2(M * N);
        for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { float s2 = 0; for (int k = 0; k < K2; ++k) s2 += bff(A2[m * K2 + k]) * bff(B2[n * K2 + k]); ref2[m * N + n] = s2; }
        Named vb[] = {
            {"bf16 MN-major SW128 (LBO=2048 chunk, SBO=1024 k-atom)", {1, 1, 2, 2048, 1024, 1, 1}, 0},
            {"bf16 K-major  SW128 (SBO=1024)                       ", {0, 0, 2, 0, 1024, 1, 1}, 2},
        };
        for (auto& nv : vb) {
            std::vector<uint8_t> ia(32768, 0), ib(32768, 0);
            auto place = [&](std::vector<uint8_t>& img, const std::vector<uint16_t>& X, int rows) {
                for (int m = 0; m < rows; ++m) for (int k = 0; k < K2; ++k) {
                    size_t off;
                    if (nv.kind == 0) { int mc = m / 64, g = (m % 64) / 8, e = m % 8; off = (size_t)mc * 2048 + (k / 8) * 1024 + (k % 8) * 128 + ((g ^ (k % 8)) * 16) + e * 2; }
                    else { off = (size_t)(m / 8) * 1024 + (m % 8) * 128 + (((k / 8) ^ (m % 8)) * 16) + (k % 8) * 2; }
                    memcpy(&img[off], &X[m * K2 + k], 2);
                }
            };
            place(ia, A2, M); place(ib, B2, N);
            CK(cudaMemcpy(da, ia.data(), 32768, cudaMemcpyHostToDevice));
            CK(cudaMemcpy(db, ib.data(), 32768, cudaMemcpyHostToDevice));
            CK(cudaMemset(dd, 0, M * N * 4));
            probe<<<1, 128, 70000>>>(da, db, 32768, 32768, nv.v, dd);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s: CUDA error %s\n", nv.name, cudaGetErrorString(e)); return 1; }
            std::vector<float> out(M * N);
            CK(cudaMemcpy(out.data(), dd, M * N * 4, cudaMemcpyDeviceToHost));
            double maxerr = 0, maxabs = 0; int zeros = 0;
            for (int i = 0; i < M * N; ++i) { maxerr = fmax(maxerr, fabs(out[i] - ref2[i])); maxabs = fmax(maxabs, fabs(out[i])); zeros += out[i] == 0.0f; }
            printf("%s: maxerr=%.4g out_absmax=%.4g zeros=%d out[0][0..1]=%g %g ref=%g %g\n", nv.name, maxerr, maxabs, zeros, out[0], out[1], ref2[0], ref2[1]);
        }
    }
    return 0;
}
