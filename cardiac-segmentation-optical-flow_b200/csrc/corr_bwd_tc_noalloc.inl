// Allocation-free variant of the tensor-core correlation-volume backward (cf_corr_volume_bwd_noalloc; no counterpart in the
// reference: added for callers that cannot call cudaMalloc -- code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).  Same kernels and the same host path as launch_corr_volume_bwd_tc;
// the only difference is where the four bf16 hi / lo planes of the feature maps live: in the static device array below
// instead of a caller-provided workspace.
//
//   * every GPU that loads the library holds these CF_NOALLOC_SCRATCH_BYTES (8 MiB) until the process exits;
//   * problems whose planes do not fit (4 planes of B*C*H*W bf16 values, each rounded up to 256 bytes) are rejected;
//   * every call on a device shares the array and finds the previous call's planes in it.  Nothing in it has to start at zero:
//     the pre-pass overwrites the part of every plane that this call's kernel reads.  For the same reason two calls that
//     OVERLAP on the GPU (different streams, or graphs replayed concurrently) race on it: keep the calls of one device in order;
//   * `g_noalloc_scratch` names a device symbol: in host code it is a stand-in, not a device address, so the address a kernel
//     may use comes from cudaGetSymbolAddress, asked once per device with that device current, and is checked.
#include <mutex>

namespace cf {
namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[CF_NOALLOC_SCRATCH_BYTES];

int noalloc_scratch_ptr(void** out) {
    static std::mutex mu;
    static void* ptr[64] = {};
    int dev = 0;
    CF_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(CF_ERR_UNSUPPORTED, "corr_volume_bwd_noalloc: device index %d not in [0, 64)", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void* p = nullptr;
        CF_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));      // the current device's instance of the array
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u))
            return fail(CF_ERR_CUDA, "corr_volume_bwd_noalloc: cudaGetSymbolAddress returned %p (null or not 256-byte aligned)", p);
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return CF_OK;
}

}  // namespace

int launch_corr_volume_bwd_tc_noalloc(const float* f1, const float* f2, const float* g, int B, int C, int HW, float* gf1, float* gf2,
                                      cudaStream_t s) {
    if (!corr_bwd_tc_supported(B, C, HW))
        return fail(CF_ERR_UNSUPPORTED, "corr_volume_bwd_noalloc: needs H*W %% 8 == 0 (tensor-core path); got H*W = %d", HW);
    const size_t need = corr_bwd_tc_workspace_bytes(B, C, HW);
    if (need > (size_t)CF_NOALLOC_SCRATCH_BYTES)
        return fail(CF_ERR_UNSUPPORTED,
                    "corr_volume_bwd_noalloc: B=%d C=%d H*W=%d needs %zu bytes of scratch, the static scratch array holds %zu "
                    "(CF_NOALLOC_SCRATCH_BYTES); use cf_corr_volume_bwd_ex with a workspace",
                    B, C, HW, need, (size_t)CF_NOALLOC_SCRATCH_BYTES);
    void* ws = nullptr;
    if (int rc = noalloc_scratch_ptr(&ws)) return rc;
    return launch_corr_volume_bwd_tc(f1, f2, g, B, C, HW, gf1, gf2, ws, (size_t)CF_NOALLOC_SCRATCH_BYTES, s);
}

}  // namespace cf
