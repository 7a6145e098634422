// Host side of corr_bwd_tc.cu: workspace layout, tensor maps, launch.
namespace cf {

namespace {
inline size_t plane_bytes(int B, int C, int HW) { return (((size_t)B * C * HW * 2) + 255) & ~(size_t)255; }
}  // namespace

bool corr_bwd_tc_supported(int B, int C, int HW) {
    // TMA: the row pitch of the bf16 planes (HW * 2 bytes) must be a multiple of 16; the loaders read G in 8-value chunks
    return HW % 8 == 0 && C >= 1 && B <= 65535 && tmap_encode_fn() != nullptr;
}

size_t corr_bwd_tc_workspace_bytes(int B, int C, int HW) { return 4 * plane_bytes(B, C, HW); }

int launch_corr_volume_bwd_tc(const float* f1, const float* f2, const float* g, int B, int C, int HW, float* gf1, float* gf2,
                              void* ws, size_t ws_bytes, cudaStream_t s) {
    using namespace tc;
    if (!corr_bwd_tc_supported(B, C, HW)) return fail(CF_ERR_UNSUPPORTED, "corr_volume_bwd (tensor cores): HW %d not a multiple of 8", HW);
    if (!ws || ws_bytes < corr_bwd_tc_workspace_bytes(B, C, HW))
        return fail(CF_ERR_BAD_ARG, "corr_volume_bwd (tensor cores): workspace %zu < %zu bytes", ws_bytes, corr_bwd_tc_workspace_bytes(B, C, HW));
    if (!aligned16(f1) || !aligned16(f2) || !aligned16(g) || (reinterpret_cast<uintptr_t>(ws) & 255u))
        return fail(CF_ERR_BAD_ARG, "corr_volume_bwd (tensor cores): pointers must be 16-byte aligned (workspace: 256)");
    const size_t pb = plane_bytes(B, C, HW);
    char* w = static_cast<char*>(ws);
    __nv_bfloat16* planes[2][2] = {{reinterpret_cast<__nv_bfloat16*>(w), reinterpret_cast<__nv_bfloat16*>(w + pb)},
                                   {reinterpret_cast<__nv_bfloat16*>(w + 2 * pb), reinterpret_cast<__nv_bfloat16*>(w + 3 * pb)}};
    const float* maps[2] = {f1, f2};
    const size_t n8 = (size_t)B * C * HW / 8;      // HW % 8 == 0
    const int need[2] = {gf2 != nullptr, gf1 != nullptr};        // fmap1 feeds gf2, fmap2 feeds gf1
    for (int i = 0; i < 2; ++i) {
        if (!need[i]) continue;
        const unsigned grid = (unsigned)std::min<size_t>((n8 + 255) / 256, (size_t)sm_count() * 16);
        split_planes_kernel<<<grid, 256, 0, s>>>(maps[i], planes[i][0], planes[i][1], n8);
        CF_CHECK_LAUNCH("split_planes_kernel");
    }

    BwdMaps M;
    const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)C, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)HW * 2, (uint64_t)C * HW * 2};
    const uint32_t box[3] = {(uint32_t)BK, (uint32_t)BN, 1u};
    for (int gemm = 0; gemm < 2; ++gemm)
        for (int part = 0; part < 2; ++part) {
            // gemm 0 (gf1) contracts over q with fmap2; gemm 1 (gf2) contracts over p with fmap1
            const int which = gemm == 0 ? 1 : 0;
            if (int rc = make_tmap_bf16(&M.f[gemm][part], planes[which][part], 3, dims, strides, box)) return rc;
        }

    BwdParams P{};
    P.g = g;
    P.out[0] = gf1;
    P.out[1] = gf2;
    P.B = B;
    P.C = C;
    P.HW = HW;
    P.k_blocks = cdiv(HW, BK);
    P.m_tiles = cdiv(HW, BM);
    P.n_tiles = cdiv(C, BN);
    P.div = sqrtf((float)C);
    int slots = 0;
    if (gf1) P.gemm_of[slots++] = 0;
    if (gf2) P.gemm_of[slots++] = 1;
    const long gx = (long)slots * P.m_tiles * P.n_tiles;
    if (gx > 0x7fffffffL) return fail(CF_ERR_UNSUPPORTED, "corr_volume_bwd (tensor cores): grid too large");
    if (int rc = smem_optin(reinterpret_cast<const void*>(&corr_bwd_tc_kernel), kSmemBytes, 12)) return rc;
    corr_bwd_tc_kernel<<<dim3((unsigned)gx, (unsigned)B), NTHREADS, kSmemBytes, s>>>(M, P);
    CF_CHECK_LAUNCH("corr_bwd_tc_kernel");
    return CF_OK;
}

}  // namespace cf
