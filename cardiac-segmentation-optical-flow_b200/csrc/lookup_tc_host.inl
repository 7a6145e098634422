// Host side of lookup_tc.cu: operand buffer geometry, preparation (pooled fmap2 pyramid -> bf16 channels-last), launch.
namespace cf {

namespace {
struct OtfGeom {
    int L;
    int h[CF_MAX_LEVELS], w[CF_MAX_LEVELS];
    size_t off_f1, off_f2[CF_MAX_LEVELS], total;
};
bool otf_geom(int B, int C, int H, int W, int L, OtfGeom* g) {
    g->L = L;
    size_t off = 0;
    g->off_f1 = off;
    off += round_up((size_t)B * H * W * C * 2, 256);
    int h = H, w = W;
    for (int l = 0; l < L; ++l) {
        if (h < 1 || w < 1) return false;
        g->h[l] = h; g->w[l] = w;
        g->off_f2[l] = off;
        off += round_up((size_t)B * h * w * C * 2, 256);
        h /= 2; w /= 2;
    }
    g->total = off;
    return true;
}
}  // namespace

bool lookup_tc_supported(int B, int C, int H, int W, int L) {
    if (B < 1 || C < 8 || C % 8 != 0 || C > 64 * MAX_KB || H % QB_H != 0 || W % QB_W != 0 || L < 1 || L > CF_MAX_LEVELS) return false;
    if ((long)H * W > (1L << 24)) return false;
    return tmap_encode_fn() != nullptr;
}

size_t lookup_tc_operand_bytes(int B, int C, int H, int W, int L) {
    OtfGeom g;
    return otf_geom(B, C, H, W, L, &g) ? g.total : 0;
}

// workspace: the fp32 pooled pyramid of fmap2 ([B, C, h_l, w_l] per level, fmap_level_dims)
int launch_lookup_tc_prepare(const float* f1, const float* f2, int B, int C, int H, int W, int L, void* operands, void* workspace,
                             size_t ws_bytes, cudaStream_t s) {
    OtfGeom g;
    LevelDims d;
    if (!otf_geom(B, C, H, W, L, &g) || !fmap_level_dims(B, C, H, W, L, &d)) return fail(CF_ERR_UNSUPPORTED, "empty fmap pyramid level");
    if (!workspace || ws_bytes < d.off[L] || ((uintptr_t)workspace & 255u))
        return fail(CF_ERR_BAD_ARG, "on-the-fly prepare: workspace of %zu bytes (256-byte aligned) required, got %zu", d.off[L], ws_bytes);
    int rc = launch_fmap_pyramid(f2, B, C, H, W, L, reinterpret_cast<float*>(workspace), s);
    if (rc) return rc;
    char* op = reinterpret_cast<char*>(operands);
    if ((rc = launch_prep_transpose_one(f1, op + g.off_f1, B, C, H * W, s))) return rc;
    for (int l = 0; l < L; ++l) {
        const float* lv = reinterpret_cast<const float*>(reinterpret_cast<const char*>(workspace) + d.off[l]);
        if ((rc = launch_prep_transpose_one(lv, op + g.off_f2[l], B, C, g.h[l] * g.w[l], s))) return rc;
    }
    return CF_OK;
}

namespace {
template <int R>
int launch_lookup_tc_r(const OtfMaps& M, const OtfParams& P, cudaStream_t s) {
    const size_t smem = smem_bytes(P.k_blocks);
    int rc = smem_optin(reinterpret_cast<const void*>(&lookup_tc_kernel<R>), smem_bytes(MAX_KB), 4 + R);
    if (rc) return rc;
    const int grid = (int)std::min<long>((long)sm_count(), P.n_blocks);
    lookup_tc_kernel<R><<<grid, NTHREADS, smem, s>>>(M, P);
    CF_CHECK_LAUNCH("lookup_tc_kernel");
    return CF_OK;
}
}  // namespace

int launch_lookup_tc(const void* operands, const float* coords, int B, int C, int H, int W, int L, int r, float* out,
                     cudaStream_t s) {
    OtfGeom g;
    if (!otf_geom(B, C, H, W, L, &g)) return fail(CF_ERR_UNSUPPORTED, "empty fmap pyramid level");
    OtfParams P{};
    OtfMaps M;
    P.coords = coords; P.out = out;
    P.B = B; P.C = C; P.H = H; P.W = W; P.HW = H * W; P.L = L;
    P.k_blocks = (C + BK - 1) / BK;
    P.qb_x = W / QB_W;
    P.qb_per_img = P.qb_x * (H / QB_H);
    P.n_blocks = (long)B * P.qb_per_img;
    P.div = sqrtf((float)C);
    int e;
    P.mul = (frexpf(P.div, &e) == 0.5f) ? 1.0f / P.div : 0.0f;
    const char* op = reinterpret_cast<const char*>(operands);
    const uint64_t rb = (uint64_t)C * 2;
    int rc;
    {
        const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
        const uint64_t str[3] = {rb, rb * W, rb * H * W};
        const uint32_t box[4] = {BK, QB_W, QB_H, 1};
        if ((rc = make_tmap_bf16(&M.f1, op + g.off_f1, 4, dims, str, box))) return rc;
    }
    for (int l = 0; l < CF_MAX_LEVELS; ++l) {
        const int ll = l < L ? l : L - 1;           // unused slots repeat the last level (a valid descriptor)
        P.lh[l] = g.h[ll]; P.lw[l] = g.w[ll];
        P.rcp_h[l] = 1.0f / (float)(g.h[ll] - 1);
        P.rcp_w[l] = 1.0f / (float)(g.w[ll] - 1);
        const uint64_t dims[4] = {(uint64_t)C, (uint64_t)g.w[ll], (uint64_t)g.h[ll], (uint64_t)B};
        const uint64_t str[3] = {rb, rb * g.w[ll], rb * g.h[ll] * g.w[ll]};
        const uint32_t box[4] = {BK, CH_W, CH_H, 1};
        if ((rc = make_tmap_bf16(&M.f2[l], op + g.off_f2[ll], 4, dims, str, box))) return rc;
    }
    switch (r) {
        case 1: return launch_lookup_tc_r<1>(M, P, s);
        case 2: return launch_lookup_tc_r<2>(M, P, s);
        case 3: return launch_lookup_tc_r<3>(M, P, s);
        case 4: return launch_lookup_tc_r<4>(M, P, s);
    }
    return fail(CF_ERR_UNSUPPORTED, "lookup radius %d not in [1,%d]", r, CF_MAX_RADIUS);
}

}  // namespace cf
