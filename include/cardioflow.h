/* cardioflow.h -- C ABI of libcardioflow.so: the B200 (sm_100a) optical-flow hot path.
 *
 * This is the drop-in boundary for ONE path of nicolas1805961/Cardiac-Segmentation-Optical-flow: the RAFT
 * correlation block (all-pairs correlation -> avg-pooled pyramid -> radius-r bilinear lookup) and the
 * voxelmorph-style dense warp.  The reference has no FFI of its own (it is pure Python on torch); each entry
 * point below names the reference interface it replaces.  Python host code binds these with ctypes
 * (cardiac-segmentation-optical-flow_b200/_cabi.py); INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into caller-owned memory (torch storage); the library never
 *     allocates or frees caller-visible memory;
 *   - all tensors are dense row-major ("contiguous NCHW"), float32 unless said otherwise;
 *   - `stream` is a cudaStream_t cast to void*; every call only ENQUEUES work on it (no host sync);
 *   - return value: 0 = ok, CF_ERR_* < 0 otherwise; cf_last_error() gives a thread-local message;
 *   - there is no CPU fallback: without a CUDA device every compute entry point returns CF_ERR_CUDA.
 */
#ifndef CARDIOFLOW_H_
#define CARDIOFLOW_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CF_API __attribute__((visibility("default")))
#else
#define CF_API
#endif

#define CF_OK 0
#define CF_ERR_BAD_ARG (-1)     /* null pointer, non-positive size, misaligned pointer            */
#define CF_ERR_UNSUPPORTED (-2) /* shape / radius / level count outside what the kernels implement */
#define CF_ERR_CUDA (-3)        /* CUDA runtime / driver error (message in cf_last_error)          */

#define CF_MAX_LEVELS 8
#define CF_MAX_RADIUS 4

/* cf_corr_build precision */
#define CF_PREC_BF16 0 /* tcgen05.mma kind::f16: feature maps rounded to bf16, fp32 accumulate in TMEM          */
#define CF_PREC_FP32 1 /* CUDA-core fp32 FMA (exact-precision mode; also serves shapes the TMA path rejects) */
#define CF_PREC_AUTO 2 /* fp32-grade: BF16X3 on the tensor cores when the shape allows it, FP32 otherwise       */
#define CF_PREC_BF16X3 3 /* tcgen05.mma on hi/lo bf16 splits: hi*hi + hi*lo + lo*hi, fp32 accumulate (~1e-6)   */

/* cf_corr_build_ex input element type / memory layout of the feature maps */
#define CF_DTYPE_F32 0
#define CF_DTYPE_BF16 1
#define CF_LAYOUT_NCHW 0 /* [B, C, H, W] contiguous (what the reference's encoders emit)                        */
#define CF_LAYOUT_NHWC 1 /* [B, H, W, C] contiguous = torch.channels_last storage; already K-major for the MMA  */

/* cf_warp2d interpolation */
#define CF_MODE_BILINEAR 0
#define CF_MODE_NEAREST 1

CF_API int cf_version(void);
CF_API const char* cf_last_error(void);

/* ---- pyramid geometry -------------------------------------------------------------------------------
 * The pyramid of one CorrBlock is ONE packed buffer.  Level l is the dense tensor [B*H*W, 1, h_l, w_l]
 * (exactly torchvision's / RAFT's corr_pyramid[l] layout) with h_l = floor(h_{l-1} / 2); level offsets
 * are rounded up to 256 bytes.  (reference: raft.py:385-392 `corr_pyramid`) */
CF_API int cf_pyramid_level_dims(int H, int W, int level, int* h_out, int* w_out);
CF_API size_t cf_corr_pyramid_level_offset(int B, int H, int W, int level); /* byte offset of level `level` */
CF_API size_t cf_corr_pyramid_bytes(int B, int H, int W, int L);            /* bytes for levels 0..L-1      */

/* ---- a1: coords_grid(batch, ht, wd, device) ------------------------------------------------------------
 * replaces `coords_grid` imported at nnunet/network_architecture/SegFlowGaussian.py:42, used :839-840
 * (torchvision twin: _utils.py:22-26).  out [B,2,H,W]: channel 0 = x (column), channel 1 = y (row). */
CF_API int cf_coords_grid(float* out, int B, int H, int W, void* stream);

/* ---- a2 + a3: CorrBlock.__init__ -------------------------------------------------------------------------
 * replaces `CorrBlock(fmap1, fmap2, radius=4)` (SegFlowGaussian.py:929 and 9 more call sites) =
 * torchvision raft.py:424-431 (`matmul(f1^T, f2) / sqrt(C)`) + raft.py:387-392 (3x avg_pool2d).
 * f1, f2 [B,C,H,W] fp32.  Writes all L levels into `pyramid` (cf_corr_pyramid_bytes).
 * `workspace` (256-byte aligned, cf_corr_build_workspace_bytes; may be NULL when that is 0) holds the bf16
 * copies of the feature maps for the tensor-core path.
 * `used_precision` (optional) receives the precision actually used (CF_PREC_BF16 / CF_PREC_FP32 / CF_PREC_BF16X3). */
CF_API size_t cf_corr_build_workspace_bytes(int B, int C, int H, int W, int precision);
CF_API int cf_corr_build(const float* f1, const float* f2, int B, int C, int H, int W, int L, int precision,
                  float* pyramid, void* workspace, size_t workspace_bytes, int* used_precision, void* stream);

/* Layout-aware variants.  With CF_BUILD_ALLOW_TILED the builder may store pyramid levels as 4x4 tiles (CF_LAYOUT_TILED4:
 * element (y, x) of a query's h_l x w_l image at ((y/4) * (w_l/4) + x/4) * 16 + (y%4) * 4 + x%4; same level offsets and
 * sizes) -- a lookup window then touches ~10 instead of ~16 DRAM atoms.  `tiled_mask` receives bit l = 1 for every level
 * written that way (0 = torchvision's row-major layout); pass it unchanged to cf_corr_lookup_layout.  Every other entry
 * point (pooling, backward passes) works on row-major pyramids only: keep flags = 0 for pyramids they will see. */
#define CF_BUILD_ALLOW_TILED 1
CF_API int cf_corr_build_layout(const float* f1, const float* f2, int B, int C, int H, int W, int L, int precision, int flags,
                         float* pyramid, void* workspace, size_t workspace_bytes, int* used_precision,
                         unsigned* tiled_mask, void* stream);
/* Full form.  `in_dtype` / `in_layout` describe f1 and f2 (CF_DTYPE_*, CF_LAYOUT_*): bf16 channels-last maps are read by
 * the TMA directly (no pre-pass, no workspace); every other combination goes through one conversion pass into `workspace`
 * (cf_corr_build_workspace_bytes).  The reference computes this volume in fp32 (RAFT casts the maps with .float();
 * SegFlowGaussian.py:2909 runs without autocast), so CF_PREC_AUTO selects the fp32-grade tensor-core mode CF_PREC_BF16X3;
 * plain CF_PREC_BF16 (2-3e-3 relative error) is opt-in.  A bf16 input has no low part: BF16X3 degrades to BF16 and
 * `used_precision` says so. */
CF_API int cf_corr_build_ex(const void* f1, const void* f2, int in_dtype, int in_layout, int B, int C, int H, int W, int L,
                     int precision, int flags, float* pyramid, void* workspace, size_t workspace_bytes,
                     int* used_precision, unsigned* tiled_mask, void* stream);
CF_API int cf_corr_lookup_layout(const float* pyramid, const float* coords, int B, int H, int W, int L, int r,
                          unsigned tiled_mask, float* out, void* stream);

/* Rebuild levels from_level+1 .. L-1 from level from_level (F.avg_pool2d(x, 2, stride=2), floor mode). */
CF_API int cf_corr_pool(float* pyramid, int B, int H, int W, int L, int from_level, void* stream);

/* ---- a4: CorrBlock.__call__(coords) ------------------------------------------------------------------------
 * replaces `corr_fn(coords1)` (SegFlowGaussian.py:935, 12x per frame pair) = torchvision raft.py:394-422.
 * coords [B,2,H,W] absolute pixel coordinates (ch0 = x, ch1 = y); out [B, L*(2r+1)^2, H, W];
 * channel l*(2r+1)^2 + i*(2r+1) + j samples level l at (x/2^l + i - r, y/2^l + j - r), bilinear,
 * align_corners=True, zero padding.  1 <= r <= CF_MAX_RADIUS. */
CF_API int cf_corr_lookup(const float* pyramid, const float* coords, int B, int H, int W, int L, int r, float* out,
                   void* stream);

/* Parity hook: the fp32 sampling coordinates the lookup kernel derives (after the normalise ->
 * un-normalise round trip) and their floors.  taps float[L][B*H*W][2][2r+1] (axis 2: 0 = x, 1 = y),
 * idx int32 same shape.  Same device code as cf_corr_lookup. */
CF_API int cf_corr_lookup_taps(const float* coords, int B, int H, int W, int L, int r, float* taps, int32_t* idx,
                        void* stream);

/* On-the-fly variant (optional path of north_star): lookup without a materialised volume.  f2_pyramid holds
 * fmap2 average-pooled L-1 times, packed like [B,C,h_l,w_l] per level at cf_fmap_pyramid_level_offset.
 * Uses linearity: avg_pool over q of <f1[p], f2[q]> == <f1[p], avg_pool(f2)[q]>. */
CF_API size_t cf_fmap_pyramid_level_offset(int B, int C, int H, int W, int level);
CF_API size_t cf_fmap_pyramid_bytes(int B, int C, int H, int W, int L);
CF_API int cf_fmap_pyramid_build(const float* f2, int B, int C, int H, int W, int L, float* f2_pyramid, void* stream);
CF_API int cf_corr_lookup_fused(const float* f1, const float* f2_pyramid, const float* coords, int B, int C, int H,
                         int W, int L, int r, float* out, void* stream);

/* The same on the tensor cores (tcgen05): the window correlations of a block of 128 queries (an 8 x 16 patch of the fmap) are
 * one small GEMM per 8 x 16 chunk of the pooled fmap2 covering the block's windows; bf16 operands, fp32 accumulate in TMEM,
 * bilinear taps with the materialised lookup's exact coordinate arithmetic.  `operands` (cf_corr_otf_operand_bytes, 256-byte
 * aligned) receives f1 and the pooled f2 levels as bf16 channels-last; `workspace` = cf_fmap_pyramid_bytes(B, C, H, W, L) bytes
 * of scratch for the fp32 pooled pyramid.  Shapes: H % 8 == 0, W % 16 == 0, C % 8 == 0, C <= 256 (cf_corr_otf_supported);
 * output within 1e-2 of the fp32 volume's lookup (bf16 operands).  Same contract as corr_fn(coords1), SegFlowGaussian.py:935. */
CF_API int cf_corr_otf_supported(int B, int C, int H, int W, int L);
CF_API size_t cf_corr_otf_operand_bytes(int B, int C, int H, int W, int L);
CF_API int cf_corr_otf_prepare(const float* f1, const float* f2, int B, int C, int H, int W, int L, void* operands,
                        void* workspace, size_t workspace_bytes, void* stream);
CF_API int cf_corr_lookup_otf(const void* operands, const float* coords, int B, int C, int H, int W, int L, int r, float* out,
                       void* stream);

/* ---- a5: SpatialTransformer.forward(flow, original) ----------------------------------------------------------
 * replaces nnunet/network_architecture/integration.py:61-79.  flow [B,2,H,W] (ch0 displaces rows, ch1
 * displaces columns), src [B,C,H,W], out [B,C,H,W]; align_corners=True, zero padding. */
CF_API int cf_warp2d(const float* flow, const float* src, int B, int C, int H, int W, int mode, float* out,
              void* stream);

/* Parity hook: sampling coordinates of the warp, taps [B,2,H,W] (ch0 = row coordinate, ch1 = column). */
CF_API int cf_warp2d_taps(const float* flow, int B, int H, int W, float* taps, void* stream);

/* ---- a6: one_hot -> warp -> argmax ------------------------------------------------------------------------------
 * replaces the `warp_linear` pattern, SegFlowGaussian.py:3571-3580 (and trainer :2430-2457).
 * label int64 [B,1,H,W] with values in [0, num_classes); out int64 [B,1,H,W]. */
CF_API int cf_warp2d_onehot_argmax(const float* flow, const int64_t* label, int num_classes, int B, int H, int W,
                            int64_t* out, void* stream);
/* Same, and counts[b][c] (int64 [B, num_classes], overwritten) = number of pixels of image b whose warped label is c:
 * the per-frame class areas the cohort scripts take from the registered masks (`np.count_nonzero(registered == c)`
 * patterns downstream of SegFlowGaussian.py:3571-3580), produced by the same launch instead of num_classes
 * compare-and-reduce passes over the int64 map. */
CF_API int cf_warp2d_onehot_argmax_counts(const float* flow, const int64_t* label, int num_classes, int B, int H, int W,
                                   int64_t* out, int64_t* counts, void* stream);

/* Same with a uint8 result (num_classes <= 256): the registered label maps as the cohort scripts keep them
 * (compute_SSIM.py:26-104 loops per (patient, slice, frame)) and as the multi-GPU driver gathers them -- the kernel writes
 * straight into the rank's slice of the NCCL all-gather buffer, 1 byte per pixel instead of 8.  counts (may be NULL) is
 * int64 [B, num_classes] (counts_total = 0: per frame) or int64 [num_classes] (counts_total = 1: class areas of the whole
 * batch, e.g. of one cine); overwritten. */
CF_API int cf_warp2d_onehot_argmax_u8(const float* flow, const int64_t* label, int num_classes, int B, int H, int W,
                               uint8_t* out, int64_t* counts, int counts_total, void* stream);

/* ---- 8(f) next rows ---------------------------------------------------------------------------------------------
 * Backward passes (autograd): gradients of the warp w.r.t. src and flow, of the lookup w.r.t. the pyramid,
 * of the pooling chain and of the correlation w.r.t. both feature maps.  grad_* outputs that accumulate
 * (atomics) must be zero-initialised by the caller. */
CF_API int cf_warp2d_bwd(const float* flow, const float* src, const float* grad_out, int B, int C, int H, int W,
                  float* grad_src /* zeroed, may be NULL */, float* grad_flow /* may be NULL */, void* stream);
/* Nearest-mode warp (SpatialTransformer(size, mode='nearest')): the output gradient goes to the nearest texel (grad_src
 * zeroed by the caller, accumulated with atomics); the gradient w.r.t. the flow is identically zero, as in ATen. */
CF_API int cf_warp2d_nearest_bwd(const float* flow, const float* grad_out, int B, int C, int H, int W, float* grad_src,
                          void* stream);
CF_API int cf_corr_lookup_bwd(const float* coords, const float* grad_out, int B, int H, int W, int L, int r,
                       float* grad_pyramid /* packed like the pyramid; ACCUMULATED into (+=): zero it once, then every
                                              lookup of the same CorrBlock may add its gradient to the same buffer */,
                       void* stream);
CF_API int cf_corr_pool_bwd(float* grad_pyramid, int B, int H, int W, int L, void* stream);
CF_API int cf_corr_volume_bwd(const float* f1, const float* f2, const float* grad_corr /* level 0 */, int B, int C,
                       int H, int W, float* grad_f1, float* grad_f2, void* stream);

/* The same two products on the tensor cores (tcgen05, fp32-grade through the bf16x3 split: |error| <= ~1e-5 of the result's scale;
 * corr_bwd_tc.cu).  precision: CF_PREC_AUTO (tensor cores when H*W % 8 == 0, else the fp32 CUDA-core GEMMs), CF_PREC_FP32,
 * CF_PREC_BF16X3 (error when unsupported).  workspace: cf_corr_volume_bwd_workspace_bytes (0 = the tensor-core path does not
 * apply), 256-byte aligned device memory.  *used_precision reports the path taken.  Reference: autograd through
 * SegFlowGaussian.py:929-935 (training loop SegFlowGaussian.py:2225 of network_training). */
CF_API size_t cf_corr_volume_bwd_workspace_bytes(int B, int C, int H, int W);
CF_API int cf_corr_volume_bwd_ex(const float* f1, const float* f2, const float* grad_corr /* level 0 */, int B, int C, int H, int W,
                          int precision, float* grad_f1 /* may be NULL */, float* grad_f2 /* may be NULL */, void* workspace,
                          size_t workspace_bytes, int* used_precision /* may be NULL */, void* stream);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate device memory
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  cf_corr_volume_bwd_ex's tensor-core
 * path with its scratch (the four bf16 hi / lo planes of the feature maps) in a static device array of CF_NOALLOC_SCRATCH_BYTES
 * that every GPU loading the library holds until exit: the call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine).  Same results as cf_corr_volume_bwd_ex(..., CF_PREC_BF16X3, ...).  Errors (CF_ERR_UNSUPPORTED, nothing is
 * launched): H*W % 8 != 0, or 4 * round_up_256(B*C*H*W*2) > CF_NOALLOC_SCRATCH_BYTES (e.g. up to B = 4 at C = 256, 32 x 32).
 * All calls on a device share the array, which keeps the previous call's contents (nothing in it needs clearing: each call
 * overwrites what it reads); calls that overlap on the GPU -- different streams, concurrently replayed graphs -- RACE on it. */
#define CF_NOALLOC_SCRATCH_BYTES (8 * 1024 * 1024)
CF_API int cf_corr_volume_bwd_noalloc(const float* f1, const float* f2, const float* grad_corr /* level 0 */, int B, int C, int H,
                               int W, float* grad_f1 /* may be NULL */, float* grad_f2 /* may be NULL */, void* stream);

/* VecInt (integration.py:82-99): vec/2^n then n x (vec += warp(vec, vec)); work = scratch [B,2,H,W]. */
CF_API int cf_vecint(const float* vec, int B, int H, int W, int nsteps, float* out, float* work, void* stream);

/* SpatialTransformerContour (integration.py:5-34): sample src [B,C,H,W] at P points; locs [B,2,P] with
 * ch0 = x, ch1 = y (the reference's un-swapped order), in pixels of a (size_h, size_w) transformer: the reference
 * normalises channel i by `self.shape[~i] - 1` (x by size_w - 1, y by size_h - 1) and grid_sample un-normalises
 * with the image's own (H, W); out [B,C,P]. */
CF_API int cf_sample_points(const float* locs, const float* src, int B, int C, int H, int W, int P, int size_h,
                     int size_w, int mode, float* out, void* stream);

/* Backward of cf_sample_points: the reference back-propagates its MSE point loss through SpatialTransformerContour into
 * `original` (= the predicted flow; training/network_training/SegFlowGaussian.py:1497-1503).  grad_out [B,C,P];
 * grad_src [B,C,H,W] accumulates with atomics (zero-initialised by the caller; may be NULL); grad_locs [B,2,P] (gradient
 * w.r.t. the un-normalised point coordinates; written; may be NULL).  Nearest mode: grad_locs = 0 (as ATen). */
CF_API int cf_sample_points_bwd(const float* locs, const float* src, const float* grad_out, int B, int C, int H, int W, int P,
                         int size_h, int size_w, int mode, float* grad_src, float* grad_locs, void* stream);

/* lib/utils.warp (8(f) row 4): replaces `warp(x, flo)` of nnunet/lib/utils.py:29-59 (call site SegFlowGaussian.py:1044).
 * flow [B,2,H,W] with ch0 displacing COLUMNS and ch1 rows; n = 2*(p+f)/max(S-1,1) - 1; grid_sample defaults:
 * bilinear, zero padding, align_corners=False.  src / out [B,C,H,W]. */
CF_API int cf_warp2d_xy(const float* flow, const float* src, int B, int C, int H, int W, float* out, void* stream);

/* Backward of cf_warp2d_xy (the reference warps the looked-up correlation features inside the training graph,
 * SegFlowGaussian.py:1044): grad_src accumulates with atomics (zero-initialised by the caller; may be NULL), grad_flow is
 * written (may be NULL). */
CF_API int cf_warp2d_xy_bwd(const float* flow, const float* src, const float* grad_out, int B, int C, int H, int W,
                     float* grad_src, float* grad_flow, void* stream);

/* Local correlation layer (8(f) row 4, first half): `CorrVolume(radius, stride)(cur, prev)` of the reference's skip connections
 * (SegFlowGaussian.py:41 import, :259 construction, :1376 call).  PARITY UNPINNED: its source nnunet/lib/raft.py was never
 * committed; only the call-site contract [B,C,H,W] x [B,C,H,W] -> [B,(2r+1)^2,H,W] is known.  Implements the standard local cost
 * volume out[b, (dy+r)*(2r+1) + (dx+r), y, x] = scale * sum_c cur[b,c,y,x] * prev[b,c,y+dy*stride,x+dx*stride], zero padding
 * (x_major != 0: channel (dx+r)*(2r+1) + (dy+r), RAFT's lookup order).  1 <= radius <= CF_MAX_RADIUS.
 * Backward: grad_cur / grad_prev [B,C,H,W] are written (either may be NULL). */
CF_API int cf_local_corr(const float* cur, const float* prev, int B, int C, int H, int W, int radius, int stride, float scale,
                  int x_major, float* out, void* stream);
CF_API int cf_local_corr_bwd(const float* cur, const float* prev, const float* grad_out, int B, int C, int H, int W, int radius,
                      int stride, float scale, int x_major, float* grad_cur, float* grad_prev, void* stream);

/* Convex 8x upsampling (8(f) row 3): replaces `upsample_flow` (SegFlowGaussian.py:846-857, C = 2) and `upsample_seg`
 * (:860-871, C = 4): out[n,c,8y+i,8x+j] = sum_k softmax_k(mask[n,k*64+i*8+j,y,x]) * scale * field[n,c,y+ky-1,x+kx-1],
 * k = ky*3+kx, zeros outside.  field [B,C,H,W], mask [B,576,H,W], out [B,C,8H,8W]; the reference passes scale = 8
 * for both.  Only factor 8 exists in the reference (CF_ERR_UNSUPPORTED otherwise). */
CF_API int cf_upsample_convex(const float* field, const float* mask, int B, int C, int H, int W, int factor, float scale,
                       float* out, void* stream);

/* Backward of cf_upsample_convex (autograd of the same two reference methods): grad_out [B,C,8H,8W] ->
 * grad_field [B,C,H,W] (accumulated with atomics: zero-initialised by the caller; may be NULL) and grad_mask [B,576,H,W]
 * (written; may be NULL). */
CF_API int cf_upsample_convex_bwd(const float* field, const float* mask, const float* grad_out, int B, int C, int H, int W,
                           int factor, float scale, float* grad_field, float* grad_mask, void* stream);

/* Local normalised cross-correlation, the registration loss taken on the warped frame (adjacent row of SURVEY 8(f)):
 * replaces `NCC.forward` of nnunet/lib/loss.py:15-76 (constructed at training/network_training/SegFlowGaussian.py:337,
 * FinalFlowRaft.py:259, nnMTLTrainerV2Raft.py:275, Optical_flow_model_successive.py:71).  I, J [N,1,H,W] fp32 (the ones
 * kernel of the reference has one input channel); five zero-padded win x win box sums, then
 * cc = clamp(cross^2 / (I_var * J_var + eps), 0.001, 0.999).  loss_map [N,H,W] (may be NULL) receives 1 - cc (what reduction=None
 * returns); cc_sum [1] (may be NULL) receives sum(cc) in a fixed summation order (the host forms 1 - cc_sum / (N*H*W)); workspace of
 * cf_ncc_workspace_bytes(N, H, W, 0) bytes is needed when cc_sum is given.  Only win = 9 exists in the reference
 * (CF_ERR_UNSUPPORTED otherwise). */
CF_API size_t cf_ncc_workspace_bytes(int N, int H, int W, int backward);
CF_API int cf_ncc(const float* I, const float* J, int N, int H, int W, int win, float eps, float* loss_map, float* cc_sum,
           void* workspace, size_t ws_bytes, void* stream);

/* Backward of cf_ncc (the loss is back-propagated through the warp into the flow, trainer SegFlowGaussian.py:2225).
 * d(loss)/d(cc[p]) = mult * grad[p] when per_pixel (grad [N,H,W]: reduction=None, mult = -1 for `1 - cc`), else
 * mult * grad[0] (grad = the scalar upstream gradient on the device: reduction='mean', mult = -1/(N*H*W)); clamped pixels
 * pass no gradient.  grad_I / grad_J [N,1,H,W] are written (either may be NULL); workspace of
 * cf_ncc_workspace_bytes(N, H, W, 1) bytes holds the five coefficient maps between the two launches. */
CF_API int cf_ncc_bwd(const float* I, const float* J, int N, int H, int W, int win, float eps, const float* grad, int per_pixel,
               float mult, float* grad_I, float* grad_J, void* workspace, size_t ws_bytes, void* stream);

/* Deterministic float64 sum of an fp32 device array (summation order depends on n alone): the compact per-unit results the
 * cohort driver gathers over NCCL (shard.py; the reference's analysis scripts reduce warped frames / flows to per-frame numbers
 * the same way, e.g. compute_SSIM.py:89-104) are checked bit for bit against a recomputation.  x must be 16-byte aligned;
 * workspace of cf_sum_f64_workspace_bytes() bytes; out [1] double. */
CF_API size_t cf_sum_f64_workspace_bytes(void);
CF_API int cf_sum_f64(const float* x, size_t n, double* out, void* workspace, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CARDIOFLOW_H_ */
