/*
 * raft_corr.h -- C ABI of libraftcorr: RAFT's correlation hot path on B200 (sm_100a).
 *
 * Drop-in boundary for athaddius/RAFT_STIR (paths below are relative to that
 * repository).  Plain pointers and sizes only; every pointer is a DEVICE
 * pointer on the current CUDA device unless said otherwise.  The library
 * allocates nothing in these calls: outputs and workspaces are the caller's.
 * All launches are asynchronous on `stream`.
 *
 * Return value: 0 on success, a negative RC_E* code on failure;
 * rc_last_error() then returns a thread-local message.  No exceptions cross
 * this boundary.  There is no CPU path: without a CUDA device every compute
 * entry point fails with RC_ECUDA.
 */
#ifndef RAFT_CORR_H_
#define RAFT_CORR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define RC_API __attribute__((visibility("default")))
#else
#define RC_API
#endif

#define RC_ABI_VERSION 6
#define RC_MAX_LEVELS 6

/* error codes */
#define RC_OK 0
#define RC_EINVAL (-1)   /* bad argument (null pointer, size, dtype, radius ...) */
#define RC_ECUDA (-2)    /* a CUDA runtime / driver call failed */
#define RC_ESIZE (-3)    /* workspace or static scratch too small for the request */

/* storage type of the correlation volume / pyramid */
#define RC_VOL_F32 0
#define RC_VOL_BF16 1
#define RC_VOL_F16 2     /* IEEE half: the bytes of bf16 with three more mantissa bits; saturates at +-65 504 */

/* arithmetic used to form the feature dot products of the volume build */
#define RC_MATH_F32 0    /* CUDA-core fp32 FMA: what the reference's SGEMM does */
#define RC_MATH_BF16 1   /* tcgen05 kind::f16, bf16 operands, fp32 accumulate in TMEM */
#define RC_MATH_TF32 2   /* tcgen05 kind::tf32: the fp32 features rounded to tf32 (10 mantissa bits), fp32 accumulate */
#define RC_MATH_F16 3    /* tcgen05 kind::f16 on IEEE-half operands (10 mantissa bits, range +-65 504), fp32 accumulate */

typedef struct CUstream_st *rc_stream_t;   /* == cudaStream_t */

/*
 * Memory layout of the pooled correlation pyramid (all levels in ONE buffer).
 *
 * Level l holds, for every query (b, q), a slab: the H[l] x W[l] map of that
 * query's correlations (H[l] = H >> l, W[l] = W >> l: what three
 * floor-truncating avg_pool2d(2, 2) calls give, core/corr.py:24-27), cut into
 * TILES of tile_h = 4 rows x tile_w columns, tile_w = 16 bytes of elements
 * (8 bf16 / 4 fp32), so that one tile is 64 contiguous, 64-byte-aligned bytes:
 * the unit in which B200's L2 fetches from HBM.  A (2r+2)^2 lookup window then
 * touches ~7 tiles instead of 10 half-used row pieces.  Tiles are stored row
 * by row; element (b, q, y, x) of level l lives at (units: elements)
 *     offset[l] + ((size_t)b * N + q) * slab[l]
 *               + ((y / 4) * tiles_x[l] + x / tile_w) * (4 * tile_w)
 *               + (y % 4) * tile_w + x % tile_w.
 * Elements of edge tiles beyond H[l] / W[l] are PADDING; every producer in this
 * library writes zeros there and the lookup relies on it.
 */
typedef struct rc_layout {
  int32_t num_levels;
  int32_t B, N;                    /* N = H * W queries per pair */
  int32_t tile_h, tile_w;          /* 4, and 8 (bf16) or 4 (fp32) */
  int32_t H[RC_MAX_LEVELS];
  int32_t W[RC_MAX_LEVELS];
  int32_t tiles_y[RC_MAX_LEVELS];  /* ceil(H[l] / tile_h) */
  int32_t tiles_x[RC_MAX_LEVELS];  /* ceil(W[l] / tile_w) */
  int64_t slab[RC_MAX_LEVELS];     /* tiles_y * tiles_x * tile_h * tile_w */
  int64_t offset[RC_MAX_LEVELS];
  int64_t total_elems;             /* size of the whole buffer, elements */
  int32_t elem_bytes;              /* 4 (RC_VOL_F32) or 2 (RC_VOL_BF16) */
} rc_layout;

RC_API int rc_abi_version(void);
RC_API const char *rc_last_error(void);

/* Fills *out for a pyramid of `num_levels` levels over B pairs of H x W
 * feature maps.  Host-only arithmetic; needs no GPU. */
RC_API int rc_pyramid_layout(int B, int H, int W, int num_levels, int vol_dtype, rc_layout *out);

/* Bytes of scratch rc_build_pyramid needs for its repacked feature maps
 * (0 for RC_MATH_F32).  Host-only arithmetic. */
RC_API int64_t rc_build_workspace_bytes(int B, int D, int H, int W, int math);

/*
 * Volume build + pooled pyramid.  Replaces CorrBlock.corr (core/corr.py:52-60:
 * matmul, then a separate / sqrt(D) pass) and the three avg_pool2d calls of
 * CorrBlock.__init__ (core/corr.py:21-27).
 *   fmap1, fmap2 : fp32 [B, D, H, W] contiguous (NCHW, as RAFT.forward passes
 *                  them, core/raft.py:102-107)
 *   pyramid      : out, rc_layout of (B, H, W, num_levels, vol_dtype)
 *   workspace    : >= rc_build_workspace_bytes(...) bytes, 256-byte aligned
 * The tensor-core modes (RC_MATH_BF16 / _F16 / _TF32) need D % 8 == 0 and num_levels <= 4; other sizes
 * return RC_EINVAL before anything is launched (RC_MATH_F32 takes any D and up to RC_MAX_LEVELS levels).
 * The tensor-core kernel runs as two-CTA clusters (tcgen05 cta_group::2); with RC_BUILD_PAIR=0 in the
 * environment (read once, at the first build of the process) it runs its single-CTA form, which gives the same bits.
 * If the device cannot co-schedule one such cluster the call returns RC_ECUDA and the message names that switch.
 */
RC_API int rc_build_pyramid(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                     int num_levels, int vol_dtype, int math, void *pyramid,
                     void *workspace, size_t workspace_bytes, rc_stream_t stream);

/*
 * rc_build_pyramid for callers that cannot allocate device memory.
 * (No counterpart in the reference: its torch ops allocate freely.)
 *
 * Some hosts of this library may not call cudaMalloc at the point where they
 * need a volume: a real-time tracker that fixed its memory plan at start-up, a
 * plugin running inside another framework's arena, code between allocator
 * checkpoints.  They still own the inputs and the output pyramid, but have
 * nowhere to put the repacked bf16 features the tensor-core build needs.  This
 * entry point keeps that scratch in a fixed array inside the library
 * (RC_STATIC_SCRATCH_BYTES, resident on every GPU that loads the library until
 * the process exits) and allocates nothing: no cudaMalloc / cudaFree, no
 * allocating library routine.
 *
 * Same arguments and result as rc_build_pyramid(..., RC_MATH_BF16, ...) minus the
 * workspace.  Limits and caveats:
 *   - rc_build_workspace_bytes(B, D, H, W, RC_MATH_BF16) must not exceed
 *     RC_STATIC_SCRATCH_BYTES (one or two Sintel-size pairs); larger requests
 *     fail with RC_ESIZE and launch nothing.
 *   - The first call on each device looks the array's device address up
 *     (cudaGetSymbolAddress); make that first call outside any stream capture.
 *   - All calls on a device share the one array.  Nothing in it has to start
 *     at zero (every call rewrites the part it uses), but two calls whose
 *     kernels overlap on the GPU -- different streams, no ordering between them
 *     -- race on it.  Keep them on one stream or order them with events.
 */
#define RC_STATIC_SCRATCH_BYTES (16u << 20)
RC_API int rc_build_pyramid_static(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                                   int num_levels, int vol_dtype, void *pyramid, rc_stream_t stream);

/*
 * The same build for a caller that already holds 16-bit channels-last features (SURVEY.md section 8(f) row 4:
 * the reference's encoder ends in a 1x1 convolution, core/extractor.py:144,184; run in channels_last bf16 it
 * emits exactly this layout, and the fp32 round trip of core/raft.py:99-103 plus the repack disappear).
 *   fmap1_nhwc16, fmap2_nhwc16 : bf16 or fp16 [B, H, W, D] contiguous, UNSCALED, 128-byte aligned, D % 64 == 0
 *   operand_dtype              : RC_VOL_BF16 or RC_VOL_F16 (the element type of the two buffers = the MMA operands)
 * No workspace: fmap1 is read as [B*N, D] rows; the 8 x 32 spatial blocks of fmap2 that form one GEMM tile are
 * cut out of the NHWC tensor by a 4-D TMA box (rows outside the map arrive as zeros), and 1/sqrt(D) is applied
 * to the fp32 accumulators.  Same result as rc_build_pyramid on the same 16-bit values.
 * rc_pack_features_nhwc16 produces such buffers from the fp32 NCHW maps (one transposing cast per map), e.g. to
 * share one packed copy between this build and rc_alt_lookup_fwd_tc (whose fmap1_16 is this fmap1_nhwc16).
 */
RC_API int rc_pack_features_nhwc16(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                            int operand_dtype, void *fmap1_nhwc16, void *fmap2_nhwc16, rc_stream_t stream);
RC_API int rc_build_pyramid_nhwc16(const void *fmap1_nhwc16, const void *fmap2_nhwc16, int operand_dtype, int B,
                            int D, int H, int W, int num_levels, int vol_dtype, void *pyramid,
                            rc_stream_t stream);

/*
 * Windowed bilinear lookup over all levels in one launch.  Replaces
 * CorrBlock.__call__ (core/corr.py:29-50) together with bilinear_sampler /
 * F.grid_sample(align_corners=True, zeros padding) (core/utils/utils.py:57-65).
 *   coords : fp32 [B, 2, H, W], channel 0 = x, 1 = y (core/utils/utils.py:74-77)
 *   out    : fp32 [B, num_levels*(2r+1)^2, H, W]; channel l*(2r+1)^2 + i*(2r+1) + j
 *            samples level l at (x/2^l + i - r, y/2^l + j - r)
 *   radius : 1..4
 * Finite coordinates far outside the map give zeros (zeros padding); NaN or infinite coordinates give NaN in
 * every sample of that query, as the reference's grid_sample does (the on-the-fly entry points do the same).
 */
RC_API int rc_lookup_fwd(const void *pyramid, int vol_dtype, const float *coords, int B, int H, int W,
                  int num_levels, int radius, float *out, rc_stream_t stream);
/*
 * The same lookup with the result rounded once to IEEE half precision: out_f16 is fp16 [B, num_levels*(2r+1)^2, H, W].
 * For callers whose next op runs under fp16 autocast (the update block, core/raft.py:127-128, with
 * --mixed_precision): the reference hands over fp32 and autocast's first convolution casts it; here the cast pass
 * and half of the lookup's output bytes disappear.  No counterpart in the reference; inference only.
 */
RC_API int rc_lookup_fwd_f16(const void *pyramid, int vol_dtype, const float *coords, int B, int H, int W,
                      int num_levels, int radius, void *out_f16, rc_stream_t stream);

/*
 * The lookup fused with the layer that consumes it: the motion encoder's
 *     cor = F.relu(self.convc1(corr))          convc1 = Conv2d(num_levels*(2r+1)^2, O, kernel 1)
 * of core/update.py:70-71 (SmallMotionEncoder) and :88-90 (BasicMotionEncoder).  The reference writes the lookup's
 * samples as fp32 [B, C, H, W], reads them back and convolves in fp32; here the lookup leaves them once, in half
 * precision, in `workspace`, and a tcgen05 GEMM (fp16 operands, fp32 accumulate, weights resident in shared memory)
 * applies weights, bias and ReLU: half the intermediate bytes, no fp32 round trip, the contraction on the tensor cores.
 *   weight_f16 : the layer's weight [O, C] converted once by rc_convc1_pack_weights (rc_convc1_weight_bytes bytes,
 *                128-byte aligned); bias: fp32 [O]; O a multiple of 16, at most 256
 *   out        : [B, O, H, W], fp32 (RC_VOL_F32) or fp16 (RC_VOL_F16, for an update block under fp16 autocast)
 *   workspace  : >= rc_lookup_convc1_workspace_bytes(...) bytes, 128-byte aligned
 * Inference only (no backward).  Same coordinate conventions as rc_lookup_fwd.
 */
RC_API int64_t rc_convc1_weight_bytes(int O, int C);
RC_API int rc_convc1_pack_weights(const float *weight, int O, int C, void *weight_f16, rc_stream_t stream);
RC_API int64_t rc_lookup_convc1_workspace_bytes(int B, int H, int W, int num_levels, int radius);
RC_API int rc_lookup_convc1_fwd(const void *pyramid, int vol_dtype, const float *coords, int B, int H, int W,
                         int num_levels, int radius, const void *weight_f16, const float *bias, int O, void *out,
                         int out_dtype, void *workspace, size_t workspace_bytes, rc_stream_t stream);

/*
 * Backward of rc_lookup_fwd: what autograd of F.grid_sample yields at
 * core/utils/utils.py:65 for every level, in one launch.
 *   grad_out     : fp32 [B, num_levels*(2r+1)^2, H, W]
 *   grad_pyramid : fp32 pyramid, rc_layout(..., RC_VOL_F32); ACCUMULATED into (+=),
 *                  so the 12 GRU iterations can share one buffer; may be NULL
 *   grad_coords  : fp32 [B, 2, H, W], overwritten; may be NULL
 */
RC_API int rc_lookup_bwd(const void *pyramid, int vol_dtype, const float *coords, const float *grad_out,
                  int B, int H, int W, int num_levels, int radius, float *grad_pyramid,
                  float *grad_coords, rc_stream_t stream);

/* Bytes of scratch rc_build_bwd needs. Host-only arithmetic. */
RC_API int64_t rc_build_bwd_workspace_bytes(int B, int D, int H, int W, int num_levels);

/*
 * Backward of rc_build_pyramid: transposes of the three avg_pool2d calls and of
 * the matmul of core/corr.py:58-60 (what autograd runs for train.py:175).  Both
 * contractions run on the tensor cores (tcgen05 kind::tf32) over the gradient
 * pyramid in place; no library call, nothing allocated, capturable.
 *   grad_pyramid   : fp32 pyramid (RC_VOL_F32 layout), read only, 16-byte aligned
 *   dfmap1, dfmap2 : fp32 [B, D, H, W], overwritten
 *   math           : RC_MATH_TF32  one tf32 pass (features rounded to nearest, gradients truncated
 *                                  to tf32): ~5e-4 relative, the setting for bf16 / tf32 volumes
 *                    RC_MATH_F32   3xTF32 (every operand split into hi + lo in shared memory, three
 *                                  MMA passes): fp32-grade, what the reference's fp32 SGEMMs give
 *   workspace      : >= rc_build_bwd_workspace_bytes(...) bytes, 256-byte aligned
 */
RC_API int rc_build_bwd(const float *grad_pyramid, const float *fmap1, const float *fmap2, int B, int D,
                 int H, int W, int num_levels, int math, float *dfmap1, float *dfmap2, void *workspace,
                 size_t workspace_bytes, rc_stream_t stream);

/*
 * On-the-fly correlation, one level: same contract as alt_cuda_corr.forward
 * (alt_cuda_corr/correlation.cpp:23-33 -> correlation_kernel.cu:260-286,
 * kernel :18-119).  Never materialises the volume.
 *   fmap1  : fp32 [B, H1, W1, C]      fmap2 : fp32 [B, H2, W2, C]
 *   coords : fp32 [B, Nc, H1, W1, 2]  (x, y) in fmap2 pixels
 *   corr   : out fp32 [B, Nc, (2r+1)^2, H1, W1], overwritten (the reference
 *            zero-fills then accumulates, :273); NOT scaled by 1/sqrt(C)
 */
RC_API int rc_alt_fwd(const float *fmap1, const float *fmap2, const float *coords, int B, int H1, int W1,
               int H2, int W2, int C, int Nc, int radius, float *corr, rc_stream_t stream);

/*
 * Same contract as alt_cuda_corr.backward (correlation.cpp:36-48 ->
 * correlation_kernel.cu:288-324, kernel :122-256).
 *   fmap1_grad [B,H1,W1,C], fmap2_grad [B,H2,W2,C], coords_grad [B,Nc,H1,W1,2]:
 *   all overwritten; coords_grad is zero-filled, as in the reference, whose
 *   kernel never writes it (:307).
 */
RC_API int rc_alt_bwd(const float *fmap1, const float *fmap2, const float *coords, const float *corr_grad,
               int B, int H1, int W1, int H2, int W2, int C, int Nc, int radius, float *fmap1_grad,
               float *fmap2_grad, float *coords_grad, rc_stream_t stream);

/*
 * AlternateCorrBlock.__call__ in one launch (core/corr.py:74-91): all levels,
 * NCHW features read in place, 1/sqrt(D) applied.
 *   fmap1      : fp32 [B, D, H, W]
 *   fmap2_pyr  : fp32, level l is [B, H>>l, W>>l, D] (NHWC) at element offset
 *                f2_offset[l]; produce it with rc_alt_pool_features
 *   coords     : fp32 [B, 2, H, W];  out: fp32 [B, num_levels*(2r+1)^2, H, W]
 */
RC_API int64_t rc_alt_feature_pyramid_elems(int B, int D, int H, int W, int num_levels, int64_t *f2_offset);
RC_API int rc_alt_pool_features(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                         int num_levels, float *fmap1_nhwc, float *fmap2_pyr, rc_stream_t stream);
RC_API int rc_alt_lookup_fwd(const float *fmap1_nhwc, const float *fmap2_pyr, const float *coords, int B,
                      int D, int H, int W, int num_levels, int radius, float *out,
                      rc_stream_t stream);

/*
 * The same AlternateCorrBlock.__call__ (core/corr.py:74-91) on the tensor cores: per block of 8 x 16
 * queries and level, the block's local volume against the bounding box of its windows is formed with
 * tcgen05 (bf16 operands, fp32 accumulate) and the windows are read out of it.  Same output as
 * rc_alt_lookup_fwd to 16-bit-operand precision; valid for any coordinates (a scattered block only costs more
 * chunks).  Needs D % 64 == 0.  operand_dtype: RC_VOL_BF16 (8-bit exponent: any feature range) or RC_VOL_F16
 * (three more mantissa bits; features beyond +-65 504 saturate).
 *   rc_alt_features_16 : converts what rc_alt_pool_features produced (fmap1 NHWC + fmap2 feature
 *                        pyramid, same element offsets) to that type
 */
RC_API int rc_alt_features_16(const float *fmap1_nhwc, const float *fmap2_pyr, int B, int D, int H, int W,
                       int num_levels, int operand_dtype, void *fmap1_16, void *fmap2_pyr_16, rc_stream_t stream);
RC_API int rc_alt_lookup_fwd_tc(const void *fmap1_16, const void *fmap2_pyr_16, int operand_dtype, const float *coords,
                         int B, int D, int H, int W, int num_levels, int radius, float *out, rc_stream_t stream);

/*
 * Flow upsampling to full resolution, the step after the lookup + update block in every refinement
 * iteration.
 *   rc_upsample_flow      replaces RAFT.upsample_flow (core/raft.py:72-83): softmax over the 9 neighbours of the
 *                         576-channel mask (channel = k*64 + sy*8 + sx), convex combination of the 3x3
 *                         neighbourhood of 8*flow (zeros outside the map), output pixel (8y+sy, 8x+sx); one
 *                         kernel instead of softmax + unfold + multiply + sum + permute
 *   rc_upsample_flow_bwd  what autograd derives for it: d/dflow [B,2,H,W] (overwritten) and d/dmask [B,576,H,W]
 *   rc_upflow8 / _bwd     replace upflow8 (core/utils/utils.py:80-82): 8 * bilinear x8, align_corners=True
 *                         (the path RAFT-small takes, core/raft.py:133)
 *   flow : fp32 [B, 2, H, W] (rc_upflow8: [B, C, H, W]);  mask: fp32 [B, 576, H, W]
 *   out / gout : fp32 [B, 2, 8H, 8W].  All contiguous.  No workspace, nothing allocated.
 */
RC_API int rc_upsample_flow(const float *flow, const float *mask, int B, int H, int W, float *out,
                     rc_stream_t stream);
/* the same reading the mask as IEEE half [B, 576, H, W], as fp16 autocast leaves it (softmax still in fp32) */
RC_API int rc_upsample_flow_f16mask(const float *flow, const void *mask_f16, int B, int H, int W, float *out,
                             rc_stream_t stream);
RC_API int rc_upsample_flow_bwd(const float *flow, const float *mask, const float *gout, int B, int H, int W,
                         float *dflow, float *dmask, rc_stream_t stream);
RC_API int rc_upflow8(const float *flow, int B, int C, int H, int W, float *out, rc_stream_t stream);
RC_API int rc_upflow8_bwd(const float *gout, int B, int C, int H, int W, float *dflow, rc_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RAFT_CORR_H_ */
