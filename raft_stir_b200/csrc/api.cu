// extern "C" entry points of libraftcorr (declared in include/raft_corr.h).
// Argument validation, layout arithmetic and dispatch only; kernels live in the
// other translation units.
#include <math.h>
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace rc {

static thread_local char g_error[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

// alt.cu
int launch_alt_fwd(const float *fmap1, const float *fmap2, const float *coords, int B, int H1, int W1,
                   int H2, int W2, int C, int Nc, int radius, float *corr, cudaStream_t s);
int launch_alt_lookup_fwd(const float *fmap1_nhwc, const float *fmap2_pyr, const float *coords, int B,
                          int D, int H, int W, int num_levels, int radius, const int64_t *off,
                          float *out, cudaStream_t s);
int launch_alt_bwd(const float *fmap1, const float *fmap2, const float *coords, const float *corr_grad,
                   int B, int H1, int W1, int H2, int W2, int C, int Nc, int radius, float *g1,
                   float *g2, cudaStream_t s);
int launch_alt_pool(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int num_levels,
                    const int64_t *off, float *f1_nhwc, float *f2_pyr, cudaStream_t s);
// alt_tc.cu
int launch_alt_features_16(const float *src, void *dst, long long elems, int half_operands, cudaStream_t s);
int launch_alt_lookup_tc(const void *f1_bf16, const void *f2_pyr_bf16, const float *coords, int B, int D, int H,
                         int W, int num_levels, int radius, const int64_t *off, int half_operands, float *out,
                         cudaStream_t s);
// upsample.cu
int launch_upsample_flow(const float *flow, const void *mask, int mask_is_half, int B, int H, int W, float *out,
                         cudaStream_t s);
int launch_upsample_flow_bwd(const float *flow, const float *mask, const float *gout, int B, int H, int W,
                             float *dflow, float *dmask, cudaStream_t s);
int launch_upflow8(const float *flow, int B, int C, int H, int W, float *out, cudaStream_t s);
int launch_upflow8_bwd(const float *gout, int B, int C, int H, int W, float *dflow, cudaStream_t s);
// build_bwd.cu
int64_t build_bwd_workspace_bytes(int B, int D, int H, int W, int num_levels);
int launch_build_bwd(const float *grad_pyramid, const float *fmap1, const float *fmap2, int B, int D,
                     int H, int W, const rc_layout &glay, int precise, float *dfmap1, float *dfmap2,
                     void *workspace, cudaStream_t s);

// Scratch of rc_build_pyramid_static: the packed bf16 operands of the tensor-core
// build.  In host code this name is only a handle for cudaGetSymbolAddress; the
// kernels get the looked-up device pointer.
static __device__ __align__(256) unsigned char g_static_scratch[RC_STATIC_SCRATCH_BYTES];

// Device address of g_static_scratch on the current device, looked up once per device.
static int static_scratch_ptr(void **out) {
  static void *cache[64] = {};
  int dev = 0;
  RC_CUDA(cudaGetDevice(&dev));
  RC_REQUIRE(dev >= 0 && dev < 64, "device index %d out of range", dev);
  void *p = __atomic_load_n(&cache[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    RC_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    RC_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&cache[dev], p, __ATOMIC_RELEASE);   // same value from any thread
  }
  *out = p;
  return RC_OK;
}

static int make_layout(int B, int H, int W, int num_levels, int vol_dtype, rc_layout *out) {
  RC_REQUIRE(out != nullptr, "rc_pyramid_layout: out is NULL");
  RC_REQUIRE(B >= 1 && H >= 1 && W >= 1, "rc_pyramid_layout: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(num_levels >= 1 && num_levels <= RC_MAX_LEVELS, "num_levels %d outside 1..%d", num_levels,
             RC_MAX_LEVELS);
  RC_REQUIRE(vol_dtype == RC_VOL_F32 || vol_dtype == RC_VOL_BF16 || vol_dtype == RC_VOL_F16, "unknown vol_dtype %d", vol_dtype);
  RC_REQUIRE((H >> (num_levels - 1)) >= 1 && (W >> (num_levels - 1)) >= 1,
             "a %dx%d map has no level %d (avg_pool2d would produce an empty tensor)", H, W,
             num_levels - 1);
  RC_REQUIRE((long long)H * W < (1ll << 31) / 64, "feature map too large");
  memset(out, 0, sizeof(*out));
  out->num_levels = num_levels;
  out->B = B;
  out->N = H * W;
  out->elem_bytes = vol_dtype == RC_VOL_F32 ? 4 : 2;
  // 64-byte tiles: 4 rows of 16 bytes
  out->tile_h = TILE_H;
  out->tile_w = 16 / out->elem_bytes;
  int64_t off = 0;
  for (int l = 0; l < num_levels; ++l) {
    out->H[l] = H >> l;
    out->W[l] = W >> l;
    out->tiles_y[l] = ceil_div(out->H[l], out->tile_h);
    out->tiles_x[l] = ceil_div(out->W[l], out->tile_w);
    out->slab[l] = (int64_t)out->tiles_y[l] * out->tiles_x[l] * out->tile_h * out->tile_w;
    out->offset[l] = off;
    off += round_up64((int64_t)B * out->N * out->slab[l], 256);
  }
  out->total_elems = off;
  return RC_OK;
}

// Limits of the tensor-core build (build_tc.cu), checked BEFORE anything is launched: the pack kernel writes
// channel pairs and the fused pyramid epilogue holds four levels.
static int check_tc_limits(const char *who, int D, int num_levels) {
  RC_REQUIRE(D % 8 == 0, "%s: the tensor-core build needs D %% 8 == 0 (got %d); use RC_MATH_F32", who, D);
  RC_REQUIRE(num_levels <= 4, "%s: the tensor-core build supports at most 4 pyramid levels (got %d); use RC_MATH_F32",
             who, num_levels);
  return RC_OK;
}

static int64_t packed_rows2(int H, int W) {
  return (int64_t)ceil_div(W, TB_W) * ceil_div(H, TB_H) * TB_N;
}

}  // namespace rc

using namespace rc;

extern "C" {

int rc_abi_version(void) { return RC_ABI_VERSION; }

const char *rc_last_error(void) { return g_error; }

int rc_pyramid_layout(int B, int H, int W, int num_levels, int vol_dtype, rc_layout *out) {
  return make_layout(B, H, W, num_levels, vol_dtype, out);
}

int64_t rc_build_workspace_bytes(int B, int D, int H, int W, int math) {
  if (math == RC_MATH_F32) return 0;
  if (B < 1 || D < 1 || H < 1 || W < 1) return RC_EINVAL;
  const int64_t es = (math == RC_MATH_BF16 || math == RC_MATH_F16) ? 2 : 4;
  const int64_t a = round_up64((int64_t)B * H * W * D * es, 1024);
  const int64_t b = round_up64((int64_t)B * packed_rows2(H, W) * D * es, 1024);
  return a + b;
}

int rc_build_pyramid(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int num_levels,
                     int vol_dtype, int math, void *pyramid, void *workspace, size_t workspace_bytes,
                     rc_stream_t stream) {
  RC_REQUIRE(fmap1 && fmap2 && pyramid, "rc_build_pyramid: NULL pointer");
  RC_REQUIRE(D >= 1, "rc_build_pyramid: D=%d", D);
  RC_REQUIRE(math == RC_MATH_F32 || math == RC_MATH_BF16 || math == RC_MATH_TF32 || math == RC_MATH_F16, "unknown math %d", math);
  rc_layout lay;
  int r = make_layout(B, H, W, num_levels, vol_dtype, &lay);
  if (r) return r;
  cudaStream_t s = (cudaStream_t)stream;
  if (math == RC_MATH_F32) return launch_build_simt(fmap1, fmap2, B, D, H, W, lay, vol_dtype, pyramid, s);
  r = check_tc_limits("rc_build_pyramid", D, num_levels);
  if (r) return r;
  const int64_t need = rc_build_workspace_bytes(B, D, H, W, math);
  if (workspace == nullptr || (int64_t)workspace_bytes < need) {
    set_error("rc_build_pyramid: workspace of %zu bytes, need %lld", workspace_bytes, (long long)need);
    return RC_ESIZE;
  }
  RC_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  RC_REQUIRE(((uintptr_t)pyramid & 255) == 0, "pyramid must be 256-byte aligned");
  const int64_t es = (math == RC_MATH_BF16 || math == RC_MATH_F16) ? 2 : 4;
  void *f1p = workspace;
  void *f2p = (char *)workspace + round_up64((int64_t)B * H * W * D * es, 1024);
  r = launch_pack(fmap1, fmap2, B, D, H, W, math, f1p, f2p, s);
  if (r) return r;
  return launch_build_tc(f1p, f2p, B, D, H, W, lay, vol_dtype, math, 0, pyramid, s);
}

int rc_build_pyramid_static(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                            int num_levels, int vol_dtype, void *pyramid, rc_stream_t stream) {
  RC_REQUIRE(fmap1 && fmap2 && pyramid, "rc_build_pyramid_static: NULL pointer");
  RC_REQUIRE(D >= 1, "rc_build_pyramid_static: D=%d", D);
  rc_layout lay;
  int r = make_layout(B, H, W, num_levels, vol_dtype, &lay);
  if (r) return r;
  r = check_tc_limits("rc_build_pyramid_static", D, num_levels);
  if (r) return r;
  const int64_t need = rc_build_workspace_bytes(B, D, H, W, RC_MATH_BF16);
  if (need < 0 || need > (int64_t)RC_STATIC_SCRATCH_BYTES) {
    set_error("rc_build_pyramid_static: B=%d D=%d H=%d W=%d needs %lld bytes of scratch, the built-in "
              "array holds %u; use rc_build_pyramid with a caller-owned workspace",
              B, D, H, W, (long long)need, (unsigned)RC_STATIC_SCRATCH_BYTES);
    return RC_ESIZE;
  }
  RC_REQUIRE(((uintptr_t)pyramid & 255) == 0, "pyramid must be 256-byte aligned");
  void *scratch = nullptr;
  r = static_scratch_ptr(&scratch);
  if (r) return r;
  cudaStream_t s = (cudaStream_t)stream;
  void *f1p = scratch;
  void *f2p = (char *)scratch + round_up64((int64_t)B * H * W * D * 2, 1024);
  r = launch_pack(fmap1, fmap2, B, D, H, W, RC_MATH_BF16, f1p, f2p, s);
  if (r) return r;
  return launch_build_tc(f1p, f2p, B, D, H, W, lay, vol_dtype, RC_MATH_BF16, 0, pyramid, s);
}

int rc_pack_features_nhwc16(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int operand_dtype,
                            void *fmap1_nhwc16, void *fmap2_nhwc16, rc_stream_t stream) {
  RC_REQUIRE(fmap1 && fmap2 && fmap1_nhwc16 && fmap2_nhwc16, "rc_pack_features_nhwc16: NULL pointer");
  RC_REQUIRE(B >= 1 && B <= 65535 && D >= 2 && D % 2 == 0 && H >= 1 && W >= 1, "rc_pack_features_nhwc16: bad sizes B=%d D=%d H=%d W=%d", B, D, H, W);
  RC_REQUIRE(operand_dtype == RC_VOL_BF16 || operand_dtype == RC_VOL_F16, "rc_pack_features_nhwc16: operand_dtype %d", operand_dtype);
  return launch_pack_nhwc16(fmap1, fmap2, B, D, H, W, operand_dtype == RC_VOL_F16, fmap1_nhwc16, fmap2_nhwc16,
                            (cudaStream_t)stream);
}

int rc_build_pyramid_nhwc16(const void *fmap1_nhwc16, const void *fmap2_nhwc16, int operand_dtype, int B, int D, int H,
                            int W, int num_levels, int vol_dtype, void *pyramid, rc_stream_t stream) {
  RC_REQUIRE(fmap1_nhwc16 && fmap2_nhwc16 && pyramid, "rc_build_pyramid_nhwc16: NULL pointer");
  RC_REQUIRE(operand_dtype == RC_VOL_BF16 || operand_dtype == RC_VOL_F16, "rc_build_pyramid_nhwc16: operand_dtype %d", operand_dtype);
  RC_REQUIRE(D >= 64 && D % 64 == 0, "rc_build_pyramid_nhwc16: D=%d must be a multiple of 64", D);
  RC_REQUIRE(((uintptr_t)fmap1_nhwc16 & 127) == 0 && ((uintptr_t)fmap2_nhwc16 & 127) == 0,
             "rc_build_pyramid_nhwc16: feature buffers must be 128-byte aligned");
  RC_REQUIRE(((uintptr_t)pyramid & 255) == 0, "pyramid must be 256-byte aligned");
  rc_layout lay;
  int r = make_layout(B, H, W, num_levels, vol_dtype, &lay);
  if (r) return r;
  r = check_tc_limits("rc_build_pyramid_nhwc16", D, num_levels);
  if (r) return r;
  return launch_build_tc(fmap1_nhwc16, fmap2_nhwc16, B, D, H, W, lay, vol_dtype,
                         operand_dtype == RC_VOL_F16 ? RC_MATH_F16 : RC_MATH_BF16, 1, pyramid, (cudaStream_t)stream);
}

int rc_lookup_fwd(const void *pyramid, int vol_dtype, const float *coords, int B, int H, int W,
                  int num_levels, int radius, float *out, rc_stream_t stream) {
  RC_REQUIRE(pyramid && coords && out, "rc_lookup_fwd: NULL pointer");
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  rc_layout lay;
  int r = make_layout(B, H, W, num_levels, vol_dtype, &lay);
  if (r) return r;
  return launch_lookup_fwd(pyramid, vol_dtype, coords, lay, H, W, radius, out, (cudaStream_t)stream);
}

int rc_lookup_fwd_f16(const void *pyramid, int vol_dtype, const float *coords, int B, int H, int W, int num_levels,
                      int radius, void *out_f16, rc_stream_t stream) {
  RC_REQUIRE(pyramid && coords && out_f16, "rc_lookup_fwd_f16: NULL pointer");
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  rc_layout lay;
  int r = make_layout(B, H, W, num_levels, vol_dtype, &lay);
  if (r) return r;
  return launch_lookup_fwd_f16(pyramid, vol_dtype, coords, lay, radius, out_f16, 0, (cudaStream_t)stream);
}

int64_t rc_convc1_weight_bytes(int O, int C) {
  if (O < 16 || O > 256 || O % 16 != 0 || C < 1) return RC_EINVAL;
  return (int64_t)O * convc1_cp(C) * 2;
}

int rc_convc1_pack_weights(const float *weight, int O, int C, void *weight_f16, rc_stream_t stream) {
  RC_REQUIRE(weight && weight_f16, "rc_convc1_pack_weights: NULL pointer");
  RC_REQUIRE(O >= 16 && O <= 256 && O % 16 == 0 && C >= 1, "rc_convc1_pack_weights: O=%d must be a multiple of 16 in 16..256, C=%d >= 1", O, C);
  return launch_convc1_pack_weights(weight, O, C, weight_f16, (cudaStream_t)stream);
}

int64_t rc_lookup_convc1_workspace_bytes(int B, int H, int W, int num_levels, int radius) {
  if (B < 1 || H < 1 || W < 1 || num_levels < 1 || num_levels > RC_MAX_LEVELS || radius < 1 || radius > 4) return RC_EINVAL;
  const int64_t C = (int64_t)num_levels * (2 * radius + 1) * (2 * radius + 1);
  return round_up64((int64_t)B * C * convc1_npad(H * W) * 2, 256);
}

int rc_lookup_convc1_fwd(const void *pyramid, int vol_dtype, const float *coords, int B, int H, int W, int num_levels,
                         int radius, const void *weight_f16, const float *bias, int O, void *out, int out_dtype,
                         void *workspace, size_t workspace_bytes, rc_stream_t stream) {
  RC_REQUIRE(pyramid && coords && weight_f16 && bias && out, "rc_lookup_convc1_fwd: NULL pointer");
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  RC_REQUIRE(O >= 16 && O <= 256 && O % 16 == 0, "rc_lookup_convc1_fwd: O=%d must be a multiple of 16 in 16..256", O);
  RC_REQUIRE(out_dtype == RC_VOL_F32 || out_dtype == RC_VOL_F16, "rc_lookup_convc1_fwd: out_dtype %d (RC_VOL_F32 or RC_VOL_F16)", out_dtype);
  rc_layout lay;
  int r = make_layout(B, H, W, num_levels, vol_dtype, &lay);
  if (r) return r;
  const int64_t need = rc_lookup_convc1_workspace_bytes(B, H, W, num_levels, radius);
  if (workspace == nullptr || (int64_t)workspace_bytes < need) {
    set_error("rc_lookup_convc1_fwd: workspace of %zu bytes, need %lld", workspace_bytes, (long long)need);
    return RC_ESIZE;
  }
  RC_REQUIRE(((uintptr_t)workspace & 127) == 0 && ((uintptr_t)weight_f16 & 127) == 0,
             "rc_lookup_convc1_fwd: workspace and weights must be 128-byte aligned");
  const int N = H * W, C = num_levels * (2 * radius + 1) * (2 * radius + 1);
  cudaStream_t s = (cudaStream_t)stream;
  r = launch_lookup_fwd_f16(pyramid, vol_dtype, coords, lay, radius, workspace, convc1_npad(N), s);
  if (r) return r;
  return launch_convc1(workspace, weight_f16, bias, B, N, C, O, out, out_dtype == RC_VOL_F16, s);
}

int rc_lookup_bwd(const void *pyramid, int vol_dtype, const float *coords, const float *grad_out, int B,
                  int H, int W, int num_levels, int radius, float *grad_pyramid, float *grad_coords,
                  rc_stream_t stream) {
  RC_REQUIRE(coords && grad_out, "rc_lookup_bwd: NULL pointer");
  RC_REQUIRE(grad_pyramid || grad_coords, "rc_lookup_bwd: nothing to compute");
  RC_REQUIRE(!grad_coords || pyramid, "rc_lookup_bwd: grad_coords needs the pyramid");
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  rc_layout lay, glay;
  int r = make_layout(B, H, W, num_levels, vol_dtype, &lay);
  if (r) return r;
  r = make_layout(B, H, W, num_levels, RC_VOL_F32, &glay);
  if (r) return r;
  return launch_lookup_bwd(pyramid, vol_dtype, coords, grad_out, lay, glay, H, W, radius, grad_pyramid,
                           grad_coords, (cudaStream_t)stream);
}

int64_t rc_build_bwd_workspace_bytes(int B, int D, int H, int W, int num_levels) {
  return build_bwd_workspace_bytes(B, D, H, W, num_levels);
}

int rc_build_bwd(const float *grad_pyramid, const float *fmap1, const float *fmap2, int B, int D, int H,
                 int W, int num_levels, int math, float *dfmap1, float *dfmap2, void *workspace,
                 size_t workspace_bytes, rc_stream_t stream) {
  RC_REQUIRE(grad_pyramid && fmap1 && fmap2 && dfmap1 && dfmap2, "rc_build_bwd: NULL pointer");
  RC_REQUIRE(B >= 1 && D >= 1, "rc_build_bwd: bad sizes B=%d D=%d", B, D);
  RC_REQUIRE(math == RC_MATH_F32 || math == RC_MATH_TF32, "rc_build_bwd: math %d (RC_MATH_F32 = 3xTF32 or RC_MATH_TF32)", math);
  rc_layout glay;
  int r = make_layout(B, H, W, num_levels, RC_VOL_F32, &glay);
  if (r) return r;
  const int64_t need = build_bwd_workspace_bytes(B, D, H, W, num_levels);
  if (need > 0 && (workspace == nullptr || (int64_t)workspace_bytes < need)) {
    set_error("rc_build_bwd: workspace of %zu bytes, need %lld", workspace_bytes, (long long)need);
    return RC_ESIZE;
  }
  return launch_build_bwd(grad_pyramid, fmap1, fmap2, B, D, H, W, glay, math == RC_MATH_F32, dfmap1, dfmap2,
                          workspace, (cudaStream_t)stream);
}

int rc_alt_fwd(const float *fmap1, const float *fmap2, const float *coords, int B, int H1, int W1, int H2,
               int W2, int C, int Nc, int radius, float *corr, rc_stream_t stream) {
  RC_REQUIRE(fmap1 && fmap2 && coords && corr, "rc_alt_fwd: NULL pointer");
  RC_REQUIRE(B >= 1 && H1 >= 1 && W1 >= 1 && H2 >= 1 && W2 >= 1 && Nc >= 1, "rc_alt_fwd: bad sizes");
  RC_REQUIRE(C >= 4 && C % 4 == 0 && C <= 512, "rc_alt_fwd: C=%d must be a multiple of 4, at most 512", C);
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  RC_REQUIRE(Nc <= 65535, "rc_alt_fwd: at most 65535 coordinate sets per call (got %d)", Nc);
  RC_REQUIRE((((uintptr_t)fmap1 | (uintptr_t)fmap2) & 15) == 0 && ((uintptr_t)coords & 7) == 0,
             "rc_alt_fwd: feature maps must be 16-byte aligned, coords 8-byte aligned");
  return launch_alt_fwd(fmap1, fmap2, coords, B, H1, W1, H2, W2, C, Nc, radius, corr, (cudaStream_t)stream);
}

int rc_alt_bwd(const float *fmap1, const float *fmap2, const float *coords, const float *corr_grad, int B,
               int H1, int W1, int H2, int W2, int C, int Nc, int radius, float *fmap1_grad,
               float *fmap2_grad, float *coords_grad, rc_stream_t stream) {
  RC_REQUIRE(fmap1 && fmap2 && coords && corr_grad && fmap1_grad && fmap2_grad, "rc_alt_bwd: NULL pointer");
  RC_REQUIRE(B >= 1 && H1 >= 1 && W1 >= 1 && H2 >= 1 && W2 >= 1 && Nc >= 1, "rc_alt_bwd: bad sizes");
  RC_REQUIRE(C >= 4 && C % 4 == 0 && C <= 512, "rc_alt_bwd: C=%d must be a multiple of 4, at most 512", C);
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  RC_REQUIRE((((uintptr_t)fmap1 | (uintptr_t)fmap2 | (uintptr_t)fmap1_grad | (uintptr_t)fmap2_grad) & 15) == 0 &&
                 ((uintptr_t)coords & 7) == 0,
             "rc_alt_bwd: feature maps and their gradients must be 16-byte aligned, coords 8-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  RC_CUDA(cudaMemsetAsync(fmap2_grad, 0, (size_t)B * H2 * W2 * C * sizeof(float), s));
  if (coords_grad) RC_CUDA(cudaMemsetAsync(coords_grad, 0, (size_t)B * Nc * H1 * W1 * 2 * sizeof(float), s));
  return launch_alt_bwd(fmap1, fmap2, coords, corr_grad, B, H1, W1, H2, W2, C, Nc, radius, fmap1_grad,
                        fmap2_grad, s);
}

int64_t rc_alt_feature_pyramid_elems(int B, int D, int H, int W, int num_levels, int64_t *f2_offset) {
  if (B < 1 || D < 1 || H < 1 || W < 1 || num_levels < 1 || num_levels > RC_MAX_LEVELS) return RC_EINVAL;
  int64_t off = 0;
  for (int l = 0; l < num_levels; ++l) {
    if (f2_offset) f2_offset[l] = off;
    off += round_up64((int64_t)B * (H >> l) * (W >> l) * D, 64);
  }
  return off;
}

int rc_alt_pool_features(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int num_levels,
                         float *fmap1_nhwc, float *fmap2_pyr, rc_stream_t stream) {
  RC_REQUIRE(fmap1 && fmap2 && fmap1_nhwc && fmap2_pyr, "rc_alt_pool_features: NULL pointer");
  RC_REQUIRE(D >= 4 && D % 4 == 0, "rc_alt_pool_features: D=%d must be a multiple of 4", D);
  int64_t off[RC_MAX_LEVELS];
  RC_REQUIRE(rc_alt_feature_pyramid_elems(B, D, H, W, num_levels, off) > 0, "rc_alt_pool_features: bad sizes");
  return launch_alt_pool(fmap1, fmap2, B, D, H, W, num_levels, off, fmap1_nhwc, fmap2_pyr, (cudaStream_t)stream);
}

int rc_alt_lookup_fwd(const float *fmap1_nhwc, const float *fmap2_pyr, const float *coords, int B, int D,
                      int H, int W, int num_levels, int radius, float *out, rc_stream_t stream) {
  RC_REQUIRE(fmap1_nhwc && fmap2_pyr && coords && out, "rc_alt_lookup_fwd: NULL pointer");
  RC_REQUIRE(D >= 4 && D % 4 == 0 && D <= 512, "rc_alt_lookup_fwd: D=%d must be a multiple of 4, at most 512", D);
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  int64_t off[RC_MAX_LEVELS];
  RC_REQUIRE(rc_alt_feature_pyramid_elems(B, D, H, W, num_levels, off) > 0, "rc_alt_lookup_fwd: bad sizes");
  RC_REQUIRE((((uintptr_t)fmap1_nhwc | (uintptr_t)fmap2_pyr) & 15) == 0, "rc_alt_lookup_fwd: feature buffers must be 16-byte aligned");
  RC_REQUIRE((H >> (num_levels - 1)) >= 1 && (W >> (num_levels - 1)) >= 1, "map too small for %d levels", num_levels);
  return launch_alt_lookup_fwd(fmap1_nhwc, fmap2_pyr, coords, B, D, H, W, num_levels, radius, off, out,
                               (cudaStream_t)stream);
}

int rc_alt_features_16(const float *fmap1_nhwc, const float *fmap2_pyr, int B, int D, int H, int W,
                       int num_levels, int operand_dtype, void *fmap1_16, void *fmap2_pyr_16, rc_stream_t stream) {
  RC_REQUIRE(fmap1_nhwc && fmap2_pyr && fmap1_16 && fmap2_pyr_16, "rc_alt_features_16: NULL pointer");
  RC_REQUIRE(D >= 4 && D % 4 == 0, "rc_alt_features_16: D=%d must be a multiple of 4", D);
  RC_REQUIRE(operand_dtype == RC_VOL_BF16 || operand_dtype == RC_VOL_F16, "rc_alt_features_16: operand_dtype %d", operand_dtype);
  int64_t off[RC_MAX_LEVELS];
  const int64_t total = rc_alt_feature_pyramid_elems(B, D, H, W, num_levels, off);
  RC_REQUIRE(total > 0, "rc_alt_features_16: bad sizes");
  const int half = operand_dtype == RC_VOL_F16;
  int r = launch_alt_features_16(fmap1_nhwc, fmap1_16, (long long)B * H * W * D, half, (cudaStream_t)stream);
  if (r) return r;
  return launch_alt_features_16(fmap2_pyr, fmap2_pyr_16, (long long)total, half, (cudaStream_t)stream);
}

int rc_alt_lookup_fwd_tc(const void *fmap1_bf16, const void *fmap2_pyr_bf16, int operand_dtype, const float *coords,
                         int B, int D, int H, int W, int num_levels, int radius, float *out, rc_stream_t stream) {
  RC_REQUIRE(fmap1_bf16 && fmap2_pyr_bf16 && coords && out, "rc_alt_lookup_fwd_tc: NULL pointer");
  RC_REQUIRE(operand_dtype == RC_VOL_BF16 || operand_dtype == RC_VOL_F16, "rc_alt_lookup_fwd_tc: operand_dtype %d", operand_dtype);
  RC_REQUIRE(D >= 64 && D % 64 == 0, "rc_alt_lookup_fwd_tc: D=%d must be a multiple of 64", D);
  RC_REQUIRE(radius >= 1 && radius <= 4, "radius %d unsupported (1..4)", radius);
  RC_REQUIRE(((uintptr_t)fmap1_bf16 & 127) == 0 && ((uintptr_t)fmap2_pyr_bf16 & 127) == 0,
             "rc_alt_lookup_fwd_tc: feature buffers must be 128-byte aligned");
  int64_t off[RC_MAX_LEVELS];
  RC_REQUIRE(rc_alt_feature_pyramid_elems(B, D, H, W, num_levels, off) > 0, "rc_alt_lookup_fwd_tc: bad sizes");
  RC_REQUIRE((H >> (num_levels - 1)) >= 1 && (W >> (num_levels - 1)) >= 1, "map too small for %d levels", num_levels);
  return launch_alt_lookup_tc(fmap1_bf16, fmap2_pyr_bf16, coords, B, D, H, W, num_levels, radius, off,
                              operand_dtype == RC_VOL_F16, out, (cudaStream_t)stream);
}

int rc_upsample_flow(const float *flow, const float *mask, int B, int H, int W, float *out, rc_stream_t stream) {
  RC_REQUIRE(flow && mask && out, "rc_upsample_flow: NULL pointer");
  RC_REQUIRE(B >= 1 && B <= 65535 && H >= 1 && H <= 65535 && W >= 1, "rc_upsample_flow: bad sizes B=%d H=%d W=%d", B, H, W);
  return launch_upsample_flow(flow, mask, 0, B, H, W, out, (cudaStream_t)stream);
}

int rc_upsample_flow_f16mask(const float *flow, const void *mask_f16, int B, int H, int W, float *out,
                             rc_stream_t stream) {
  RC_REQUIRE(flow && mask_f16 && out, "rc_upsample_flow_f16mask: NULL pointer");
  RC_REQUIRE(B >= 1 && B <= 65535 && H >= 1 && H <= 65535 && W >= 1, "rc_upsample_flow_f16mask: bad sizes B=%d H=%d W=%d", B, H, W);
  return launch_upsample_flow(flow, mask_f16, 1, B, H, W, out, (cudaStream_t)stream);
}

int rc_upsample_flow_bwd(const float *flow, const float *mask, const float *gout, int B, int H, int W, float *dflow,
                         float *dmask, rc_stream_t stream) {
  RC_REQUIRE(flow && mask && gout && dflow && dmask, "rc_upsample_flow_bwd: NULL pointer");
  RC_REQUIRE(B >= 1 && B <= 65535 && H >= 1 && H <= 65535 && W >= 1, "rc_upsample_flow_bwd: bad sizes B=%d H=%d W=%d", B, H, W);
  return launch_upsample_flow_bwd(flow, mask, gout, B, H, W, dflow, dmask, (cudaStream_t)stream);
}

int rc_upflow8(const float *flow, int B, int C, int H, int W, float *out, rc_stream_t stream) {
  RC_REQUIRE(flow && out, "rc_upflow8: NULL pointer");
  RC_REQUIRE(B >= 1 && C >= 1 && (long long)B * C <= 65535 && H >= 1 && 8 * (long long)H <= 65535 && W >= 1,
             "rc_upflow8: bad sizes B=%d C=%d H=%d W=%d", B, C, H, W);
  return launch_upflow8(flow, B, C, H, W, out, (cudaStream_t)stream);
}

int rc_upflow8_bwd(const float *gout, int B, int C, int H, int W, float *dflow, rc_stream_t stream) {
  RC_REQUIRE(gout && dflow, "rc_upflow8_bwd: NULL pointer");
  RC_REQUIRE(B >= 1 && C >= 1 && (long long)B * C <= 65535 && H >= 1 && 8 * (long long)H <= 65535 && W >= 1,
             "rc_upflow8_bwd: bad sizes B=%d C=%d H=%d W=%d", B, C, H, W);
  return launch_upflow8_bwd(gout, B, C, H, W, dflow, (cudaStream_t)stream);
}

}  // extern "C"
