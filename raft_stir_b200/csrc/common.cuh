// Shared declarations for libraftcorr's translation units (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <stdio.h>

#include "../../include/raft_corr.h"

namespace rc {

// ---- error reporting -------------------------------------------------------
void set_error(const char *fmt, ...);   // api.cu; thread-local message

#define RC_REQUIRE(cond, ...)            \
  do {                                   \
    if (!(cond)) {                       \
      rc::set_error(__VA_ARGS__);        \
      return RC_EINVAL;                  \
    }                                    \
  } while (0)

#define RC_CUDA(call)                                                              \
  do {                                                                             \
    cudaError_t e__ = (call);                                                      \
    if (e__ != cudaSuccess) {                                                      \
      rc::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),       \
                    __FILE__, __LINE__);                                           \
      return RC_ECUDA;                                                             \
    }                                                                              \
  } while (0)

#define RC_LAUNCH_CHECK(name)                                                      \
  do {                                                                             \
    cudaError_t e__ = cudaGetLastError();                                          \
    if (e__ != cudaSuccess) {                                                      \
      rc::set_error("launch of %s failed: %s", name, cudaGetErrorString(e__));     \
      return RC_ECUDA;                                                             \
    }                                                                              \
  } while (0)

// ---- small helpers ---------------------------------------------------------
__host__ __device__ inline int64_t round_up64(int64_t v, int64_t m) { return (v + m - 1) / m * m; }
__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// 1/sqrt(D) = pre * post with pre an exact power of two (applied to fmap1 by the
// pack: scaling bf16 by a power of two commutes with rounding) and post the rest,
// applied to the fp32 accumulators.  For D a power of 4 (RAFT-basic's 256) post
// is exactly 1 and the build's epilogue skips the multiply.
inline void split_scale(int D, float *pre, float *post) {
  int k = 0;
  while ((1 << (2 * (k + 1))) <= D) ++k;          // 4^k <= D
  if ((1 << (2 * k)) == D) {
    *pre = 1.0f / (float)(1 << k);
    *post = 1.0f;
  } else {
    *pre = 1.0f;
    *post = 1.0f / sqrtf((float)D);
  }
}

// Device-side copy of the layout: passed by value to kernels.
struct Levels {
  int num_levels;
  int H[RC_MAX_LEVELS];
  int W[RC_MAX_LEVELS];
  int tiles_y[RC_MAX_LEVELS];
  int tiles_x[RC_MAX_LEVELS];
  long long slab[RC_MAX_LEVELS];
  long long offset[RC_MAX_LEVELS];
};

// Pyramid tiles (rc_layout in raft_corr.h): TILE_H rows of 16 bytes = 64 bytes.
constexpr int TILE_H = 4;
template <typename VolT>
struct Tile {
  static constexpr int W = 16 / (int)sizeof(VolT);     // elements per 16-byte tile row
  static constexpr int LOG_W = sizeof(VolT) == 2 ? 3 : 2;
  static constexpr int ELEMS = TILE_H * W;             // elements per tile
};
// Conversions between fp32 and the stored element type: two 16-bit elements per 32-bit word (first element low).
template <typename VolT>
__device__ __forceinline__ uint32_t pack2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 p = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&p);
}
// fp16 storage saturates at the largest finite half instead of overflowing to infinity
__device__ __forceinline__ float clamp_half(float v) { return fminf(fmaxf(v, -65504.f), 65504.f); }
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  __half2 p = __floats2half2_rn(clamp_half(a), clamp_half(b));
  return *reinterpret_cast<uint32_t *>(&p);
}
template <>
__device__ __forceinline__ uint32_t pack2<float>(float a, float) { return __float_as_uint(a); }   // unused for 4-byte storage
template <typename VolT>
__device__ __forceinline__ float2 unpack2(uint32_t w);
template <>
__device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
template <>
__device__ __forceinline__ float2 unpack2<__half>(uint32_t w) { return __half22float2(*reinterpret_cast<__half2 *>(&w)); }
template <>
__device__ __forceinline__ float2 unpack2<float>(uint32_t w) { return make_float2(__uint_as_float(w), 0.f); }   // unused

// element offset of (y, x) inside one query's slab
template <typename VolT>
__host__ __device__ inline long long tile_offset(int y, int x, int tiles_x) {
  return ((long long)(y >> 2) * tiles_x + (x >> Tile<VolT>::LOG_W)) * Tile<VolT>::ELEMS +
         (y & 3) * Tile<VolT>::W + (x & (Tile<VolT>::W - 1));
}

inline Levels to_levels(const rc_layout &lay) {
  Levels lv{};
  lv.num_levels = lay.num_levels;
  for (int l = 0; l < lay.num_levels; ++l) {
    lv.H[l] = lay.H[l];
    lv.W[l] = lay.W[l];
    lv.tiles_y[l] = lay.tiles_y[l];
    lv.tiles_x[l] = lay.tiles_x[l];
    lv.slab[l] = lay.slab[l];
    lv.offset[l] = lay.offset[l];
  }
  return lv;
}

// Geometry of the target (fmap2) blocking used by the tensor-core build:
// targets are grouped in spatial blocks of TB_H rows x TB_W columns so that one
// GEMM N-tile (256 columns) is one block and the 2x2/4x4/8x8 pooling of the
// pyramid stays inside a tile.
constexpr int TB_H = 8;
constexpr int TB_W = 32;
constexpr int TB_N = TB_H * TB_W;   // 256 targets per block

// ---- launchers implemented in the kernel files ----------------------------
// pack.cu
int launch_pack(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int math,
                void *f1p, void *f2p, cudaStream_t s);
// build_simt.cu
int launch_build_simt(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                      const rc_layout &lay, int vol_dtype, void *pyramid, cudaStream_t s);
int launch_pool_levels(const rc_layout &lay, int vol_dtype, void *pyramid, int first_level,
                       cudaStream_t s);
// build_tc.cu
int launch_build_tc(const void *f1p, const void *f2p, int B, int D, int H, int W,
                    const rc_layout &lay, int vol_dtype, int math, int nhwc, void *pyramid, cudaStream_t s);
// pack.cu: fp32 NCHW -> 16-bit NHWC, both maps, unscaled
int launch_pack_nhwc16(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int half_operands,
                       void *f1_nhwc, void *f2_nhwc, cudaStream_t s);
// lookup.cu
int launch_lookup_fwd_f16(const void *pyramid, int vol_dtype, const float *coords, const rc_layout &lay, int radius,
                          void *out_f16, int out_pitch, cudaStream_t s);
// convc1_tc.cu
int convc1_cp(int C);
int convc1_npad(int N);
int launch_convc1_pack_weights(const float *w, int O, int C, void *w16, cudaStream_t s);
int launch_convc1(const void *samples, const void *w16, const float *bias, int B, int N, int C, int O, void *out,
                  int out_f16, cudaStream_t s);
int launch_lookup_fwd(const void *pyramid, int vol_dtype, const float *coords, const rc_layout &lay,
                      int H, int W, int radius, float *out, cudaStream_t s);
int launch_lookup_bwd(const void *pyramid, int vol_dtype, const float *coords, const float *grad_out,
                      const rc_layout &lay, const rc_layout &glay, int H, int W, int radius,
                      float *grad_pyramid, float *grad_coords, cudaStream_t s);

}  // namespace rc
