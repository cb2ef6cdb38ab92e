// Exact-fp32 volume build on the CUDA cores (RC_MATH_F32).
//
// This is the arithmetic the reference gets from torch.matmul on fp32 inputs
// with TF32 off (core/corr.py:58): fp32 FMA accumulation.  It exists for users
// who need the reference's precision bit-for-class (tolerance 1e-4 and better)
// and as the in-library cross-check of the tcgen05 build; the fast path is
// build_tc.cu.  Level 0 is written by a 64x64-tile SGEMM that reads the NCHW
// maps in place (both operands are pixel-contiguous, so no repack is needed);
// levels 1.. are pooled by a second kernel.  The pyramid buffer is cleared
// first so that the padding inside edge tiles is zero (rc_layout's invariant).
#include "common.cuh"

namespace rc {
namespace {

constexpr int SM_T = 64;   // tile edge (queries x targets)
constexpr int SM_K = 16;

template <typename VolT>
__device__ __forceinline__ VolT to_vol(float v);
template <>
__device__ __forceinline__ float to_vol<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 to_vol<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ __half to_vol<__half>(float v) { return __float2half_rn(clamp_half(v)); }

__device__ __forceinline__ float from_vol(float v) { return v; }
__device__ __forceinline__ float from_vol(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float from_vol(__half v) { return __half2float(v); }

// grid: (ceil(N/64) targets, ceil(N/64) queries, B); block 256 = 16x16, 4x4 outputs each
template <typename VolT>
__global__ void __launch_bounds__(256)
build_level0_simt_kernel(const float *__restrict__ fmap1, const float *__restrict__ fmap2,
                         VolT *__restrict__ lvl0, int D, int N, int W, int tiles_x, long long slab,
                         float scale) {
  __shared__ float As[SM_K][SM_T];
  __shared__ float Bs[SM_K][SM_T];
  const int b = blockIdx.z;
  const int q0 = blockIdx.y * SM_T;
  const int t0 = blockIdx.x * SM_T;
  const float *A = fmap1 + (size_t)b * D * N;
  const float *Bm = fmap2 + (size_t)b * D * N;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;

  float acc[4][4] = {};
  for (int k0 = 0; k0 < D; k0 += SM_K) {
    // 16 x 64 floats per operand, 256 threads -> 4 each
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int idx = threadIdx.x + 256 * i;
      int kk = idx >> 6, c = idx & 63;
      int k = k0 + kk;
      As[kk][c] = (k < D && q0 + c < N) ? __ldg(A + (size_t)k * N + q0 + c) : 0.f;
      Bs[kk][c] = (k < D && t0 + c < N) ? __ldg(Bm + (size_t)k * N + t0 + c) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SM_K; ++kk) {
      float a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bb[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int q = q0 + ty * 4 + i;
    if (q >= N) continue;
    VolT *row = lvl0 + ((size_t)b * N + q) * slab;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int t = t0 + tx * 4 + j;
      if (t < N) row[tile_offset<VolT>(t / W, t % W, tiles_x)] = to_vol<VolT>(acc[i][j] * scale);
    }
  }
}

// level l from level l-1: one thread per output element
template <typename VolT>
__global__ void __launch_bounds__(256)
pool_level_kernel(const VolT *__restrict__ src, VolT *__restrict__ dst, long long slabs, int Hd,
                  int Wd, int tiles_x_s, long long slab_s, int tiles_x_d, long long slab_d) {
  const long long per = (long long)Hd * Wd;
  const long long total = slabs * per;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long s = i / per;
    int r = (int)(i - s * per);
    int y = r / Wd, x = r - y * Wd;
    const VolT *p = src + s * slab_s;
    float v = 0.25f * (from_vol(p[tile_offset<VolT>(2 * y, 2 * x, tiles_x_s)]) +
                       from_vol(p[tile_offset<VolT>(2 * y, 2 * x + 1, tiles_x_s)]) +
                       from_vol(p[tile_offset<VolT>(2 * y + 1, 2 * x, tiles_x_s)]) +
                       from_vol(p[tile_offset<VolT>(2 * y + 1, 2 * x + 1, tiles_x_s)]));
    dst[s * slab_d + tile_offset<VolT>(y, x, tiles_x_d)] = to_vol<VolT>(v);
  }
}

template <typename VolT>
int build_simt_t(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                 const rc_layout &lay, VolT *pyr, cudaStream_t s) {
  const int N = H * W;
  RC_CUDA(cudaMemsetAsync(pyr, 0, (size_t)lay.total_elems * sizeof(VolT), s));
  dim3 grid(ceil_div(N, SM_T), ceil_div(N, SM_T), B);
  build_level0_simt_kernel<VolT><<<grid, 256, 0, s>>>(fmap1, fmap2, pyr + lay.offset[0], D, N, W,
                                                      lay.tiles_x[0], lay.slab[0],
                                                      1.0f / sqrtf((float)D));
  RC_LAUNCH_CHECK("build_level0_simt_kernel");
  return RC_OK;
}

template <typename VolT>
int pool_levels_t(const rc_layout &lay, VolT *pyr, int first_level, cudaStream_t s) {
  const long long slabs = (long long)lay.B * lay.N;
  for (int l = first_level; l < lay.num_levels; ++l) {
    long long total = slabs * lay.H[l] * lay.W[l];
    if (total == 0) continue;
    int blocks = (int)((total + 255) / 256 < 148LL * 32 ? (total + 255) / 256 : 148LL * 32);
    pool_level_kernel<VolT><<<blocks, 256, 0, s>>>(pyr + lay.offset[l - 1], pyr + lay.offset[l],
                                                   slabs, lay.H[l], lay.W[l], lay.tiles_x[l - 1],
                                                   lay.slab[l - 1], lay.tiles_x[l], lay.slab[l]);
    RC_LAUNCH_CHECK("pool_level_kernel");
  }
  return RC_OK;
}

}  // namespace

int launch_build_simt(const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                      const rc_layout &lay, int vol_dtype, void *pyramid, cudaStream_t s) {
  int rcode = vol_dtype == RC_VOL_BF16  ? build_simt_t<__nv_bfloat16>(fmap1, fmap2, B, D, H, W, lay, (__nv_bfloat16 *)pyramid, s)
              : vol_dtype == RC_VOL_F16 ? build_simt_t<__half>(fmap1, fmap2, B, D, H, W, lay, (__half *)pyramid, s)
                                        : build_simt_t<float>(fmap1, fmap2, B, D, H, W, lay, (float *)pyramid, s);
  if (rcode != RC_OK) return rcode;
  return launch_pool_levels(lay, vol_dtype, pyramid, 1, s);
}

int launch_pool_levels(const rc_layout &lay, int vol_dtype, void *pyramid, int first_level,
                       cudaStream_t s) {
  return vol_dtype == RC_VOL_BF16  ? pool_levels_t<__nv_bfloat16>(lay, (__nv_bfloat16 *)pyramid, first_level, s)
         : vol_dtype == RC_VOL_F16 ? pool_levels_t<__half>(lay, (__half *)pyramid, first_level, s)
                                   : pool_levels_t<float>(lay, (float *)pyramid, first_level, s);
}

}  // namespace rc
