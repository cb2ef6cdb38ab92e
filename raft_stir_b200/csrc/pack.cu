// Feature repack for the tensor-core volume build.
//
// RAFT hands the correlation block two fp32 NCHW maps (core/raft.py:102-107):
// for the contraction over D both are "pixel-contiguous, D-strided".  The
// tcgen05 build wants K-major operands (D contiguous per pixel) in bf16 (or
// fp32 for kind::tf32), and wants fmap2's pixels ordered by 8x32 spatial block
// so that a 256-column GEMM tile is one block (common.cuh: TB_H, TB_W).
//
//   f1p[b][q][d]                      q = y*W + x                 (N rows), times the
//                                          exact power-of-two part of 1/sqrt(D)
//   f2p[b][blk*256 + (y%8)*32 + x%32][d]   blk = (y/8)*nbx + x/32   (nblk*256 rows,
//                                          rows outside the map are zero)
//
// One launch packs both maps.  HBM-bound: reads 2*B*D*N*4 bytes once, writes
// half (bf16) or the same (fp32).
#include "common.cuh"

namespace rc {
namespace {

constexpr int PK_ROWS = 32;   // pixels per block (one x-run)
constexpr int PK_D = 64;      // channels per block

template <typename OutT>
__device__ __forceinline__ void store_pair(OutT *p, float a, float b);
template <>
__device__ __forceinline__ void store_pair<__nv_bfloat16>(__nv_bfloat16 *p, float a, float b) {
  *reinterpret_cast<__nv_bfloat162 *>(p) = __floats2bfloat162_rn(a, b);
}
template <>
__device__ __forceinline__ void store_pair<__half>(__half *p, float a, float b) {
  *reinterpret_cast<uint32_t *>(p) = pack2<__half>(a, b);     // saturating
}
// fp32 operands feed kind::tf32, which reads the top 19 bits and ignores the rest (truncation): round to nearest here
__device__ __forceinline__ float round_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}
template <>
__device__ __forceinline__ void store_pair<float>(float *p, float a, float b) {
  *reinterpret_cast<float2 *>(p) = make_float2(round_tf32(a), round_tf32(b));
}

// grid: (row_groups_f1 + row_groups_f2, D/PK_D, B); block: 256 threads
template <typename OutT>
__global__ void __launch_bounds__(256)
pack_features_kernel(const float *__restrict__ fmap1, const float *__restrict__ fmap2,
                     OutT *__restrict__ f1p, OutT *__restrict__ f2p, int D, int H, int W,
                     int groups1, int nbx, long long rows2, float scale1) {
  __shared__ float tile[PK_D][PK_ROWS + 1];
  const int N = H * W;
  const int b = blockIdx.z;
  const int d0 = blockIdx.y * PK_D;
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;   // 0..7

  const bool second = blockIdx.x >= groups1;
  const int g = second ? blockIdx.x - groups1 : blockIdx.x;
  const long long row0 = (long long)g * PK_ROWS;

  // source pixel of this lane's row
  int src = -1;
  if (!second) {
    long long q = row0 + lane;
    if (q < N) src = (int)q;
  } else {
    int blk = (int)(row0 / TB_N);
    int rr = (int)(row0 % TB_N) + lane;   // row0 is a multiple of 32 -> same y for all lanes
    int y = (blk / nbx) * TB_H + rr / TB_W;
    int x = (blk % nbx) * TB_W + rr % TB_W;
    if (y < H && x < W) src = y * W + x;
  }
  const float *in = (second ? fmap2 : fmap1) + (size_t)b * D * N;
  const float mul = second ? 1.f : scale1;     // exact power of two (common.cuh: split_scale)
#pragma unroll
  for (int k = 0; k < PK_D / 8; ++k) {
    int d = wid + 8 * k;
    float v = 0.f;
    if (src >= 0 && d0 + d < D) v = mul * __ldg(in + (size_t)(d0 + d) * N + src);
    tile[d][lane] = v;
  }
  __syncthreads();

  OutT *out = second ? f2p + ((size_t)b * rows2 + row0) * D : f1p + ((size_t)b * N + row0) * D;
  const long long limit = second ? rows2 : (long long)N;
#pragma unroll
  for (int k = 0; k < PK_ROWS / 8; ++k) {
    int r = wid + 8 * k;
    if (row0 + r < limit && d0 + 2 * lane < D)
      store_pair<OutT>(out + (size_t)r * D + d0 + 2 * lane, tile[2 * lane][r], tile[2 * lane + 1][r]);
  }
}

}  // namespace

int launch_pack(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int math,
                void *f1p, void *f2p, cudaStream_t s) {
  const int N = H * W;
  const int nbx = ceil_div(W, TB_W), nby = ceil_div(H, TB_H);
  const long long rows2 = (long long)nbx * nby * TB_N;
  const int groups1 = ceil_div(N, PK_ROWS);
  const int groups2 = (int)(rows2 / PK_ROWS);
  dim3 grid(groups1 + groups2, ceil_div(D, PK_D), B);
  float pre = 1.f, post = 1.f;
  split_scale(D, &pre, &post);
  if (math == RC_MATH_BF16) {
    pack_features_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(
        fmap1, fmap2, (__nv_bfloat16 *)f1p, (__nv_bfloat16 *)f2p, D, H, W, groups1, nbx, rows2, pre);
  } else if (math == RC_MATH_F16) {
    pack_features_kernel<__half><<<grid, 256, 0, s>>>(fmap1, fmap2, (__half *)f1p, (__half *)f2p, D, H, W, groups1,
                                                       nbx, rows2, pre);
  } else {
    pack_features_kernel<float><<<grid, 256, 0, s>>>(fmap1, fmap2, (float *)f1p, (float *)f2p, D, H,
                                                     W, groups1, nbx, rows2, pre);
  }
  RC_LAUNCH_CHECK("pack_features_kernel");
  return RC_OK;
}

// fp32 NCHW -> 16-bit NHWC [B][H][W][D] for both maps (the layout a channels_last 16-bit encoder emits): the same
// transposing kernel with every row group in plain pixel order and no scale.
int launch_pack_nhwc16(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int half_operands,
                       void *f1_nhwc, void *f2_nhwc, cudaStream_t s) {
  const int N = H * W;
  const int groups = ceil_div(N, PK_ROWS);
  dim3 grid(groups, ceil_div(D, PK_D), B);
  for (int which = 0; which < 2; ++which) {
    const float *src = which ? fmap2 : fmap1;
    void *dst = which ? f2_nhwc : f1_nhwc;
    if (half_operands)
      pack_features_kernel<__half><<<grid, 256, 0, s>>>(src, nullptr, (__half *)dst, nullptr, D, H, W, groups, 1, 0, 1.f);
    else
      pack_features_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(src, nullptr, (__nv_bfloat16 *)dst, nullptr, D, H, W,
                                                               groups, 1, 0, 1.f);
    RC_LAUNCH_CHECK("pack_features_kernel (NHWC)");
  }
  return RC_OK;
}

}  // namespace rc
