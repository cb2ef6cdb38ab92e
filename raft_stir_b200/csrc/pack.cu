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

// 16-bit outputs: one block = 64 consecutive output rows x ALL channels, so that every output row (D * 2 bytes, 512 for
// D = 256) leaves whole, as consecutive 128-byte stores of one warp, and every channel is read in 256-byte runs.  (The
// 64-channel tiles above write 128-byte pieces of a row from four blocks that run at different times.)  The transpose
// goes through shared memory as packed channel PAIRS: tile[c][r] = (in[2c][r], in[2c+1][r]), row pitch 65 words, both
// the pixel-major writes and the channel-major reads conflict-free.  D <= 256, D even.
constexpr int PW_ROWS = 64;
constexpr int PW_MAXPAIRS = 128;

template <typename OutT>
__global__ void __launch_bounds__(256)
pack_rows16_kernel(const float *__restrict__ fmap1, const float *__restrict__ fmap2, OutT *__restrict__ f1p,
                   OutT *__restrict__ f2p, int D, int H, int W, int groups1, int nbx, long long rows2, float scale1) {
  __shared__ uint32_t tile[PW_MAXPAIRS][PW_ROWS + 1];
  const int N = H * W;
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const bool second = (int)blockIdx.x >= groups1;
  const int g = second ? blockIdx.x - groups1 : blockIdx.x;
  const long long row0 = (long long)g * PW_ROWS;
  const int pairs = D >> 1;

  // source pixels of this lane's two rows (lane, lane + 32)
  int src[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    src[h] = -1;
    const long long r = row0 + 32 * h + lane;
    if (!second) {
      if (r < N) src[h] = (int)r;
    } else {
      const int blk = (int)(r / TB_N), rr = (int)(r % TB_N);
      const int y = (blk / nbx) * TB_H + rr / TB_W, x = (blk % nbx) * TB_W + rr % TB_W;
      if (r < rows2 && y < H && x < W) src[h] = y * W + x;
    }
  }
  const float *in = (second ? fmap2 : fmap1) + (size_t)b * D * N;
  const float mul = second ? 1.f : scale1;     // exact power of two (common.cuh: split_scale)
  for (int c0 = 0; c0 < pairs; c0 += 32) {     // 4 channel pairs per warp and round, 16 loads in flight per thread
    float v[4][2][2];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = c0 + wid + 8 * k;
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int e = 0; e < 2; ++e)
          v[k][h][e] = (c < pairs && src[h] >= 0) ? __ldg(in + (size_t)(2 * c + e) * N + src[h]) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = c0 + wid + 8 * k;
      if (c < pairs) {
#pragma unroll
        for (int h = 0; h < 2; ++h) tile[c][32 * h + lane] = pack2<OutT>(mul * v[k][h][0], mul * v[k][h][1]);
      }
    }
  }
  __syncthreads();
  OutT *out = second ? f2p + ((size_t)b * rows2 + row0) * D : f1p + ((size_t)b * N + row0) * D;
  const long long limit = second ? rows2 : (long long)N;
#pragma unroll 2
  for (int k = 0; k < PW_ROWS / 8; ++k) {
    const int r = wid + 8 * k;
    if (row0 + r >= limit) continue;
    uint32_t *o = reinterpret_cast<uint32_t *>(out + (size_t)r * D);
    for (int c = lane; c < pairs; c += 32) o[c] = tile[c][r];
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
  if ((math == RC_MATH_BF16 || math == RC_MATH_F16) && D % 2 == 0 && D <= 2 * PW_MAXPAIRS) {
    const int g1 = ceil_div(N, PW_ROWS), g2 = (int)(rows2 / PW_ROWS);
    dim3 gw(g1 + g2, B);
    if (math == RC_MATH_BF16)
      pack_rows16_kernel<__nv_bfloat16><<<gw, 256, 0, s>>>(fmap1, fmap2, (__nv_bfloat16 *)f1p, (__nv_bfloat16 *)f2p, D, H, W,
                                                           g1, nbx, rows2, pre);
    else
      pack_rows16_kernel<__half><<<gw, 256, 0, s>>>(fmap1, fmap2, (__half *)f1p, (__half *)f2p, D, H, W, g1, nbx, rows2, pre);
  } else if (math == RC_MATH_BF16) {
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
