// Flow upsampling to full resolution: the step that follows the correlation lookup + update block in
// every refinement iteration.
//
//   convex upsampling  RAFT.upsample_flow            core/raft.py:72-83
//   bilinear x8        upflow8 (RAFT-small, no mask)  core/utils/utils.py:80-82
//
// The reference spells the convex combination as softmax + F.unfold + broadcast multiply + sum +
// permute + reshape: six kernels and four full-size temporaries for something that reads the
// 576-channel mask once and writes 128 floats per coarse pixel.  Here one kernel does it: a CTA takes
// 32 coarse pixels of one row, thread (x, sx) walks sy, reads its 9 mask values (warp = 32 consecutive
// x of one channel: 128-byte coalesced), forms the softmax in registers, combines the 3x3 neighbours
// of 8*flow it holds in registers, and parks the two results in a skewed shared-memory tile from which
// the 16 output rows (2 channels x 8 sub-rows, 256 floats each) leave as full coalesced lines.
// HBM-bound: 2 304 B of mask in + 512 B out per coarse pixel.  Backward: same tiling, softmax
// recomputed, d/dmask written coalesced, d/dflow folded over sub-pixels in registers, over sx through
// shared memory, then 18 atomics per coarse pixel onto the 3x3 neighbourhood.
#include "common.cuh"

namespace rc {
namespace {

constexpr int UP_TX = 32;                 // coarse pixels per CTA (one warp lane each)
constexpr int UP_ROW = 8 * UP_TX + 8;     // skewed tile row: position p of 256 sits at p + p / 32

__device__ __forceinline__ void load_neighbours(const float *__restrict__ flow, int b, int y, int x, int H, int W,
                                                bool in, float (&nb)[2][9]) {
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
      const bool ok = in && yy >= 0 && yy < H && xx >= 0 && xx < W;          // F.unfold(padding=1): zeros outside
      nb[c][k] = ok ? 8.0f * __ldg(flow + (((size_t)b * 2 + c) * H + yy) * W + xx) : 0.f;
    }
}

// mask element loads: fp32, or fp16 as fp16 autocast leaves the update block's output (core/raft.py:127-128)
__device__ __forceinline__ float ld_mask(const float *p) { return __ldg(p); }
__device__ __forceinline__ float ld_mask(const __half *p) { return __half2float(__ldg(p)); }
__device__ __forceinline__ float4 ld_mask4(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ float4 ld_mask4(const __half *p) {
  const uint2 u = __ldg(reinterpret_cast<const uint2 *>(p));
  const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ void softmax9(float (&m)[9]) {
  float mx = m[0];
#pragma unroll
  for (int k = 1; k < 9; ++k) mx = fmaxf(mx, m[k]);
  float sum = 0.f;
  // exp(m - mx) as 2^((m - mx) * log2 e) on the SFU: arguments are <= 0, the relative error (2^-22 plus
  // |m - mx| * 2^-24 from the product) is far below the 1e-5 the fp32 result is compared at
#pragma unroll
  for (int k = 0; k < 9; ++k) { m[k] = ex2_approx((m[k] - mx) * 1.4426950408889634f); sum += m[k]; }
  const float inv = __frcp_rn(sum);
#pragma unroll
  for (int k = 0; k < 9; ++k) m[k] *= inv;
}

template <typename MT>
__global__ void __launch_bounds__(256)
upsample_flow_kernel(const float *__restrict__ flow, const MT *__restrict__ mask, float *__restrict__ out, int H,
                     int W) {
  __shared__ float tile[2][8][UP_ROW];
  const int lane = threadIdx.x & 31, sx = threadIdx.x >> 5;
  const int x0 = blockIdx.x * UP_TX, y = blockIdx.y, b = blockIdx.z;
  const int x = x0 + lane;
  const bool in = x < W;
  const size_t HW = (size_t)H * W;
  float nb[2][9];
  load_neighbours(flow, b, y, x, H, W, in, nb);
  const MT *mp = mask + (size_t)b * 576 * HW + (size_t)y * W + (in ? x : 0);
#pragma unroll 2
  for (int sy = 0; sy < 8; ++sy) {
    float m[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) m[k] = in ? ld_mask(mp + (size_t)(k * 64 + sy * 8 + sx) * HW) : 0.f;
    softmax9(m);
    float o0 = 0.f, o1 = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      o0 = fmaf(m[k], nb[0][k], o0);
      o1 = fmaf(m[k], nb[1][k], o1);
    }
    const int p = 8 * lane + sx;
    tile[0][sy][p + (p >> 5)] = o0;
    tile[1][sy][p + (p >> 5)] = o1;
  }
  __syncthreads();
  // 16 output rows of 256 floats; thread t writes column t of each (coalesced 128-byte lines)
  const int p = threadIdx.x;
  if (x0 + (p >> 3) < W) {
    const size_t W8 = (size_t)8 * W;
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int c = r >> 3, sy = r & 7;
      out[(((size_t)b * 2 + c) * 8 * H + 8 * y + sy) * W8 + (size_t)8 * x0 + p] = tile[c][sy][p + (p >> 5)];
    }
  }
}

// The same for W % 4 == 0 (Sintel, 1080p; every width a multiple of 32 pixels): 512 threads, thread = (four
// consecutive coarse pixels, sx, sy), so the nine mask values of four pixels arrive as nine 16-byte loads issued back
// to back (a warp reads 128 contiguous bytes of four channel planes per instruction), and the neighbourhood of
// 8*flow is staged once per CTA in shared memory instead of 18 registers per thread: 64 registers, two CTAs per SM,
// enough bytes in flight to keep HBM busy while the other CTA computes and stores.
constexpr int UPV_ROW = 8 * UP_TX + 5 * 8;   // skew 5 per 32 positions

template <typename MT>
__global__ void __launch_bounds__(512, 2)
upsample_flow_v4_kernel(const float *__restrict__ flow, const MT *__restrict__ mask, float *__restrict__ out,
                        int H, int W) {
  __shared__ float tile[2][8][UPV_ROW];
  __shared__ float fl[2][3][UP_TX + 4];       // 8*flow at rows y-1..y+1, columns x0-1..x0+32 (index 0 = x0-1)
  const int t = threadIdx.x;
  const int q = t & 7, sx = (t >> 3) & 7, sy = t >> 6;
  const int x0 = blockIdx.x * UP_TX, y = blockIdx.y, b = blockIdx.z;
  const size_t HW = (size_t)H * W;
  const bool in = x0 + 4 * q < W;              // W % 4 == 0: a quad is inside or outside as a whole
  float4 m[9];
  {
    const MT *mp = mask + (size_t)b * 576 * HW + (size_t)y * W + x0 + 4 * q;
#pragma unroll
    for (int k = 0; k < 9; ++k)
      m[k] = in ? ld_mask4(mp + (size_t)(k * 64 + sy * 8 + sx) * HW) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (t < 2 * 3 * (UP_TX + 2)) {
    const int c = t / (3 * (UP_TX + 2)), r = (t / (UP_TX + 2)) % 3, i = t % (UP_TX + 2);
    const int yy = y + r - 1, xx = x0 + i - 1;
    const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;                // F.unfold(padding=1): zeros outside
    fl[c][r][i] = ok ? 8.0f * __ldg(flow + (((size_t)b * 2 + c) * H + yy) * W + xx) : 0.f;
  }
  __syncthreads();
  float o[2][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float w[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) w[k] = i == 0 ? m[k].x : i == 1 ? m[k].y : i == 2 ? m[k].z : m[k].w;
    softmax9(w);
    float o0 = 0.f, o1 = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      o0 = fmaf(w[k], fl[0][k / 3][4 * q + i + k % 3], o0);
      o1 = fmaf(w[k], fl[1][k / 3][4 * q + i + k % 3], o1);
    }
    o[0][i] = o0; o[1][i] = o1;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = 8 * (4 * q + i) + sx;
    tile[0][sy][p + 5 * (p >> 5)] = o[0][i];
    tile[1][sy][p + 5 * (p >> 5)] = o[1][i];
  }
  __syncthreads();
  // 16 output rows of 256 floats: threads 0..255 take channel 0, 256..511 channel 1, column t % 256 of 8 rows each
  const int p = t & 255, c = t >> 8;
  if (x0 + (p >> 3) < W) {
    const size_t W8 = (size_t)8 * W;
#pragma unroll
    for (int r = 0; r < 8; ++r)
      out[(((size_t)b * 2 + c) * 8 * H + 8 * y + r) * W8 + (size_t)8 * x0 + p] = tile[c][r][p + 5 * (p >> 5)];
  }
}

// dflow must be zero on entry.
__global__ void __launch_bounds__(256)
upsample_flow_bwd_kernel(const float *__restrict__ flow, const float *__restrict__ mask,
                         const float *__restrict__ gout, float *__restrict__ dflow, float *__restrict__ dmask, int H,
                         int W) {
  __shared__ float tile[2][8][UP_ROW];
  __shared__ float red[8][18][UP_TX + 1];
  const int lane = threadIdx.x & 31, sx = threadIdx.x >> 5;
  const int x0 = blockIdx.x * UP_TX, y = blockIdx.y, b = blockIdx.z;
  const int x = x0 + lane;
  const bool in = x < W;
  const size_t HW = (size_t)H * W;
  {
    const int p = threadIdx.x;
    const bool ok = x0 + (p >> 3) < W;
    const size_t W8 = (size_t)8 * W;
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int c = r >> 3, sy = r & 7;
      tile[c][sy][p + (p >> 5)] =
          ok ? __ldg(gout + (((size_t)b * 2 + c) * 8 * H + 8 * y + sy) * W8 + (size_t)8 * x0 + p) : 0.f;
    }
  }
  float nb[2][9];
  load_neighbours(flow, b, y, x, H, W, in, nb);
  __syncthreads();
  float dnb[2][9];
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int k = 0; k < 9; ++k) dnb[c][k] = 0.f;
  const size_t moff = (size_t)b * 576 * HW + (size_t)y * W + (in ? x : 0);
#pragma unroll 2
  for (int sy = 0; sy < 8; ++sy) {
    float m[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) m[k] = in ? __ldg(mask + moff + (size_t)(k * 64 + sy * 8 + sx) * HW) : 0.f;
    softmax9(m);
    const int p = 8 * lane + sx;
    const float g0 = tile[0][sy][p + (p >> 5)], g1 = tile[1][sy][p + (p >> 5)];
    float dw[9], dot = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      dw[k] = fmaf(g0, nb[0][k], g1 * nb[1][k]);
      dot = fmaf(m[k], dw[k], dot);
      dnb[0][k] = fmaf(g0, m[k], dnb[0][k]);
      dnb[1][k] = fmaf(g1, m[k], dnb[1][k]);
    }
    if (in) {
#pragma unroll
      for (int k = 0; k < 9; ++k) dmask[moff + (size_t)(k * 64 + sy * 8 + sx) * HW] = m[k] * (dw[k] - dot);
    }
  }
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int k = 0; k < 9; ++k) red[sx][c * 9 + k][lane] = dnb[c][k];
  __syncthreads();
  for (int i = threadIdx.x; i < 18 * UP_TX; i += 256) {
    const int ck = i / UP_TX, l = i % UP_TX;
    const int c = ck / 9, k = ck % 9;
    const int xs = x0 + l, yy = y + k / 3 - 1, xx = xs + k % 3 - 1;
    if (xs < W && yy >= 0 && yy < H && xx >= 0 && xx < W) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w][ck][l];
      atomicAdd(dflow + (((size_t)b * 2 + c) * H + yy) * W + xx, 8.0f * s);
    }
  }
}

// align_corners=True source position of output index i (torch's area_pixel_compute_source_index)
__device__ __forceinline__ void lerp_src(int i, int n_in, float scale, int &i0, int &i1, float &f) {
  const float src = scale * (float)i;
  i0 = min((int)src, n_in - 1);
  i1 = min(i0 + 1, n_in - 1);
  f = src - (float)i0;
}

__global__ void __launch_bounds__(256)
upflow8_kernel(const float *__restrict__ flow, float *__restrict__ out, int C, int H, int W, float sy, float sx) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x, Y = blockIdx.y, bc = blockIdx.z;
  if (X >= 8 * W) return;
  int y0, y1, x0, x1;
  float fy, fx;
  lerp_src(Y, H, sy, y0, y1, fy);
  lerp_src(X, W, sx, x0, x1, fx);
  const float *f = flow + (size_t)bc * H * W;
  const float a = __ldg(f + (size_t)y0 * W + x0), bq = __ldg(f + (size_t)y0 * W + x1);
  const float c = __ldg(f + (size_t)y1 * W + x0), d = __ldg(f + (size_t)y1 * W + x1);
  // torch's upsample_bilinear2d: h0lambda * (w0lambda * a + w1lambda * b) + h1lambda * (...)
  const float top = (1.f - fx) * a + fx * bq, bot = (1.f - fx) * c + fx * d;
  out[((size_t)bc * 8 * H + Y) * 8 * W + X] = 8.0f * ((1.f - fy) * top + fy * bot);
}

// dflow must be zero on entry.
__global__ void __launch_bounds__(256)
upflow8_bwd_kernel(const float *__restrict__ gout, float *__restrict__ dflow, int H, int W, float sy, float sx) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x, Y = blockIdx.y, bc = blockIdx.z;
  if (X >= 8 * W) return;
  int y0, y1, x0, x1;
  float fy, fx;
  lerp_src(Y, H, sy, y0, y1, fy);
  lerp_src(X, W, sx, x0, x1, fx);
  const float g = 8.0f * __ldg(gout + ((size_t)bc * 8 * H + Y) * 8 * W + X);
  float *d = dflow + (size_t)bc * H * W;
  atomicAdd(d + (size_t)y0 * W + x0, g * (1.f - fy) * (1.f - fx));
  atomicAdd(d + (size_t)y0 * W + x1, g * (1.f - fy) * fx);
  atomicAdd(d + (size_t)y1 * W + x0, g * fy * (1.f - fx));
  atomicAdd(d + (size_t)y1 * W + x1, g * fy * fx);
}

}  // namespace

template <typename MT>
static int upsample_flow_t(const float *flow, const MT *mask, int B, int H, int W, float *out, cudaStream_t s) {
  dim3 grid(ceil_div(W, UP_TX), H, B);
  // four pixels per load: 16 bytes of fp32, 8 of fp16; the plane stride H*W keeps that alignment when W % 4 == 0
  if (W % 4 == 0 && ((uintptr_t)mask & (4 * sizeof(MT) - 1)) == 0) {
    upsample_flow_v4_kernel<MT><<<grid, 512, 0, s>>>(flow, mask, out, H, W);
    RC_LAUNCH_CHECK("upsample_flow_v4_kernel");
    return RC_OK;
  }
  upsample_flow_kernel<MT><<<grid, 256, 0, s>>>(flow, mask, out, H, W);
  RC_LAUNCH_CHECK("upsample_flow_kernel");
  return RC_OK;
}

int launch_upsample_flow(const float *flow, const void *mask, int mask_is_half, int B, int H, int W, float *out,
                         cudaStream_t s) {
  return mask_is_half ? upsample_flow_t<__half>(flow, (const __half *)mask, B, H, W, out, s)
                      : upsample_flow_t<float>(flow, (const float *)mask, B, H, W, out, s);
}

int launch_upsample_flow_bwd(const float *flow, const float *mask, const float *gout, int B, int H, int W,
                             float *dflow, float *dmask, cudaStream_t s) {
  RC_CUDA(cudaMemsetAsync(dflow, 0, (size_t)B * 2 * H * W * sizeof(float), s));
  dim3 grid(ceil_div(W, UP_TX), H, B);
  upsample_flow_bwd_kernel<<<grid, 256, 0, s>>>(flow, mask, gout, dflow, dmask, H, W);
  RC_LAUNCH_CHECK("upsample_flow_bwd_kernel");
  return RC_OK;
}

static float align_corners_scale(int n_in, int n_out) { return n_out > 1 ? (float)(n_in - 1) / (float)(n_out - 1) : 0.f; }

int launch_upflow8(const float *flow, int B, int C, int H, int W, float *out, cudaStream_t s) {
  dim3 grid(ceil_div(8 * W, 256), 8 * H, B * C);
  upflow8_kernel<<<grid, 256, 0, s>>>(flow, out, C, H, W, align_corners_scale(H, 8 * H), align_corners_scale(W, 8 * W));
  RC_LAUNCH_CHECK("upflow8_kernel");
  return RC_OK;
}

int launch_upflow8_bwd(const float *gout, int B, int C, int H, int W, float *dflow, cudaStream_t s) {
  RC_CUDA(cudaMemsetAsync(dflow, 0, (size_t)B * C * H * W * sizeof(float), s));
  dim3 grid(ceil_div(8 * W, 256), 8 * H, B * C);
  upflow8_bwd_kernel<<<grid, 256, 0, s>>>(gout, dflow, H, W, align_corners_scale(H, 8 * H),
                                          align_corners_scale(W, 8 * W));
  RC_LAUNCH_CHECK("upflow8_bwd_kernel");
  return RC_OK;
}

}  // namespace rc
