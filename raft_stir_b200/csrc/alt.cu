// On-the-fly correlation ("alternate" path): feature dot products fused with
// the bilinear taps; the N x N volume is never materialised.
//
// Replaces alt_cuda_corr/correlation_kernel.cu (forward :18-119, backward
// :122-256) and the per-call glue of AlternateCorrBlock (core/corr.py:63-91).
//
// The reference kernel runs one warp per 4x8 query block, refills two
// 32-channel shared tiles per tap and splats every partial dot product into
// global memory with four read-modify-writes.  Here a CTA owns a strip of 32
// consecutive queries (so the result can be staged and stored 128 bytes at a
// time), a warp walks the (2r+2)^2 integer taps of one query at a time with
// the query's feature vector held in registers, lanes split the channels
// (512-byte coalesced reads of a tap's NHWC feature row), 32 taps at a time
// are reduced with a transposing butterfly (31 shuffles for 32 dot products),
// and the bilinear interpolation runs once on the finished window.
// fp32 FMA throughout, like the reference.
#include "common.cuh"

namespace rc {
namespace {

constexpr int ALT_WARPS = 8;
constexpr int ALT_Q = 32;        // queries per CTA
constexpr int ALT_MAXC = 512;    // channels (4 float4 per lane)
constexpr int ALT_KV = ALT_MAXC / 128;

struct AltLevels {
  int count;                       // entries along the "mid" axis (levels, or Nc)
  int H2[RC_MAX_LEVELS], W2[RC_MAX_LEVELS];
  long long off[RC_MAX_LEVELS];    // element offset of fmap2 level inside fmap2 base
  float cscale[RC_MAX_LEVELS];     // multiply coords by this (2^-l)
};

// sum over lanes of v[t], delivered to lane t (t = 0..31)
__device__ __forceinline__ float transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? v[i] : v[i + s];
      const float keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

// PLANAR: coords [B,2,N1] (x plane, y plane), shared by all mid entries and scaled per level
// else  : coords [B,count,N1,2] interleaved, one set per mid entry
template <int R, bool PLANAR>
__global__ void __launch_bounds__(ALT_WARPS * 32)
alt_fwd_kernel(const float *__restrict__ fmap1, const float *__restrict__ fmap2,
               const float *__restrict__ coords, float *__restrict__ out, AltLevels lv, int N1, int C,
               float out_scale) {
  constexpr int NW = 2 * R + 1, T = 2 * R + 2, TT = T * T, NG = (TT + 31) / 32;
  __shared__ float dots[ALT_WARPS][NG * 32];
  __shared__ float stage[NW * NW][ALT_Q + 1];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.z, mid = blockIdx.y;
  const int q0 = blockIdx.x * ALT_Q;
  const int li = PLANAR ? mid : 0;       // one-level calls (any number of coordinate sets) share entry 0
  const int H2 = lv.H2[li], W2 = lv.W2[li];
  const float *f2 = fmap2 + lv.off[li] + (size_t)b * H2 * W2 * C;

  for (int qi = warp; qi < ALT_Q; qi += ALT_WARPS) {
    const int q = q0 + qi;
    if (q >= N1) break;   // warp-uniform
    float cx, cy;
    if (PLANAR) {
      cx = __ldg(coords + ((size_t)b * 2 + 0) * N1 + q) * lv.cscale[mid];
      cy = __ldg(coords + ((size_t)b * 2 + 1) * N1 + q) * lv.cscale[mid];
    } else {
      const float2 c2 = __ldg(reinterpret_cast<const float2 *>(coords) + ((size_t)b * lv.count + mid) * N1 + q);
      cx = c2.x; cy = c2.y;
    }
    const bool sane = fabsf(cx) < 1.0e7f && fabsf(cy) < 1.0e7f;
    const float fx = floorf(cx), fy = floorf(cy);
    const int x0 = sane ? (int)fx - R : -(1 << 24);
    const int y0 = sane ? (int)fy - R : -(1 << 24);
    // NaN / infinite coordinates: NaN in every sample of the query (what the reference kernel's dx = x - floor(x) gives)
    const float off = ((cx - cx == 0.f) && (cy - cy == 0.f)) ? 0.f : __int_as_float(0x7fc00000);
    const float ax = sane ? cx - fx : off, ay = sane ? cy - fy : off;

    // this query's feature vector: lane holds channels lane*4 + 128*k .. +3
    float4 a[ALT_KV];
    const float *f1 = fmap1 + ((size_t)b * N1 + q) * C;
#pragma unroll
    for (int k = 0; k < ALT_KV; ++k) {
      const int c = lane * 4 + 128 * k;
      a[k] = (c < C) ? __ldg(reinterpret_cast<const float4 *>(f1 + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }

#pragma unroll 1
    for (int g = 0; g < NG; ++g) {
      float part[32];
#pragma unroll
      for (int tt = 0; tt < 32; ++tt) {
        const int t = g * 32 + tt;
        const int j = t / T, i = t - j * T;
        const int yy = y0 + j, xx = x0 + i;
        float p = 0.f;
        if (t < TT && yy >= 0 && yy < H2 && xx >= 0 && xx < W2) {   // warp-uniform
          const float *row = f2 + ((size_t)yy * W2 + xx) * C;
#pragma unroll
          for (int k = 0; k < ALT_KV; ++k) {
            const int c = lane * 4 + 128 * k;
            if (c < C) {
              const float4 v = __ldg(reinterpret_cast<const float4 *>(row + c));
              p = fmaf(a[k].x, v.x, p);
              p = fmaf(a[k].y, v.y, p);
              p = fmaf(a[k].z, v.z, p);
              p = fmaf(a[k].w, v.w, p);
            }
          }
        }
        part[tt] = p;
      }
      dots[warp][g * 32 + lane] = transpose_reduce32(part, lane);
    }
    __syncwarp();
    // bilinear on the finished window; output index = iy + NW*ix (correlation_kernel.cu:92-95)
    for (int o = lane; o < NW * NW; o += 32) {
      const int ix = o / NW, iy = o - ix * NW;
      const float v00 = dots[warp][iy * T + ix], v01 = dots[warp][iy * T + ix + 1];
      const float v10 = dots[warp][(iy + 1) * T + ix], v11 = dots[warp][(iy + 1) * T + ix + 1];
      const float h0 = fmaf(ax, v01 - v00, v00), h1 = fmaf(ax, v11 - v10, v10);
      stage[o][qi] = fmaf(ay, h1 - h0, h0) * out_scale;
    }
    __syncwarp();
  }
  __syncthreads();
  const int nq = min(ALT_Q, N1 - q0);
  float *o = out + (((size_t)b * lv.count + mid) * NW * NW) * N1 + q0;
  for (int idx = threadIdx.x; idx < NW * NW * ALT_Q; idx += ALT_WARPS * 32) {
    const int ch = idx >> 5, qi = idx & 31;
    if (qi < nq) o[(size_t)ch * N1 + qi] = stage[ch][qi];
  }
}

// Backward of one level (contract of alt_cuda_corr.backward).  A warp owns a
// query: it folds the (2r+1)^2 output gradients into (2r+2)^2 tap gradients,
// then for every in-bounds tap accumulates fmap1_grad in registers and adds
// g * f1 into fmap2_grad with vector atomics (several queries hit one tap).
template <int R>
__global__ void __launch_bounds__(ALT_WARPS * 32)
alt_bwd_kernel(const float *__restrict__ fmap1, const float *__restrict__ fmap2,
               const float *__restrict__ coords, const float *__restrict__ corr_grad,
               float *__restrict__ g1, float *__restrict__ g2, int N1, int H2, int W2, int C, int Nc) {
  constexpr int NW = 2 * R + 1, T = 2 * R + 2, TT = T * T;
  __shared__ float gout[ALT_WARPS][NW * NW];
  __shared__ float gtap[ALT_WARPS][TT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int q = blockIdx.x * ALT_WARPS + warp;
  if (q >= N1) return;
  const float *f2 = fmap2 + (size_t)b * H2 * W2 * C;
  float *g2b = g2 + (size_t)b * H2 * W2 * C;

  float4 a[ALT_KV], acc[ALT_KV];
  const float *f1 = fmap1 + ((size_t)b * N1 + q) * C;
#pragma unroll
  for (int k = 0; k < ALT_KV; ++k) {
    const int c = lane * 4 + 128 * k;
    a[k] = (c < C) ? __ldg(reinterpret_cast<const float4 *>(f1 + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
    acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  }

  for (int n = 0; n < Nc; ++n) {
    const float2 c2 = __ldg(reinterpret_cast<const float2 *>(coords) + ((size_t)b * Nc + n) * N1 + q);
    const float cx = c2.x, cy = c2.y;
    const bool sane = fabsf(cx) < 1.0e7f && fabsf(cy) < 1.0e7f;
    const float fx = floorf(cx), fy = floorf(cy);
    const int x0 = sane ? (int)fx - R : -(1 << 24);
    const int y0 = sane ? (int)fy - R : -(1 << 24);
    // NaN / infinite coordinates: NaN in every sample of the query (what the reference kernel's dx = x - floor(x) gives)
    const float off = ((cx - cx == 0.f) && (cy - cy == 0.f)) ? 0.f : __int_as_float(0x7fc00000);
    const float ax = sane ? cx - fx : off, ay = sane ? cy - fy : off;

    const float *g = corr_grad + (((size_t)b * Nc + n) * NW * NW) * N1 + q;
    for (int o = lane; o < NW * NW; o += 32) gout[warp][o] = __ldg(g + (size_t)o * N1);
    __syncwarp();
    for (int t = lane; t < TT; t += 32) {
      const int j = t / T, i = t - j * T;   // tap (row j <-> y, col i <-> x)
      // outputs (ix, iy) with ix in {i-1, i}, iy in {j-1, j}; index iy + NW*ix
      auto G = [&](int ix, int iy) -> float {
        return (ix >= 0 && ix < NW && iy >= 0 && iy < NW) ? gout[warp][iy + NW * ix] : 0.f;
      };
      gtap[warp][t] = (1.f - ay) * ((1.f - ax) * G(i, j) + ax * G(i - 1, j)) +
                      ay * ((1.f - ax) * G(i, j - 1) + ax * G(i - 1, j - 1));
    }
    __syncwarp();
#pragma unroll 1
    for (int t = 0; t < TT; ++t) {
      const int j = t / T, i = t - j * T;
      const int yy = y0 + j, xx = x0 + i;
      if (yy < 0 || yy >= H2 || xx < 0 || xx >= W2) continue;   // warp-uniform
      const float gt = gtap[warp][t];
      const size_t rowoff = ((size_t)yy * W2 + xx) * C;
#pragma unroll
      for (int k = 0; k < ALT_KV; ++k) {
        const int c = lane * 4 + 128 * k;
        if (c < C) {
          const float4 v = __ldg(reinterpret_cast<const float4 *>(f2 + rowoff + c));
          acc[k].x = fmaf(gt, v.x, acc[k].x);
          acc[k].y = fmaf(gt, v.y, acc[k].y);
          acc[k].z = fmaf(gt, v.z, acc[k].z);
          acc[k].w = fmaf(gt, v.w, acc[k].w);
          atomicAdd(reinterpret_cast<float4 *>(g2b + rowoff + c),
                    make_float4(gt * a[k].x, gt * a[k].y, gt * a[k].z, gt * a[k].w));
        }
      }
    }
    __syncwarp();
  }
  float *o = g1 + ((size_t)b * N1 + q) * C;
#pragma unroll
  for (int k = 0; k < ALT_KV; ++k) {
    const int c = lane * 4 + 128 * k;
    if (c < C) *reinterpret_cast<float4 *>(o + c) = acc[k];
  }
}

// NCHW -> NHWC transpose of both maps (level 0 of the feature pyramid)
__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const float *__restrict__ a, const float *__restrict__ bsrc, float *__restrict__ oa,
                    float *__restrict__ ob, int D, int N) {
  __shared__ float tile[32][33];
  const bool second = blockIdx.z & 1;
  const int b = blockIdx.z >> 1;
  const float *in = (second ? bsrc : a) + (size_t)b * D * N;
  float *out = (second ? ob : oa) + (size_t)b * D * N;
  const int n0 = blockIdx.x * 32, d0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int d = d0 + ty + 8 * k, n = n0 + tx;
    tile[ty + 8 * k][tx] = (d < D && n < N) ? __ldg(in + (size_t)d * N + n) : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int n = n0 + ty + 8 * k, d = d0 + tx;
    if (n < N && d < D) out[(size_t)n * D + d] = tile[tx][ty + 8 * k];
  }
}

// NHWC 2x2 mean, floor-truncating (F.avg_pool2d at core/corr.py:70-71)
__global__ void __launch_bounds__(256)
pool_nhwc_kernel(const float *__restrict__ src, float *__restrict__ dst, int B, int Hs, int Ws, int D) {
  const int Hd = Hs >> 1, Wd = Ws >> 1;
  const long long total = (long long)B * Hd * Wd * (D / 4);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % (D / 4));
    long long r = i / (D / 4);
    const int x = (int)(r % Wd); r /= Wd;
    const int y = (int)(r % Hd);
    const int b = (int)(r / Hd);
    const float4 *p = reinterpret_cast<const float4 *>(src + (((size_t)b * Hs + 2 * y) * Ws + 2 * x) * D) + c4;
    const float4 v0 = p[0], v1 = p[D / 4], v2 = p[(size_t)Ws * (D / 4)], v3 = p[(size_t)Ws * (D / 4) + D / 4];
    float4 o;
    o.x = 0.25f * (v0.x + v1.x + v2.x + v3.x);
    o.y = 0.25f * (v0.y + v1.y + v2.y + v3.y);
    o.z = 0.25f * (v0.z + v1.z + v2.z + v3.z);
    o.w = 0.25f * (v0.w + v1.w + v2.w + v3.w);
    reinterpret_cast<float4 *>(dst + (((size_t)b * Hd + y) * Wd + x) * D)[c4] = o;
  }
}

template <int R, bool PLANAR>
int alt_fwd_t(const float *f1, const float *f2, const float *coords, float *out, const AltLevels &lv,
              int B, int N1, int C, float scale, cudaStream_t s) {
  dim3 grid(ceil_div(N1, ALT_Q), lv.count, B);
  alt_fwd_kernel<R, PLANAR><<<grid, ALT_WARPS * 32, 0, s>>>(f1, f2, coords, out, lv, N1, C, scale);
  RC_LAUNCH_CHECK("alt_fwd_kernel");
  return RC_OK;
}

template <int R>
int alt_bwd_t(const float *f1, const float *f2, const float *coords, const float *cg, float *g1,
              float *g2, int B, int N1, int H2, int W2, int C, int Nc, cudaStream_t s) {
  dim3 grid(ceil_div(N1, ALT_WARPS), B);
  alt_bwd_kernel<R><<<grid, ALT_WARPS * 32, 0, s>>>(f1, f2, coords, cg, g1, g2, N1, H2, W2, C, Nc);
  RC_LAUNCH_CHECK("alt_bwd_kernel");
  return RC_OK;
}

}  // namespace

#define RC_ALT_SWITCH(EXPR)                                          \
  switch (radius) {                                                  \
    case 1: { constexpr int R_ = 1; return EXPR; }                   \
    case 2: { constexpr int R_ = 2; return EXPR; }                   \
    case 3: { constexpr int R_ = 3; return EXPR; }                   \
    case 4: { constexpr int R_ = 4; return EXPR; }                   \
    default: set_error("radius %d unsupported (1..4)", radius); return RC_EINVAL; \
  }

int launch_alt_fwd(const float *fmap1, const float *fmap2, const float *coords, int B, int H1, int W1,
                   int H2, int W2, int C, int Nc, int radius, float *corr, cudaStream_t s) {
  AltLevels lv{};
  lv.count = Nc;                          // coordinate sets: grid.y, no per-set table (all sets sample the same fmap2)
  lv.H2[0] = H2; lv.W2[0] = W2; lv.off[0] = 0; lv.cscale[0] = 1.f;
  RC_ALT_SWITCH((alt_fwd_t<R_, false>(fmap1, fmap2, coords, corr, lv, B, H1 * W1, C, 1.0f, s)))
}

int launch_alt_lookup_fwd(const float *fmap1_nhwc, const float *fmap2_pyr, const float *coords, int B,
                          int D, int H, int W, int num_levels, int radius, const int64_t *off,
                          float *out, cudaStream_t s) {
  AltLevels lv{};
  lv.count = num_levels;
  for (int l = 0; l < num_levels; ++l) {
    lv.H2[l] = H >> l; lv.W2[l] = W >> l; lv.off[l] = off[l];
    lv.cscale[l] = 1.0f / (float)(1 << l);
  }
  const float scale = 1.0f / sqrtf((float)D);
  RC_ALT_SWITCH((alt_fwd_t<R_, true>(fmap1_nhwc, fmap2_pyr, coords, out, lv, B, H * W, D, scale, s)))
}

int launch_alt_bwd(const float *fmap1, const float *fmap2, const float *coords, const float *corr_grad,
                   int B, int H1, int W1, int H2, int W2, int C, int Nc, int radius, float *g1,
                   float *g2, cudaStream_t s) {
  RC_ALT_SWITCH((alt_bwd_t<R_>(fmap1, fmap2, coords, corr_grad, g1, g2, B, H1 * W1, H2, W2, C, Nc, s)))
}

int launch_alt_pool(const float *fmap1, const float *fmap2, int B, int D, int H, int W, int num_levels,
                    const int64_t *off, float *f1_nhwc, float *f2_pyr, cudaStream_t s) {
  const int N = H * W;
  dim3 grid(ceil_div(N, 32), ceil_div(D, 32), 2 * B);
  nchw_to_nhwc_kernel<<<grid, 256, 0, s>>>(fmap1, fmap2, f1_nhwc, f2_pyr + off[0], D, N);
  RC_LAUNCH_CHECK("nchw_to_nhwc_kernel");
  for (int l = 1; l < num_levels; ++l) {
    const int Hs = H >> (l - 1), Ws = W >> (l - 1);
    const long long total = (long long)B * (Hs >> 1) * (Ws >> 1) * (D / 4);
    if (total == 0) continue;
    const int blocks = (int)((total + 255) / 256 < 148LL * 16 ? (total + 255) / 256 : 148LL * 16);
    pool_nhwc_kernel<<<blocks, 256, 0, s>>>(f2_pyr + off[l - 1], f2_pyr + off[l], B, Hs, Ws, D);
    RC_LAUNCH_CHECK("pool_nhwc_kernel");
  }
  return RC_OK;
}

}  // namespace rc
