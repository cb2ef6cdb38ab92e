// Windowed bilinear lookup over the correlation pyramid, forward and backward.
//
// Replaces CorrBlock.__call__ (core/corr.py:29-50) + bilinear_sampler /
// F.grid_sample(align_corners=True, zeros padding) (core/utils/utils.py:57-65)
// and, for the backward, what autograd runs for those lines.
//
// Work decomposition: one CTA per strip of 32 consecutive queries of one pair,
// one warp per pyramid level.  All (2r+1)^2 samples of a query at a level share
// one fractional offset (ax, ay) and read a (2r+2)^2 window of integer taps
// (SURVEY 7.0-4), so each warp
//   1. stages the 32 windows in shared memory with vector loads: 4 (or 3)
//      lanes cover one 10-tap window row, 8 rows per load instruction, so a
//      warp-wide load touches ~8 cache lines instead of 32;
//   2. switches to lane == query: separable interpolation (along x, then y)
//      out of shared memory (odd row stride -> no bank conflicts) and
//      128-byte coalesced stores into out[b][channel][q0..q0+31].
// HBM-bound: per query and level it reads (2r+2)^2 volume elements and writes
// (2r+1)^2 floats; no intermediate ever reaches global memory (the reference
// materialises the sampling grid, four per-level outputs, a cat and a
// permute().contiguous() copy).
#include "common.cuh"

namespace rc {
namespace {

template <typename VolT>
struct Vec4;
template <>
struct Vec4<float> {
  using type = float4;
  static __device__ __forceinline__ void unpack(const float4 &v, float *f) {
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
  static __device__ __forceinline__ float4 zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
};
template <>
struct Vec4<__nv_bfloat16> {
  using type = uint2;
  static __device__ __forceinline__ void unpack(const uint2 &v, float *f) {
    f[0] = __uint_as_float(v.x << 16);
    f[1] = __uint_as_float(v.x & 0xffff0000u);
    f[2] = __uint_as_float(v.y << 16);
    f[3] = __uint_as_float(v.y & 0xffff0000u);
  }
  static __device__ __forceinline__ uint2 zero() { return make_uint2(0u, 0u); }
};

template <int R>
struct Win {
  static constexpr int NW = 2 * R + 1;        // samples per axis
  static constexpr int T = 2 * R + 2;         // integer taps per axis
  static constexpr int VPR = (T + 3 + 3) / 4; // aligned 4-vectors that cover any T-tap row
  static constexpr int STRIDE = (T * T) | 1;  // words per query in smem (odd)
  static constexpr int SLOTS = T * VPR;       // vector loads per query
  static constexpr int U = (SLOTS % 8 == 0) ? 8 : ((SLOTS % 6 == 0) ? 6 : 2);   // loads in flight per lane
};

struct QueryWin {
  int x0, y0;       // first tap of the window (may be far outside)
  float ax, ay;     // fractional offsets
};

// Window origin of a query at a level.  x / 2^l is an exact scaling, identical
// to the reference's division (core/corr.py:41).  Non-finite or absurd
// coordinates are treated as "window entirely outside" -> zeros.
template <int R>
__device__ __forceinline__ QueryWin make_window(float x, float y, int level, bool active) {
  const float inv = __int_as_float((127 - level) << 23);   // 2^-level
  const float cx = x * inv, cy = y * inv;
  const float fx = floorf(cx), fy = floorf(cy);
  QueryWin w;
  const bool sane = active && fabsf(cx) < 1.0e7f && fabsf(cy) < 1.0e7f;
  w.x0 = sane ? (int)fx - R : -(1 << 24);
  w.y0 = sane ? (int)fy - R : -(1 << 24);
  w.ax = sane ? cx - fx : 0.f;
  w.ay = sane ? cy - fy : 0.f;
  return w;
}

// Stage the T x T windows of the warp's 32 queries into `taps`
// (taps[qq*STRIDE + j*T + i] = level[y0+j][x0+i], zero outside the map).
template <typename VolT, int R>
__device__ __forceinline__ void stage_windows(const VolT *__restrict__ lvl, float *taps,
                                              const QueryWin &w, long long slab0, long long slab,
                                              int pitch, int Hl, int Wl, int lane) {
  using W_ = Win<R>;
  using V = Vec4<VolT>;
  constexpr int U = W_::U;
  static_assert(W_::SLOTS % U == 0, "unroll must divide slots");
#pragma unroll 1
  for (int it0 = 0; it0 < W_::SLOTS; it0 += U) {
    typename V::type v[U];
    int dst[U], xb[U], qx0[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int s = (it0 + u) * 32 + lane;
      const int qq = s / W_::SLOTS;
      const int rem = s - qq * W_::SLOTS;
      const int j = rem / W_::VPR;
      const int k = rem - j * W_::VPR;
      const int x0 = __shfl_sync(0xffffffffu, w.x0, qq);
      const int y0 = __shfl_sync(0xffffffffu, w.y0, qq);
      const int y = y0 + j;
      const int xbase = ((x0 >> 2) + k) * 4;
      const bool ok = y >= 0 && y < Hl && xbase >= 0 && xbase < pitch;
      v[u] = V::zero();
      if (ok) {
        const VolT *p = lvl + (slab0 + qq) * slab + (size_t)y * pitch + xbase;
        v[u] = __ldg(reinterpret_cast<const typename V::type *>(p));
      }
      dst[u] = qq * W_::STRIDE + j * W_::T;
      xb[u] = xbase;
      qx0[u] = x0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float f[4];
      V::unpack(v[u], f);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int x = xb[u] + e;
        const int i = x - qx0[u];
        if (i >= 0 && i < W_::T) taps[dst[u] + i] = (x < Wl) ? f[e] : 0.f;
      }
    }
  }
}

// grid: B * ceil(N/32) strips; block: 32 * num_levels threads; dyn smem: L*32*STRIDE floats
template <typename VolT, int R>
__global__ void __launch_bounds__(32 * RC_MAX_LEVELS)
lookup_fwd_kernel(const VolT *__restrict__ pyr, const float *__restrict__ coords,
                  float *__restrict__ out, Levels lv, int N, int strips_per_pair) {
  using W_ = Win<R>;
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31;
  const int l = threadIdx.x >> 5;
  float *taps = smem + (size_t)l * 32 * W_::STRIDE;

  const int b = blockIdx.x / strips_per_pair;
  const int q0 = (blockIdx.x - b * strips_per_pair) * 32;
  const int q = q0 + lane;
  const bool active = q < N;
  float x = 0.f, y = 0.f;
  if (active) {
    x = __ldg(coords + ((size_t)b * 2 + 0) * N + q);
    y = __ldg(coords + ((size_t)b * 2 + 1) * N + q);
  }
  const QueryWin w = make_window<R>(x, y, l, active);

  stage_windows<VolT, R>(pyr + lv.offset[l], taps, w, (long long)b * N + q0, lv.slab[l],
                         lv.pitch[l], lv.H[l], lv.W[l], lane);
  __syncwarp();

  if (!active) return;
  const float *t = taps + lane * W_::STRIDE;
  const int C = lv.num_levels * W_::NW * W_::NW;
  float *o = out + ((size_t)b * C + (size_t)l * W_::NW * W_::NW) * N + q;
  float hprev[W_::NW];
#pragma unroll
  for (int j = 0; j < W_::T; ++j) {
    float v[W_::T], h[W_::NW];
#pragma unroll
    for (int i = 0; i < W_::T; ++i) v[i] = t[j * W_::T + i];
#pragma unroll
    for (int i = 0; i < W_::NW; ++i) h[i] = fmaf(w.ax, v[i + 1] - v[i], v[i]);
    if (j > 0) {
#pragma unroll
      for (int i = 0; i < W_::NW; ++i)
        o[(size_t)(i * W_::NW + (j - 1)) * N] = fmaf(w.ay, h[i] - hprev[i], hprev[i]);
    }
#pragma unroll
    for (int i = 0; i < W_::NW; ++i) hprev[i] = h[i];
  }
}

// Backward.  Same decomposition.  Per warp: stage the forward windows (needed
// for d/dcoords), then lane == query folds the (2r+1)^2 output gradients into
// the (2r+2)^2 tap gradients IN PLACE in shared memory (row j of the window is
// dead once output rows j-1 and j have used it), then the staging pattern is
// run in reverse to add the tap gradients into the fp32 gradient pyramid.
// Every query owns its slabs, so the read-modify-write needs no atomics.
template <typename VolT, int R>
__global__ void __launch_bounds__(32 * RC_MAX_LEVELS)
lookup_bwd_kernel(const VolT *__restrict__ pyr, const float *__restrict__ coords,
                  const float *__restrict__ grad_out, float *__restrict__ gpyr,
                  float *__restrict__ gcoords, Levels lv, Levels glv, int N, int strips_per_pair) {
  using W_ = Win<R>;
  extern __shared__ float smem[];
  __shared__ float dc[RC_MAX_LEVELS][2][32];
  const int lane = threadIdx.x & 31;
  const int l = threadIdx.x >> 5;
  float *taps = smem + (size_t)l * 32 * W_::STRIDE;

  const int b = blockIdx.x / strips_per_pair;
  const int q0 = (blockIdx.x - b * strips_per_pair) * 32;
  const int q = q0 + lane;
  const bool active = q < N;
  float x = 0.f, y = 0.f;
  if (active) {
    x = __ldg(coords + ((size_t)b * 2 + 0) * N + q);
    y = __ldg(coords + ((size_t)b * 2 + 1) * N + q);
  }
  const QueryWin w = make_window<R>(x, y, l, active);
  const bool want_dc = gcoords != nullptr;

  if (want_dc) {
    stage_windows<VolT, R>(pyr + lv.offset[l], taps, w, (long long)b * N + q0, lv.slab[l],
                           lv.pitch[l], lv.H[l], lv.W[l], lane);
    __syncwarp();
  }

  float dx = 0.f, dy = 0.f;
  {
    float *t = taps + lane * W_::STRIDE;
    const int C = lv.num_levels * W_::NW * W_::NW;
    const float *g = grad_out + ((size_t)b * C + (size_t)l * W_::NW * W_::NW) * N + q;
    float hg_prev[W_::T];   // x-folded gradient of output row j-1
    float cur[W_::T];       // forward taps, row j
#pragma unroll
    for (int i = 0; i < W_::T; ++i) {
      hg_prev[i] = 0.f;
      cur[i] = want_dc ? t[i] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < W_::NW; ++j) {
      float gj[W_::NW], nxt[W_::T], hg[W_::T];
#pragma unroll
      for (int i = 0; i < W_::NW; ++i) gj[i] = active ? __ldg(g + (size_t)(i * W_::NW + j) * N) : 0.f;
#pragma unroll
      for (int i = 0; i < W_::T; ++i) nxt[i] = want_dc ? t[(j + 1) * W_::T + i] : 0.f;
      // d/dcoords
#pragma unroll
      for (int i = 0; i < W_::NW; ++i) {
        const float d0 = cur[i + 1] - cur[i], d1 = nxt[i + 1] - nxt[i];
        dx = fmaf(gj[i], fmaf(w.ay, d1 - d0, d0), dx);
        const float e0 = nxt[i] - cur[i], e1 = nxt[i + 1] - cur[i + 1];
        dy = fmaf(gj[i], fmaf(w.ax, e1 - e0, e0), dy);
      }
      // fold along x: hg[ii] = (1-ax) g[ii] + ax g[ii-1]
#pragma unroll
      for (int i = 0; i < W_::T; ++i) {
        const float a = (i < W_::NW) ? gj[i] : 0.f;
        const float c = (i > 0) ? gj[i - 1] : 0.f;
        hg[i] = fmaf(w.ax, c - a, a);
      }
      // tap-gradient row j is final: (1-ay) hg_j + ay hg_{j-1}
#pragma unroll
      for (int i = 0; i < W_::T; ++i) t[j * W_::T + i] = fmaf(w.ay, hg_prev[i] - hg[i], hg[i]);
#pragma unroll
      for (int i = 0; i < W_::T; ++i) {
        hg_prev[i] = hg[i];
        cur[i] = nxt[i];
      }
    }
#pragma unroll
    for (int i = 0; i < W_::T; ++i) t[(W_::T - 1) * W_::T + i] = w.ay * hg_prev[i];
  }
  __syncwarp();

  if (gpyr != nullptr) {
    float *glvl = gpyr + glv.offset[l];
    const long long slab0 = (long long)b * N + q0;
    const int Hl = glv.H[l], Wl = glv.W[l], pitch = glv.pitch[l];
    const long long slab = glv.slab[l];
#pragma unroll 4
    for (int it = 0; it < W_::SLOTS; ++it) {
      const int s = it * 32 + lane;
      const int qq = s / W_::SLOTS;
      const int rem = s - qq * W_::SLOTS;
      const int j = rem / W_::VPR;
      const int k = rem - j * W_::VPR;
      const int x0 = __shfl_sync(0xffffffffu, w.x0, qq);
      const int y0 = __shfl_sync(0xffffffffu, w.y0, qq);
      const int yy = y0 + j;
      const int xbase = ((x0 >> 2) + k) * 4;
      // touch the vector only if it intersects the window and the map
      const bool ok = yy >= 0 && yy < Hl && xbase >= 0 && xbase < pitch && xbase < Wl &&
                      xbase + 3 >= x0 && xbase < x0 + W_::T;
      if (ok) {
        float4 *p = reinterpret_cast<float4 *>(glvl + (slab0 + qq) * slab + (size_t)yy * pitch + xbase);
        float4 v = *p;
        float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int xx = xbase + e;
          const int i = xx - x0;
          if (i >= 0 && i < W_::T && xx < Wl) f[e] += taps[qq * W_::STRIDE + j * W_::T + i];
        }
        *p = make_float4(f[0], f[1], f[2], f[3]);
      }
    }
  }

  if (want_dc) {
    const float inv = __int_as_float((127 - l) << 23);
    dc[l][0][lane] = dx * inv;
    dc[l][1][lane] = dy * inv;
    __syncthreads();
    if (l == 0 && active) {
      float sx = 0.f, sy = 0.f;
      for (int k = 0; k < lv.num_levels; ++k) {
        sx += dc[k][0][lane];
        sy += dc[k][1][lane];
      }
      gcoords[((size_t)b * 2 + 0) * N + q] = sx;
      gcoords[((size_t)b * 2 + 1) * N + q] = sy;
    }
  }
}

template <typename VolT, int R>
int lookup_fwd_t(const VolT *pyr, const float *coords, const rc_layout &lay, float *out,
                 cudaStream_t s) {
  using W_ = Win<R>;
  const int N = lay.N;
  const int strips = ceil_div(N, 32);
  const size_t smem = (size_t)lay.num_levels * 32 * W_::STRIDE * sizeof(float);
  auto kern = lookup_fwd_kernel<VolT, R>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<lay.B * strips, 32 * lay.num_levels, smem, s>>>(pyr, coords, out, to_levels(lay), N, strips);
  RC_LAUNCH_CHECK("lookup_fwd_kernel");
  return RC_OK;
}

template <typename VolT, int R>
int lookup_bwd_t(const VolT *pyr, const float *coords, const float *grad_out, const rc_layout &lay,
                 const rc_layout &glay, float *gpyr, float *gcoords, cudaStream_t s) {
  using W_ = Win<R>;
  const int N = lay.N;
  const int strips = ceil_div(N, 32);
  const size_t smem = (size_t)lay.num_levels * 32 * W_::STRIDE * sizeof(float);
  auto kern = lookup_bwd_kernel<VolT, R>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<lay.B * strips, 32 * lay.num_levels, smem, s>>>(pyr, coords, grad_out, gpyr, gcoords,
                                                         to_levels(lay), to_levels(glay), N, strips);
  RC_LAUNCH_CHECK("lookup_bwd_kernel");
  return RC_OK;
}

}  // namespace

#define RC_DISPATCH_R(FN, VT, ...)                                  \
  switch (radius) {                                                 \
    case 1: return FN<VT, 1>(__VA_ARGS__);                          \
    case 2: return FN<VT, 2>(__VA_ARGS__);                          \
    case 3: return FN<VT, 3>(__VA_ARGS__);                          \
    case 4: return FN<VT, 4>(__VA_ARGS__);                          \
    default: set_error("radius %d unsupported (1..4)", radius); return RC_EINVAL; \
  }

int launch_lookup_fwd(const void *pyramid, int vol_dtype, const float *coords, const rc_layout &lay,
                      int H, int W, int radius, float *out, cudaStream_t s) {
  (void)H; (void)W;
  if (vol_dtype == RC_VOL_BF16) {
    RC_DISPATCH_R(lookup_fwd_t, __nv_bfloat16, (const __nv_bfloat16 *)pyramid, coords, lay, out, s)
  } else {
    RC_DISPATCH_R(lookup_fwd_t, float, (const float *)pyramid, coords, lay, out, s)
  }
}

int launch_lookup_bwd(const void *pyramid, int vol_dtype, const float *coords, const float *grad_out,
                      const rc_layout &lay, const rc_layout &glay, int H, int W, int radius,
                      float *grad_pyramid, float *grad_coords, cudaStream_t s) {
  (void)H; (void)W;
  if (vol_dtype == RC_VOL_BF16) {
    RC_DISPATCH_R(lookup_bwd_t, __nv_bfloat16, (const __nv_bfloat16 *)pyramid, coords, grad_out, lay,
                  glay, grad_pyramid, grad_coords, s)
  } else {
    RC_DISPATCH_R(lookup_bwd_t, float, (const float *)pyramid, coords, grad_out, lay, glay,
                  grad_pyramid, grad_coords, s)
  }
}

}  // namespace rc
