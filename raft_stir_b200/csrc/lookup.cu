// Windowed bilinear lookup over the correlation pyramid, forward and backward.
//
// Replaces CorrBlock.__call__ (core/corr.py:29-50) + bilinear_sampler /
// F.grid_sample(align_corners=True, zeros padding) (core/utils/utils.py:57-65)
// and, for the backward, what autograd runs for those lines.
//
// Work decomposition: one CTA per strip of 32 consecutive queries of one pair,
// one warp per pyramid level.  All (2r+1)^2 samples of a query at a level share
// one fractional offset (ax, ay) and read a (2r+2)^2 window of integer taps
// (SURVEY 7.0-4).  The pyramid is stored in 64-byte tiles of 4 rows x 16 bytes
// (rc_layout), the unit B200's L2 fetches from HBM, so a window is ~7 tiles.
//   1. STAGE: the warp copies the 16-byte tile rows ("units") that cover its 32
//      windows into shared memory, unshifted, with cp.async: unit (row j, tile
//      column c) of a window is one 16-byte copy of one lane; consecutive lanes
//      take the units of one window, so the 4 rows of a tile (and its
//      x-neighbour) sit in the same instruction and coalesce.  All copies of a
//      warp are in flight at once, none holds a register.  Out-of-map units are
//      zero-filled (= grid_sample's zeros padding); padding columns inside edge
//      tiles are zero in memory.
//   2. INTERPOLATE: lane == query.  The lane reads the units of a window row
//      with 16-byte shared loads, shifts them by its own sub-tile offset
//      (x0 mod tile width) in registers, runs the separable interpolation
//      (along x, then y) and stores 128 bytes per warp and channel straight into
//      out[b][channel][q0..q0+31].
// Shared memory holds one row of units per query; the unit position is XORed
// with bits of the query so that the 16-byte reads of 8 consecutive queries hit
// 8 different bank groups.
// HBM-bound: per query and level it reads (2r+2)^2 volume elements and writes
// (2r+1)^2 floats; no intermediate reaches global memory (the reference
// materialises the sampling grid, four per-level outputs, a cat and a
// permute().contiguous() copy).
#include "common.cuh"
#include <cuda_fp16.h>

#include <atomic>

namespace rc {
namespace {

// Geometry of one (volume type, radius) combination.
template <typename VolT, int R>
struct Win {
  static constexpr int NW = 2 * R + 1;                 // samples per axis
  static constexpr int T = 2 * R + 2;                  // integer taps per axis
  static constexpr int EPV = Tile<VolT>::W;            // elements per 16-byte unit
  static constexpr int LOG_EPV = Tile<VolT>::LOG_W;
  static constexpr int SPAN = EPV - 1 + T;             // widest reach from the first unit's start
  // When the widest window overshoots NCOL units by exactly one element, that
  // element is fetched by the query's own lane instead of a whole extra unit column.
  static constexpr int EXTRA = (SPAN % EPV == 1) ? 1 : 0;
  static constexpr int NCOL = EXTRA ? SPAN / EPV : (SPAN + EPV - 1) / EPV;
  static constexpr int SL = T * NCOL;                  // units per window
  static constexpr int ROWW = NCOL * 4;                // 32-bit words per staged window row
  static constexpr int WPQ = SL * 4;                   // words per window
  static constexpr int K = SL | 1;                     // odd bank-rotation multiplier
  static constexpr int ITER = SL / 2;                  // copy rounds per group of 16 queries
  // staged forward window: units per query row, padded to 4 (mod 8) so that rows of
  // consecutive queries alternate between the two halves of the 8 bank groups
  static constexpr int SU = SL + ((4 - SL % 8) + 8) % 8;
  static constexpr int WIN_BYTES = 32 * SU * 16;       // per level (32 queries)
  // words of a row a lane may need: NEED, read from W[0 .. NEED+2]
  static constexpr int NEED = sizeof(VolT) == 2 ? T / 2 + 1 : T;
  static_assert(SL % 2 == 0, "T is even");
  static_assert(NEED + 3 <= ROWW + 1, "window row does not fit the staged units");
};

struct QueryWin {
  int x0, y0;       // first tap of the window (may be far outside)
  float ax, ay;     // fractional offsets
};

// Window origin of a query at a level.  x / 2^l is an exact scaling, identical
// to the reference's division (core/corr.py:41).  Finite coordinates far outside
// the map give an all-zero window (zeros padding).  NaN or infinite coordinates
// give NaN in every sample of the query, as the reference's grid_sample does (a
// diverging training run must not turn into finite zeros): the window is taken
// as outside and the fractional offset as NaN, which every lerp propagates.
template <int R>
__device__ __forceinline__ QueryWin make_window(float x, float y, int level, bool active) {
  const float inv = __int_as_float((127 - level) << 23);   // 2^-level
  const float cx = x * inv, cy = y * inv;
  const float fx = floorf(cx), fy = floorf(cy);
  QueryWin w;
  const bool sane = active && fabsf(cx) < 1.0e7f && fabsf(cy) < 1.0e7f;
  const bool finite = (cx - cx == 0.f) && (cy - cy == 0.f);      // false for NaN and +-Inf
  const float off = (active && !finite) ? __int_as_float(0x7fc00000) : 0.f;
  w.x0 = sane ? (int)fx - R : -(1 << 24);
  w.y0 = sane ? (int)fy - R : -(1 << 24);
  w.ax = sane ? cx - fx : off;
  w.ay = sane ? cy - fy : off;
  return w;
}

// 16-byte asynchronous copy global -> shared, zero-filled when !ok (src-size 0).
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, bool ok) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
// One element as raw bits in the low part of a word (bf16: low half), 0 when !ok.
template <typename VolT>
__device__ __forceinline__ uint32_t ldg_elem(const VolT *p, bool ok) {
  uint32_t v;
  if (sizeof(VolT) == 2) {
    unsigned short h;
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.b16 %0, 0;\n\t"
        "@p ld.global.nc.b16 %0, [%1];\n\t}"
        : "=h"(h) : "l"(p), "r"((int)ok) : "memory");
    v = h;
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.b32 %0, 0;\n\t"
        "@p ld.global.nc.b32 %0, [%1];\n\t}"
        : "=r"(v) : "l"(p), "r"((int)ok) : "memory");
  }
  return v;
}

// word index (within a window) -> shared-memory word of query qq
template <typename W_>
__device__ __forceinline__ int smem_word(int word, int qq) {
  return word * 32 + ((qq * W_::K + (word >> 2)) & 31);
}

// Geometry of one level as the kernels need it.
struct LevelGeom {
  int Hl, Wl, tiles_x;
  long long slab;      // elements per query
};

// byte offset of unit `slot` of query qq inside a level's staged windows
template <typename W_>
__device__ __forceinline__ uint32_t unit_off(int qq, int slot) {
  return (uint32_t)(qq * W_::SU + (slot ^ ((qq >> 1) & 3))) * 16u;
}

// STAGE (see the file comment): copy the units of the warp's 32 windows into `win`
// (shared-memory byte address, W_::WIN_BYTES).  `lvl` points at the level,
// slab0 = b*N + q0.  Completion: cp_async_wait_all() + __syncwarp().
template <typename VolT, int R>
__device__ __forceinline__ void stage_windows(const VolT *__restrict__ lvl, uint32_t win, const QueryWin &w,
                                              long long slab0, const LevelGeom &g, int lane) {
  using W_ = Win<VolT, R>;
#pragma unroll 1
  for (int grp = 0; grp < 2; ++grp) {
#pragma unroll
    for (int it = 0; it < W_::ITER; ++it) {
      const int s = it * 32 + lane;                 // 0 .. 16*SL-1
      const int ql = s / W_::SL;                    // query within the group
      const int slot = s - ql * W_::SL;             // unit within the window: j * NCOL + c
      const int j = slot / W_::NCOL;
      const int c = slot - j * W_::NCOL;
      const int qq = grp * 16 + ql;
      const int x0 = __shfl_sync(0xffffffffu, w.x0, qq);
      const int y0 = __shfl_sync(0xffffffffu, w.y0, qq);
      const int y = y0 + j;
      const int tc = (x0 >> W_::LOG_EPV) + c;
      const bool ok = (unsigned)y < (unsigned)g.Hl && (unsigned)tc < (unsigned)g.tiles_x;
      const long long unit = ((long long)(y >> 2) * g.tiles_x + tc) * 4 + (y & 3);
      const VolT *p = lvl + (slab0 + qq) * g.slab + unit * W_::EPV;
      cp_async16(win + unit_off<W_>(qq, slot), ok ? p : lvl, ok);
    }
  }
}

// The one element a widest window needs beyond its NCOL units, per window row
// (raw bits; zero when this query does not need it or it lies outside the map).
template <typename VolT, int R>
__device__ __forceinline__ void load_extras(const VolT *__restrict__ lvl, const QueryWin &w, long long slab_q,
                                            const LevelGeom &g, uint32_t (&ex)[Win<VolT, R>::T]) {
  using W_ = Win<VolT, R>;
  if (W_::EXTRA) {
    const bool need = (w.x0 & (W_::EPV - 1)) == W_::EPV - 1;
    const int tc = (w.x0 >> W_::LOG_EPV) + W_::NCOL;
#pragma unroll
    for (int j = 0; j < W_::T; ++j) {
      const int y = w.y0 + j;
      const bool ok = need && (unsigned)y < (unsigned)g.Hl && (unsigned)tc < (unsigned)g.tiles_x;
      const long long unit = ((long long)(y >> 2) * g.tiles_x + tc) * 4 + (y & 3);
      ex[j] = ldg_elem<VolT>(lvl + slab_q * g.slab + unit * W_::EPV, ok);
    }
  } else {
#pragma unroll
    for (int j = 0; j < W_::T; ++j) ex[j] = 0u;
  }
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

// Row j of this lane's window as T floats, out of the staged units: read the
// row's NCOL units, shift by the lane's sub-tile offset o in registers.
template <typename VolT, int R>
__device__ __forceinline__ void read_row(uint32_t win, int lane, int j, int o, uint32_t extra,
                                         float (&v)[Win<VolT, R>::T]) {
  using W_ = Win<VolT, R>;
  constexpr int T = W_::T, NEED = W_::NEED;
  uint32_t wv[W_::ROWW + 3];
#pragma unroll
  for (int c = 0; c < W_::NCOL; ++c) {
    const uint4 u = lds128(win + unit_off<W_>(lane, j * W_::NCOL + c));
    wv[4 * c + 0] = u.x; wv[4 * c + 1] = u.y; wv[4 * c + 2] = u.z; wv[4 * c + 3] = u.w;
  }
  wv[W_::ROWW] = extra;
  wv[W_::ROWW + 1] = 0u;
  wv[W_::ROWW + 2] = 0u;
  const int m = sizeof(VolT) == 2 ? (o >> 1) : o;     // word shift, 0..3
  uint32_t xs[NEED + 2], ys[NEED];
#pragma unroll
  for (int k = 0; k < NEED + 2; ++k) xs[k] = (m & 1) ? wv[k + 1] : wv[k];
#pragma unroll
  for (int k = 0; k < NEED; ++k) ys[k] = (m & 2) ? xs[k + 2] : xs[k];
  if (sizeof(VolT) == 2) {
    const int sh = (o & 1) * 16;
#pragma unroll
    for (int k = 0; k < T / 2; ++k) {
      const float2 f = unpack2<VolT>(__funnelshift_r(ys[k], ys[k + 1], sh));
      v[2 * k] = f.x;
      v[2 * k + 1] = f.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < T; ++i) v[i] = __uint_as_float(ys[i]);
  }
}

// grid: B * ceil(N/32) strips; block: 32 * num_levels threads; dyn smem: L * WIN_BYTES
template <typename OutT>
__device__ __forceinline__ OutT to_out(float v);
template <>
__device__ __forceinline__ float to_out<float>(float v) { return v; }
template <>
__device__ __forceinline__ __half to_out<__half>(float v) { return __float2half_rn(v); }

// OutT = float: the reference's output.  OutT = __half: the same values rounded once to fp16, for callers whose
// next op runs under fp16 autocast (core/raft.py:127-128) and would cast the fp32 output itself.
template <typename VolT, int R, typename OutT>
__global__ void __launch_bounds__(32 * RC_MAX_LEVELS)
lookup_fwd_kernel(const VolT *__restrict__ pyr, const float *__restrict__ coords,
                  OutT *__restrict__ out, Levels lv, int N, int strips_per_pair, int out_pitch) {
  using W_ = Win<VolT, R>;
  extern __shared__ __align__(16) uint32_t smem_u[];
  const int lane = threadIdx.x & 31;
  const int l = threadIdx.x >> 5;
  const uint32_t win = (uint32_t)__cvta_generic_to_shared(smem_u) + l * W_::WIN_BYTES;

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
  LevelGeom g;
  g.Hl = lv.H[l]; g.Wl = lv.W[l]; g.tiles_x = lv.tiles_x[l]; g.slab = lv.slab[l];
  const VolT *lvl = pyr + lv.offset[l];
  const long long slab0 = (long long)b * N + q0;

  uint32_t ex[W_::T];
  stage_windows<VolT, R>(lvl, win, w, slab0, g, lane);
  load_extras<VolT, R>(lvl, w, slab0 + lane, g, ex);
  cp_async_wait_all();
  __syncwarp();

  if (!active) return;
  const int o = w.x0 & (W_::EPV - 1);
  const int C = lv.num_levels * W_::NW * W_::NW;
  // out_pitch: elements between consecutive channel planes (N, or N rounded up when a TMA reader needs 16-byte pitches)
  OutT *op = out + ((size_t)b * C + (size_t)l * W_::NW * W_::NW) * out_pitch + q;
  float hprev[W_::NW];
#pragma unroll
  for (int j = 0; j < W_::T; ++j) {
    float v[W_::T], h[W_::NW];
    read_row<VolT, R>(win, lane, j, o, ex[j], v);
#pragma unroll
    for (int i = 0; i < W_::NW; ++i) h[i] = fmaf(w.ax, v[i + 1] - v[i], v[i]);
    if (j > 0) {
#pragma unroll
      for (int i = 0; i < W_::NW; ++i)
        op[(size_t)(i * W_::NW + (j - 1)) * out_pitch] = to_out<OutT>(fmaf(w.ay, h[i] - hprev[i], hprev[i]));
    }
#pragma unroll
    for (int i = 0; i < W_::NW; ++i) hprev[i] = h[i];
  }
}

// words of tap-gradient staging per query of a level of the gradient pyramid (see lookup_bwd_kernel)
template <typename G_>
__host__ __device__ __forceinline__ int bwd_level_words(int Hl) { return Hl < G_::T ? Hl * G_::ROWW : G_::WPQ; }
// word offset of every level's staging inside the backward kernel's dynamic shared memory
struct BwdOffsets { int off[RC_MAX_LEVELS]; };

// Backward.  Same decomposition, one warp per level.
//   A. (only if d/dcoords is wanted) stage the forward windows as in the forward.
//   B. lane == query: fold the (2r+1)^2 output gradients into the (2r+2)^2 tap
//      gradients (separable, gather form) and park them in shared memory in the
//      unshifted unit layout of the fp32 GRADIENT pyramid; accumulate d/dcoords
//      from the forward taps on the way.
//   C. lane == unit: add the units into the gradient pyramid with 16-byte vector
//      reductions (red.global.add.v4.f32).  Every query owns its slabs, so each
//      address is hit once per launch; words of a unit outside the window (never
//      written in B) and padding columns are masked out.
template <typename VolT, int R>
__global__ void __launch_bounds__(32 * RC_MAX_LEVELS)
lookup_bwd_kernel(const VolT *__restrict__ pyr, const float *__restrict__ coords,
                  const float *__restrict__ grad_out, float *__restrict__ gpyr,
                  float *__restrict__ gcoords, Levels lv, Levels glv, const __grid_constant__ BwdOffsets bo, int N,
                  int strips_per_pair) {
  using W_ = Win<VolT, R>;     // forward window in the volume's layout
  using G_ = Win<float, R>;    // tap gradients in the fp32 gradient pyramid's layout
  constexpr int T = W_::T, NW = W_::NW;
  extern __shared__ __align__(16) uint32_t smem_u[];
  __shared__ float dc[RC_MAX_LEVELS][2][32];
  const int lane = threadIdx.x & 31;
  const int l = threadIdx.x >> 5;
  const bool want_dc = gcoords != nullptr;
  const bool want_vol = gpyr != nullptr;
  // A level with fewer rows than a window has taps (Hl < T: the coarsest level of the usual shapes) parks its tap
  // gradients by MAP row (Hl rows per query) instead of by window row (T rows): half the staging at the FlyingChairs
  // and Sintel shapes' level 3, which is what lets a fourth CTA live on an SM (54 instead of 61 KB).
  const int gHl = glv.H[l];
  const bool short_level = gHl < T;                       // warp-uniform
  const int gwin_words = want_vol ? 32 * bwd_level_words<G_>(gHl) : 0;
  uint32_t *gwin = smem_u + bo.off[l];
  const uint32_t win = (uint32_t)__cvta_generic_to_shared(gwin + gwin_words);

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
  const long long slab0 = (long long)b * N + q0;

  uint32_t ex[T];
#pragma unroll
  for (int j = 0; j < T; ++j) ex[j] = 0u;
  if (want_dc) {
    LevelGeom g;
    g.Hl = lv.H[l]; g.Wl = lv.W[l]; g.tiles_x = lv.tiles_x[l]; g.slab = lv.slab[l];
    const VolT *lvl = pyr + lv.offset[l];
    stage_windows<VolT, R>(lvl, win, w, slab0, g, lane);
    load_extras<VolT, R>(lvl, w, slab0 + lane, g, ex);
    cp_async_wait_all();
    __syncwarp();
  }

  LevelGeom gg;
  gg.Hl = glv.H[l]; gg.Wl = glv.W[l]; gg.tiles_x = glv.tiles_x[l]; gg.slab = glv.slab[l];
  float *glvl = want_vol ? gpyr + glv.offset[l] : nullptr;

  // ---- B
  float dx = 0.f, dy = 0.f;
  {
    const int o = w.x0 & (W_::EPV - 1);
    const int og = w.x0 & (G_::EPV - 1);
    const int C = lv.num_levels * NW * NW;
    const float *gp = grad_out + ((size_t)b * C + (size_t)l * NW * NW) * N + q;
    float hg_prev[T];   // x-folded gradient of output row j-1
    float cur[T];       // forward taps, row j
#pragma unroll
    for (int i = 0; i < T; ++i) { hg_prev[i] = 0.f; cur[i] = 0.f; }
    if (want_dc) read_row<VolT, R>(win, lane, 0, o, ex[0], cur);
#pragma unroll
    for (int j = 0; j <= NW; ++j) {
      float trow[T];    // tap-gradient row j
      if (j < NW) {
        float gj[NW], hg[T];
#pragma unroll
        for (int i = 0; i < NW; ++i) gj[i] = active ? __ldg(gp + (size_t)(i * NW + j) * N) : 0.f;
        if (want_dc) {
          float nxt[T];
          read_row<VolT, R>(win, lane, j + 1, o, ex[j + 1], nxt);
#pragma unroll
          for (int i = 0; i < NW; ++i) {
            const float d0 = cur[i + 1] - cur[i], d1 = nxt[i + 1] - nxt[i];
            dx = fmaf(gj[i], fmaf(w.ay, d1 - d0, d0), dx);
            const float e0 = nxt[i] - cur[i], e1 = nxt[i + 1] - cur[i + 1];
            dy = fmaf(gj[i], fmaf(w.ax, e1 - e0, e0), dy);
          }
#pragma unroll
          for (int i = 0; i < T; ++i) cur[i] = nxt[i];
        }
        // fold along x: hg[ii] = (1-ax) g[ii] + ax g[ii-1]
#pragma unroll
        for (int i = 0; i < T; ++i) {
          const float a = (i < NW) ? gj[i] : 0.f;
          const float c = (i > 0) ? gj[i - 1] : 0.f;
          hg[i] = fmaf(w.ax, c - a, a);
        }
        // tap-gradient row j is final: (1-ay) hg_j + ay hg_{j-1}
#pragma unroll
        for (int i = 0; i < T; ++i) {
          trow[i] = fmaf(w.ay, hg_prev[i] - hg[i], hg[i]);
          hg_prev[i] = hg[i];
        }
      } else {
#pragma unroll
        for (int i = 0; i < T; ++i) trow[i] = w.ay * hg_prev[i];
      }
      if (want_vol) {
        const int srow = short_level ? w.y0 + j : j;                 // staged row: the map row, or the window row
        const bool row_staged = !short_level || (unsigned)srow < (unsigned)gHl;
#pragma unroll
        for (int i = 0; i < T; ++i) {
          if (og + i < G_::ROWW) {
            if (row_staged) gwin[smem_word<G_>(srow * G_::ROWW + og + i, lane)] = __float_as_uint(trow[i]);
          } else {
            // the one tap beyond the staged units (og == 3, last tap): this lane owns it
            const int yy = w.y0 + j, xx = w.x0 + i;
            if (active && (unsigned)yy < (unsigned)gg.Hl && (unsigned)xx < (unsigned)gg.Wl)
              atomicAdd(glvl + (slab0 + lane) * gg.slab + tile_offset<float>(yy, xx, gg.tiles_x), trow[i]);
          }
        }
      }
    }
  }
  __syncwarp();

  // ---- C
  if (want_vol && short_level) {
    const int sl = gHl * G_::NCOL;                // units per query: (map row, unit column)
    for (int grp = 0; grp < 2; ++grp) {
      for (int it = 0; it < (16 * sl + 31) / 32; ++it) {      // every lane runs every round (the shuffles need all of them)
        const int s_raw = it * 32 + lane;
        const bool valid = s_raw < 16 * sl;                   // 16 * sl need not be a multiple of 32
        const int s = valid ? s_raw : 0;
        const int ql = s / sl;
        const int slot = s - ql * sl;
        const int yy = slot / G_::NCOL;
        const int c = slot - yy * G_::NCOL;
        const int qq = grp * 16 + ql;
        const int x0 = __shfl_sync(0xffffffffu, w.x0, qq);
        const int y0 = __shfl_sync(0xffffffffu, w.y0, qq);
        const int tc = (x0 >> G_::LOG_EPV) + c;
        const bool ok = valid && (unsigned)(yy - y0) < (unsigned)T && (unsigned)tc < (unsigned)gg.tiles_x;
        const long long unit = ((long long)(yy >> 2) * gg.tiles_x + tc) * 4 + (yy & 3);
        float *ptr = glvl + (slab0 + qq) * gg.slab + unit * G_::EPV;
        const int base = (slot * 4) * 32 + ((qq * G_::K + slot) & 31);
        const int xb = tc * G_::EPV;
        const int i0 = xb - x0;
        float f[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = i0 + e;
          const bool in = ok && i >= 0 && i < T && xb + e < gg.Wl;
          f[e] = in ? __uint_as_float(gwin[base + 32 * e]) : 0.f;
        }
        if (ok)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(ptr), "f"(f[0]), "f"(f[1]), "f"(f[2]),
                       "f"(f[3])
                       : "memory");
      }
    }
  } else if (want_vol) {
#pragma unroll 1
    for (int grp = 0; grp < 2; ++grp) {
#pragma unroll 5
      for (int it = 0; it < G_::ITER; ++it) {
        const int s = it * 32 + lane;
        const int ql = s / G_::SL;
        const int slot = s - ql * G_::SL;
        const int j = slot / G_::NCOL;
        const int c = slot - j * G_::NCOL;
        const int qq = grp * 16 + ql;
        const int x0 = __shfl_sync(0xffffffffu, w.x0, qq);
        const int y0 = __shfl_sync(0xffffffffu, w.y0, qq);
        const int yy = y0 + j;
        const int tc = (x0 >> G_::LOG_EPV) + c;
        const bool ok = (unsigned)yy < (unsigned)gg.Hl && (unsigned)tc < (unsigned)gg.tiles_x;
        const long long unit = ((long long)(yy >> 2) * gg.tiles_x + tc) * 4 + (yy & 3);
        float *ptr = glvl + (slab0 + qq) * gg.slab + unit * G_::EPV;
        const int base = (slot * 4) * 32 + ((qq * G_::K + slot) & 31);
        const int xb = tc * G_::EPV;          // x of the unit's first element
        const int i0 = xb - x0;               // window index of that element
        float f[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = i0 + e;
          const bool in = i >= 0 && i < T && xb + e < gg.Wl;
          f[e] = in ? __uint_as_float(gwin[base + 32 * e]) : 0.f;
        }
        // one 16-byte reduction per unit: performed at L2, nothing comes back, so
        // no load latency sits on the warp.  Each address is touched once per
        // launch (a query owns its slabs): the result does not depend on order.
        if (ok)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(ptr), "f"(f[0]), "f"(f[1]), "f"(f[2]),
                       "f"(f[3])
                       : "memory");
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

// Raise a kernel's dynamic shared memory limit to `max_bytes` once per device (`done`: one flag array per kernel
// instantiation, owned by the caller).  Setting the attribute to the launch's own size on every call would let two
// host threads with different level counts interleave set and launch.
static int raise_smem_limit_once(const void *kern, size_t max_bytes, std::atomic<int> (&done)[64]) {
  int dev = 0;
  RC_CUDA(cudaGetDevice(&dev));
  RC_REQUIRE(dev >= 0 && dev < 64, "device index %d out of range", dev);
  if (!done[dev].load(std::memory_order_acquire)) {
    RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_bytes));
    done[dev].store(1, std::memory_order_release);
  }
  return RC_OK;
}

template <typename VolT, int R, typename OutT>
int lookup_fwd_t(const VolT *pyr, const float *coords, const rc_layout &lay, OutT *out, int out_pitch,
                 cudaStream_t s) {
  using W_ = Win<VolT, R>;
  const int N = lay.N;
  const int strips = ceil_div(N, 32);
  const size_t smem = (size_t)lay.num_levels * W_::WIN_BYTES;
  auto kern = lookup_fwd_kernel<VolT, R, OutT>;
  static std::atomic<int> limit_set[64];
  if (int r = raise_smem_limit_once((const void *)kern, (size_t)RC_MAX_LEVELS * W_::WIN_BYTES, limit_set)) return r;
  kern<<<lay.B * strips, 32 * lay.num_levels, smem, s>>>(pyr, coords, out, to_levels(lay), N, strips,
                                                         out_pitch > 0 ? out_pitch : N);
  RC_LAUNCH_CHECK("lookup_fwd_kernel");
  return RC_OK;
}

template <typename VolT, int R>
int lookup_bwd_t(const VolT *pyr, const float *coords, const float *grad_out, const rc_layout &lay,
                 const rc_layout &glay, float *gpyr, float *gcoords, cudaStream_t s) {
  using W_ = Win<VolT, R>;
  using G_ = Win<float, R>;
  const int N = lay.N;
  const int strips = ceil_div(N, 32);
  BwdOffsets bo{};
  size_t smem = 0;
  for (int l = 0; l < lay.num_levels; ++l) {
    bo.off[l] = (int)(smem / 4);
    smem += (gpyr ? (size_t)32 * bwd_level_words<G_>(glay.H[l]) * 4 : 0) + (gcoords ? (size_t)W_::WIN_BYTES : 0);
  }
  auto kern = lookup_bwd_kernel<VolT, R>;
  static std::atomic<int> limit_set[64];
  const size_t max_smem = (size_t)RC_MAX_LEVELS * ((size_t)32 * G_::WPQ * 4 + (size_t)W_::WIN_BYTES);
  if (int r = raise_smem_limit_once((const void *)kern, max_smem, limit_set)) return r;
  kern<<<lay.B * strips, 32 * lay.num_levels, smem, s>>>(pyr, coords, grad_out, gpyr, gcoords,
                                                         to_levels(lay), to_levels(glay), bo, N, strips);
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

#define RC_DISPATCH_R2(FN, VT, OT, ...)                             \
  switch (radius) {                                                 \
    case 1: return FN<VT, 1, OT>(__VA_ARGS__);                      \
    case 2: return FN<VT, 2, OT>(__VA_ARGS__);                      \
    case 3: return FN<VT, 3, OT>(__VA_ARGS__);                      \
    case 4: return FN<VT, 4, OT>(__VA_ARGS__);                      \
    default: set_error("radius %d unsupported (1..4)", radius); return RC_EINVAL; \
  }

int launch_lookup_fwd(const void *pyramid, int vol_dtype, const float *coords, const rc_layout &lay,
                      int H, int W, int radius, float *out, cudaStream_t s) {
  (void)H; (void)W;
  if (vol_dtype == RC_VOL_BF16) {
    RC_DISPATCH_R2(lookup_fwd_t, __nv_bfloat16, float, (const __nv_bfloat16 *)pyramid, coords, lay, out, 0, s)
  } else if (vol_dtype == RC_VOL_F16) {
    RC_DISPATCH_R2(lookup_fwd_t, __half, float, (const __half *)pyramid, coords, lay, out, 0, s)
  } else {
    RC_DISPATCH_R2(lookup_fwd_t, float, float, (const float *)pyramid, coords, lay, out, 0, s)
  }
}

int launch_lookup_fwd_f16(const void *pyramid, int vol_dtype, const float *coords, const rc_layout &lay, int radius,
                          void *out_f16, int out_pitch, cudaStream_t s) {
  __half *out = (__half *)out_f16;
  if (vol_dtype == RC_VOL_BF16) {
    RC_DISPATCH_R2(lookup_fwd_t, __nv_bfloat16, __half, (const __nv_bfloat16 *)pyramid, coords, lay, out, out_pitch, s)
  } else if (vol_dtype == RC_VOL_F16) {
    RC_DISPATCH_R2(lookup_fwd_t, __half, __half, (const __half *)pyramid, coords, lay, out, out_pitch, s)
  } else {
    RC_DISPATCH_R2(lookup_fwd_t, float, __half, (const float *)pyramid, coords, lay, out, out_pitch, s)
  }
}

int launch_lookup_bwd(const void *pyramid, int vol_dtype, const float *coords, const float *grad_out,
                      const rc_layout &lay, const rc_layout &glay, int H, int W, int radius,
                      float *grad_pyramid, float *grad_coords, cudaStream_t s) {
  (void)H; (void)W;
  if (vol_dtype == RC_VOL_BF16) {
    RC_DISPATCH_R(lookup_bwd_t, __nv_bfloat16, (const __nv_bfloat16 *)pyramid, coords, grad_out, lay,
                  glay, grad_pyramid, grad_coords, s)
  } else if (vol_dtype == RC_VOL_F16) {
    RC_DISPATCH_R(lookup_bwd_t, __half, (const __half *)pyramid, coords, grad_out, lay, glay, grad_pyramid,
                  grad_coords, s)
  } else {
    RC_DISPATCH_R(lookup_bwd_t, float, (const float *)pyramid, coords, grad_out, lay, glay,
                  grad_pyramid, grad_coords, s)
  }
}

}  // namespace rc
