// On-the-fly correlation on the tensor cores ("alternate" path, bf16 operands).
//
// Same contract as AlternateCorrBlock.__call__ (core/corr.py:74-91 calling
// alt_cuda_corr/correlation_kernel.cu:18-119 once per level): the N x N volume is never
// materialised, every iteration recomputes the dot products its windows need.  The reference and
// alt.cu do that tap by tap on the CUDA cores (100 dot products of length D per query and level).
// Here a CTA takes a spatial block of 8 x 16 queries and ONE pyramid level, forms the bounding box
// of the block's 128 windows in that level, and computes the block's LOCAL volume
//     C[128 queries, box targets] = f1[128, D] . f2_l[box, D]^T
// with tcgen05 (bf16 operands, fp32 accumulators in TMEM), 32 x 8 targets (one 128 x 256 UMMA tile)
// at a time.  Neighbouring queries' windows overlap almost entirely when the flow is smooth, so the
// box is little larger than one window (25 x 17 targets at level 0 for a constant flow) and the
// tensor cores do a few times the reference's arithmetic at two orders of magnitude its speed;
// a box of any size is walked chunk by chunk, so the result is the same for arbitrary coordinates,
// only slower.  Both operands come in through TMA boxes of the NHWC feature maps; coordinates
// outside a map arrive as zeros, which IS grid_sample's zeros padding.
//
// Two kernels.  alt_tc_persist_kernel (D <= 256, every configuration the reference ships) is the
// one that runs: persistent CTAs, query features resident per work unit, see the comment above it.
// alt_tc_fwd_kernel is the first version, kept for D > 256: one CTA per (query block, level),
// 192 threads: warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2..5 = epilogue
// (thread = query = TMEM lane).  Per chunk its epilogue reads the accumulator one target row
// (32 columns) at a time, bounces it through a per-warp shared-memory scratch and lets every lane
// pick the <= (2r+2) taps of its own window row; the windows collect in shared memory and are
// interpolated once all chunks are done.  Two TMEM accumulators alternate, so the MMAs of chunk c+1
// run while chunk c is picked apart.
#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace {

using namespace tc;

constexpr int QB_H = 8, QB_W = 16;          // query block (128 queries = UMMA M)
constexpr int CH_W = 32, CH_H = 8;          // target chunk (256 targets = UMMA N)
constexpr int KB = 64;                      // bf16 channels per k-block (128 bytes)
constexpr int A_BYTES = 128 * 128;          // 16 KB
constexpr int B_BYTES = 256 * 128;          // 32 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int NSTAGE = 3;
constexpr int SCR_STRIDE = 36;              // floats per lane in the bounce scratch (16-byte aligned rows)
constexpr int ATC_THREADS = 192;

struct AtcParams {
  int B, D, H, W, num_levels, nbx, nby;
  int half_operands;                  // 0: bf16 operands, 1: IEEE half (same kernels, other format code)
  int Hl[RC_MAX_LEVELS], Wl[RC_MAX_LEVELS];
  float out_scale;
};

struct AtcMaps {
  CUtensorMap f1;                     // fmap1 NHWC bf16: (D, W, H, B), box (64, 16, 8, 1)
  CUtensorMap f2[RC_MAX_LEVELS];      // fmap2 level l NHWC bf16: (D, W_l, H_l, B), box (64, 32, 8, 1)
};

// Window of one query at pyramid level l: origin (x0, y0) of its (2r+2)^2 taps, the bilinear fractions, and
// whether it meets the map at all (coords / 2^l as core/corr.py:40, taps as correlation_kernel.cu:59-116).
template <int R>
__device__ __forceinline__ bool window_at_level(float cxf, float cyf, int l, int Hl, int Wl, int &x0, int &y0,
                                                float &ax, float &ay) {
  constexpr int T = 2 * R + 2;
  const float inv = __int_as_float((127 - l) << 23);
  const float cx = cxf * inv, cy = cyf * inv;
  x0 = 0; y0 = 0; ax = 0.f; ay = 0.f;
  if (!(fabsf(cx) < 1.0e7f && fabsf(cy) < 1.0e7f)) {
    if (!((cx - cx == 0.f) && (cy - cy == 0.f))) ax = ay = __int_as_float(0x7fc00000);   // NaN / Inf coordinates -> NaN samples
    return false;
  }
  const float fx = floorf(cx), fy = floorf(cy);
  x0 = (int)fx - R; y0 = (int)fy - R;
  ax = cx - fx; ay = cy - fy;
  return x0 + T > 0 && x0 < Wl && y0 + T > 0 && y0 < Hl;
}

// dyn smem: [stages][A 16K | B 32K] (1024-aligned) | scratch 4 x 32 x 36 floats | windows 128 x (T*T+1) floats | barriers
template <int R>
__global__ void __launch_bounds__(ATC_THREADS, 1)
alt_tc_fwd_kernel(const __grid_constant__ AtcMaps maps, const float *__restrict__ coords, float *__restrict__ out,
                  const AtcParams p) {
  constexpr int NW = 2 * R + 1, T = 2 * R + 2, WS = T * T + 1;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sStage = smem_u32(smem);
  float *scratch = (float *)(smem + NSTAGE * STAGE_BYTES);
  float *wins = scratch + 4 * 32 * SCR_STRIDE;
  uint8_t *tail = (uint8_t *)(wins + 128 * WS);
  tail = (uint8_t *)(((uintptr_t)tail + 15) & ~(uintptr_t)15);
  const uint32_t sBar = smem_u32(tail);
  const uint32_t bar_full = sBar, bar_empty = sBar + 8 * NSTAGE;            // [NSTAGE] each
  const uint32_t bar_t_full = sBar + 16 * NSTAGE, bar_t_empty = bar_t_full + 16;   // [2] each
  uint32_t *tmem_slot = (uint32_t *)(tail + 16 * NSTAGE + 32);
  int *bbox = (int *)(tail + 16 * NSTAGE + 48);                             // x0, y0 (min), x1, y1 (max)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int l = blockIdx.y, b = blockIdx.z;
  const int qy0 = (blockIdx.x / p.nbx) * QB_H, qx0 = (blockIdx.x % p.nbx) * QB_W;
  const int Hl = p.Hl[l], Wl = p.Wl[l];
  const int N1 = p.H * p.W;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_t_full + 8 * s, 1);
      mbar_init(bar_t_empty + 8 * s, 4);
    }
    bbox[0] = 1 << 30; bbox[1] = 1 << 30; bbox[2] = -(1 << 30); bbox[3] = -(1 << 30);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  __syncthreads();

  // ---- every epilogue thread: its query, its window at this level, its share of the bounding box
  int x0 = 0, y0 = 0, q = -1;
  float ax = 0.f, ay = 0.f;
  bool relevant = false;
  if (warp >= 2) {
    const int ql = (warp & 3) * 32 + lane;               // TMEM lane = row of the A box = qy * 16 + qx
    const int yq = qy0 + ql / QB_W, xq = qx0 + ql % QB_W;
    float *w = wins + ql * WS;
#pragma unroll 4
    for (int k = 0; k < T * T; ++k) w[k] = 0.f;
    if (yq < p.H && xq < p.W) {
      q = yq * p.W + xq;
      const float inv = __int_as_float((127 - l) << 23);
      const float cx = __ldg(coords + ((size_t)b * 2 + 0) * N1 + q) * inv;
      const float cy = __ldg(coords + ((size_t)b * 2 + 1) * N1 + q) * inv;
      if (fabsf(cx) < 1.0e7f && fabsf(cy) < 1.0e7f) {
        const float fx = floorf(cx), fy = floorf(cy);
        x0 = (int)fx - R; y0 = (int)fy - R;
        ax = cx - fx; ay = cy - fy;
        relevant = x0 + T > 0 && x0 < Wl && y0 + T > 0 && y0 < Hl;     // the window meets the map
      } else if (!((cx - cx == 0.f) && (cy - cy == 0.f))) {
        ax = ay = __int_as_float(0x7fc00000);                          // NaN / Inf coordinates -> NaN samples
      }
    }
    if (relevant) {
      atomicMin(&bbox[0], max(x0, 0));
      atomicMin(&bbox[1], max(y0, 0));
      atomicMax(&bbox[2], min(x0 + T - 1, Wl - 1));
      atomicMax(&bbox[3], min(y0 + T - 1, Hl - 1));
    }
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int bx0 = bbox[0], by0 = bbox[1];
  const int nx = bbox[2] >= bx0 ? (bbox[2] - bx0) / CH_W + 1 : 0;
  const int ny = bbox[3] >= by0 ? (bbox[3] - by0) / CH_H + 1 : 0;
  const int nchunks = nx * ny;
  const int NKB = p.D / KB;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      uint32_t st = 0, ph = 0;
      for (int c = 0; c < nchunks; ++c) {
        const int cx0 = bx0 + (c % nx) * CH_W, cy0 = by0 + (c / nx) * CH_H;
        for (int kb = 0; kb < NKB; ++kb) {
          mbar_wait(bar_empty + 8 * st, ph ^ 1);
          mbar_expect_tx(bar_full + 8 * st, STAGE_BYTES);
          const uint32_t dst = sStage + st * STAGE_BYTES;
          tma_load_4d(dst, &maps.f1, bar_full + 8 * st, kb * KB, qx0, qy0, b);
          tma_load_4d(dst + A_BYTES, &maps.f2[l], bar_full + 8 * st, kb * KB, cx0, cy0, b);
          if (++st == NSTAGE) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      const uint32_t IDESC = p.half_operands ? idesc_f16(128, 256) : idesc_bf16(128, 256);
      uint32_t st = 0, ph = 0;
      for (int c = 0; c < nchunks; ++c) {
        const uint32_t acc = c & 1;
        mbar_wait(bar_t_empty + 8 * acc, ((c >> 1) & 1) ^ 1);
        fence_after();
        for (int kb = 0; kb < NKB; ++kb) {
          mbar_wait(bar_full + 8 * st, ph);
          fence_after();
          const uint32_t al = desc_lo(sStage + st * STAGE_BYTES), bl = desc_lo(sStage + st * STAGE_BYTES + A_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base + acc * 256, desc_make(al + 2 * k), desc_make(bl + 2 * k), IDESC, (kb | k) != 0);
          umma_commit(bar_empty + 8 * st);
          if (++st == NSTAGE) { st = 0; ph ^= 1; }
        }
        umma_commit(bar_t_full + 8 * acc);
      }
    }
  } else {
    // ===================================================== epilogue: pick the window taps out of every chunk
    const int quarter = warp & 3;
    const int ql = quarter * 32 + lane;
    float *w = wins + ql * WS;
    float *scr = scratch + (quarter * 32 + lane) * SCR_STRIDE;
    for (int c = 0; c < nchunks; ++c) {
      const uint32_t acc = c & 1;
      const int cx0 = bx0 + (c % nx) * CH_W, cy0 = by0 + (c / nx) * CH_H;
      mbar_wait(bar_t_full + 8 * acc, (c >> 1) & 1);
      fence_after();
      const uint32_t taddr = tmem_base + acc * 256 + ((uint32_t)(quarter * 32) << 16);
      const int i_lo = max(cx0 - x0, 0), i_hi = min(cx0 + CH_W - x0, T);     // window columns inside this chunk
#pragma unroll 1
      for (int rr = 0; rr < CH_H; ++rr) {
        const int j = cy0 + rr - y0;                                         // window row of this target row
        const bool need = relevant && j >= 0 && j < T && i_lo < i_hi;
        if (!__any_sync(0xffffffffu, need)) continue;
        uint32_t v[32];
        tmem_ld32(taddr + rr * CH_W, v);
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 8; ++k)
          *reinterpret_cast<uint4 *>(scr + 4 * k) = make_uint4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
        __syncwarp();
        if (need) {
          float tv[T];
#pragma unroll
          for (int i = 0; i < T; ++i) tv[i] = scr[min(max(x0 + i - cx0, 0), CH_W - 1)];    // all loads first
#pragma unroll
          for (int i = 0; i < T; ++i)
            if (i >= i_lo && i < i_hi) w[j * T + i] = tv[i];
        }
        __syncwarp();
      }
      fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_t_empty + 8 * acc);
    }

    // ---- bilinear interpolation of the finished window; channel = l*NW*NW + i*NW + j (x slow, y fast)
    if (q >= 0) {
      const int C = p.num_levels * NW * NW;
      float *op = out + ((size_t)b * C + (size_t)l * NW * NW) * N1 + q;
      float hprev[NW];
#pragma unroll
      for (int j = 0; j < T; ++j) {
        float h[NW];
#pragma unroll
        for (int i = 0; i < NW; ++i) {
          const float v0 = w[j * T + i], v1 = w[j * T + i + 1];
          h[i] = fmaf(ax, v1 - v0, v0);
        }
        if (j > 0) {
#pragma unroll
          for (int i = 0; i < NW; ++i)
            op[(size_t)(i * NW + (j - 1)) * N1] = fmaf(ay, h[i] - hprev[i], hprev[i]) * p.out_scale;
        }
#pragma unroll
        for (int i = 0; i < NW; ++i) hprev[i] = h[i];
      }
    }
  }

  fence_before();
  __syncthreads();
  fence_after();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
}

// ------------------------------------------------------------------------------------------------
// Persistent variant for D <= 256 (every configuration the reference ships).
//
// The kernel above reloads the query block's features with every k-block of every chunk, and every
// CTA pays its own prologue (barrier init, TMEM allocation, coordinates, pipeline fill) for two or
// three chunks of work.  Here one CTA per SM walks a list of work units; a unit is one 8 x 16 query
// block and a run of pyramid levels.  The block's own features (128 x D bf16, up to 64 KB) are
// loaded ONCE per unit and stay in shared memory while the chunks of all its levels stream through
// a 3-slot ring; the producer warp works out the bounding boxes of the coming levels itself and
// hands them to the other warps through a small descriptor ring, so loads, MMAs and tap picking of
// consecutive levels and units overlap without a CTA-wide barrier.  Chunks are cut to the box
// (ChunkWalk below).  Eight epilogue warps, two per TMEM lane quarter: each reads 16 accumulator
// columns of a target row and stores them straight into its query's window, the lane-dependent
// horizontal shift folded into the store address (no bounce buffer, no intra-warp barrier).
constexpr int P_SLOT_N = 256;               // most targets of one chunk (accumulator columns, rows of a ring slot);
                                            // 128 (six 16 KB slots, four accumulators) measured 23 % slower
constexpr int P_SLOT_BYTES = P_SLOT_N * 128;
constexpr int P_NSTAGE = 96 * 1024 / P_SLOT_BYTES;
constexpr int P_NACC = 512 / P_SLOT_N;      // accumulators in TMEM
constexpr int HALF_W = CH_W / 2;            // accumulator columns per epilogue warp and target row; narrow chunk width
constexpr int P_MAX_KB = 4;                 // resident query features: up to 4 k-blocks = 256 channels
constexpr int P_NDESC = 4;
constexpr int P_TAIL_BYTES = 16 * P_NSTAGE + 16 * P_MAX_KB + 16 * P_NACC + 16 * P_NDESC + 16 + 16 * P_NDESC + 16;
constexpr int P_THREADS = 320;              // warp 0 producer, warp 1 MMA, warps 2..9 epilogue (two per TMEM lane quarter)

struct PersistParams {
  AtcParams a;
  int n_units, n_items, nsplit, lpu;        // items = B * blocks; units = items * nsplit; lpu = levels per unit
};

// Bounding box of a unit's windows in one level, as the producer warp publishes it.
struct LevelDesc { int bx0, by0, bw, bh; };

struct PersistMaps {
  CUtensorMap f1;                     // fmap1 NHWC bf16: (D, W, H, B), box (64, 16, 8, 1)
  CUtensorMap f2w[RC_MAX_LEVELS];     // fmap2 level l: box (64, 32, 2, 1): two target rows of a wide chunk
  CUtensorMap f2n[RC_MAX_LEVELS];     // fmap2 level l: box (64, 16, 2, 1): two target rows of a narrow chunk
};

// The chunks of one level, derived from its box the same way by all three roles.  A box at most 16 targets wide
// (the coarse levels of a smooth flow) is cut into chunks of 16 x (2..16) targets, a wider one into 32 x (2..8);
// rows come in twos (the TMA box), so the accumulator of a chunk has N = 32 .. 256 columns and neither the loads
// nor the MMAs are spent on targets far outside the box.
struct ChunkWalk {
  int cw, rows_max, nx, ny, bh;
  bool narrow;
  __device__ explicit ChunkWalk(const LevelDesc &d) {
    narrow = d.bw <= HALF_W;
    cw = narrow ? HALF_W : CH_W;
    rows_max = P_SLOT_N / cw;
    bh = d.bh;
    const bool some = d.bw > 0 && d.bh > 0;
    nx = some ? (d.bw + cw - 1) / cw : 0;
    ny = some ? (d.bh + rows_max - 1) / rows_max : 0;
  }
  __device__ int rows(int iy) const { return min(rows_max, (bh - iy * rows_max + 1) & ~1); }
};

template <int R>
__global__ void __launch_bounds__(P_THREADS, 1)
alt_tc_persist_kernel(const __grid_constant__ PersistMaps maps, const float *__restrict__ coords,
                      float *__restrict__ out, const PersistParams pp) {
  constexpr int NW = 2 * R + 1, T = 2 * R + 2, WS = T * T + 1;
  const AtcParams &p = pp.a;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sA = smem_u32(smem);                                       // P_MAX_KB x 16 KB
  const uint32_t sB = sA + P_MAX_KB * A_BYTES;                              // P_NSTAGE ring slots
  float *wins = (float *)(smem + P_MAX_KB * A_BYTES + P_NSTAGE * P_SLOT_BYTES);  // 128 windows x (T*T+1) floats
  uint8_t *tail = (uint8_t *)(wins + 128 * WS);
  tail = (uint8_t *)(((uintptr_t)tail + 15) & ~(uintptr_t)15);
  const uint32_t sBar = smem_u32(tail);
  const uint32_t b_full = sBar, b_empty = b_full + 8 * P_NSTAGE;
  const uint32_t a_full = b_empty + 8 * P_NSTAGE, a_empty = a_full + 8 * P_MAX_KB;
  const uint32_t t_full = a_empty + 8 * P_MAX_KB, t_empty = t_full + 8 * P_NACC;
  const uint32_t d_full = t_empty + 8 * P_NACC, d_empty = d_full + 8 * P_NDESC;
  constexpr int BAR_BYTES = 16 * P_NSTAGE + 16 * P_MAX_KB + 16 * P_NACC + 16 * P_NDESC;
  uint32_t *tmem_slot = (uint32_t *)(tail + BAR_BYTES);
  LevelDesc *descs = (LevelDesc *)(tail + BAR_BYTES + 16);                  // P_NDESC x 16 B

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int N1 = p.H * p.W;
  const int NKB = p.D / KB;
  const int nblk = p.nbx * p.nby;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P_NSTAGE; ++s) { mbar_init(b_full + 8 * s, 1); mbar_init(b_empty + 8 * s, 1); }
    for (int s = 0; s < P_MAX_KB; ++s) { mbar_init(a_full + 8 * s, 1); mbar_init(a_empty + 8 * s, 1); }
    for (int s = 0; s < P_NACC; ++s) { mbar_init(t_full + 8 * s, 1); mbar_init(t_empty + 8 * s, 8); }
    for (int s = 0; s < P_NDESC; ++s) { mbar_init(d_full + 8 * s, 1); mbar_init(d_empty + 8 * s, 9); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== scout + TMA producer
    uint32_t st = 0, ph = 0, dcount = 0, nu = 0;
    // this lane's share of a block: queries 4*lane .. 4*lane+3 (a quarter of one block row); their coordinates
    // are fetched one unit ahead so that the first box of a unit is known the moment the previous unit is issued
    float cxn[4], cyn[4];
    bool insn[4];
    auto fetch = [&](int u) {
      if (u >= pp.n_units) return;
      const int item = u % pp.n_items;
      const int b = item / nblk, blk = item - b * nblk;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int ql = lane * 4 + k;
        const int yq = (blk / p.nbx) * QB_H + ql / QB_W, xq = (blk % p.nbx) * QB_W + ql % QB_W;
        insn[k] = yq < p.H && xq < p.W;
        const int q = insn[k] ? yq * p.W + xq : 0;
        cxn[k] = __ldg(coords + ((size_t)b * 2 + 0) * N1 + q);
        cyn[k] = __ldg(coords + ((size_t)b * 2 + 1) * N1 + q);
      }
    };
    fetch(blockIdx.x);
    for (int u = blockIdx.x; u < pp.n_units; u += gridDim.x, ++nu) {
      const int lg = u / pp.n_items, item = u - lg * pp.n_items;      // fine levels (most chunks) first
      const int b = item / nblk, blk = item - b * nblk;
      const int qy0 = (blk / p.nbx) * QB_H, qx0 = (blk % p.nbx) * QB_W;
      const int l_begin = lg * pp.lpu, l_end = min(p.num_levels, l_begin + pp.lpu);
      float cxs[4], cys[4];
      bool inside[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) { cxs[k] = cxn[k]; cys[k] = cyn[k]; inside[k] = insn[k]; }
      fetch(u + gridDim.x);
      if (lane == 0) {
        for (int kb = 0; kb < NKB; ++kb) {
          mbar_wait(a_empty + 8 * kb, (nu & 1) ^ 1);
          mbar_expect_tx(a_full + 8 * kb, A_BYTES);
          tma_load_4d(sA + kb * A_BYTES, &maps.f1, a_full + 8 * kb, kb * KB, qx0, qy0, b);
        }
      }
      for (int l = l_begin; l < l_end; ++l) {
        const int Hl = p.Hl[l], Wl = p.Wl[l];
        int mnx = 1 << 30, mny = 1 << 30, mxx = -(1 << 30), mxy = -(1 << 30);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          int x0, y0;
          float ax, ay;
          if (inside[k] && window_at_level<R>(cxs[k], cys[k], l, Hl, Wl, x0, y0, ax, ay)) {
            mnx = min(mnx, max(x0, 0));
            mny = min(mny, max(y0, 0));
            mxx = max(mxx, min(x0 + T - 1, Wl - 1));
            mxy = max(mxy, min(y0 + T - 1, Hl - 1));
          }
        }
        mnx = __reduce_min_sync(0xffffffffu, mnx);
        mny = __reduce_min_sync(0xffffffffu, mny);
        mxx = __reduce_max_sync(0xffffffffu, mxx);
        mxy = __reduce_max_sync(0xffffffffu, mxy);
        if (lane == 0) {
          const LevelDesc d{mnx, mny, mxx >= mnx ? mxx - mnx + 1 : 0, mxy >= mny ? mxy - mny + 1 : 0};
          const uint32_t slot = dcount % P_NDESC;
          mbar_wait(d_empty + 8 * slot, ((dcount / P_NDESC) & 1) ^ 1);
          descs[slot] = d;
          mbar_arrive(d_full + 8 * slot);
          const ChunkWalk cwk(d);
          const CUtensorMap *map = cwk.narrow ? &maps.f2n[l] : &maps.f2w[l];
          const uint32_t pair_bytes = 2 * cwk.cw * 128;                 // two target rows of one k-block
          for (int iy = 0; iy < cwk.ny; ++iy) {
            const int rows = cwk.rows(iy), cy0 = d.by0 + iy * cwk.rows_max;
            for (int ix = 0; ix < cwk.nx; ++ix) {
              const int cx0 = d.bx0 + ix * cwk.cw;
              for (int kb = 0; kb < NKB; ++kb) {
                mbar_wait(b_empty + 8 * st, ph ^ 1);
                mbar_expect_tx(b_full + 8 * st, (rows / 2) * pair_bytes);
                for (int r2 = 0; r2 < rows / 2; ++r2)
                  tma_load_4d(sB + st * P_SLOT_BYTES + r2 * pair_bytes, map, b_full + 8 * st, kb * KB, cx0, cy0 + 2 * r2, b);
                if (++st == P_NSTAGE) { st = 0; ph ^= 1; }
              }
            }
          }
        }
        ++dcount;
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      uint32_t st = 0, ph = 0, dcount = 0, nu = 0, g = 0;
      for (int u = blockIdx.x; u < pp.n_units; u += gridDim.x, ++nu) {
        const int lg = u / pp.n_items;
        const int l_begin = lg * pp.lpu, l_end = min(p.num_levels, l_begin + pp.lpu);
        bool released = false, first = true;
        for (int l = l_begin; l < l_end; ++l, ++dcount) {
          const uint32_t slot = dcount % P_NDESC;
          mbar_wait(d_full + 8 * slot, (dcount / P_NDESC) & 1);
          const ChunkWalk cwk(descs[slot]);
          mbar_arrive(d_empty + 8 * slot);
          const int nchunks = cwk.nx * cwk.ny;
          for (int c = 0, ix = 0, iy = 0; c < nchunks; ++c, ++g) {
            const uint32_t acc = g % P_NACC;
            const bool last = (l == l_end - 1) && (c == nchunks - 1);
            const int Nc = cwk.rows(iy) * cwk.cw;                              // N = targets of this chunk
            const uint32_t IDESC = p.half_operands ? idesc_f16(128, Nc) : idesc_bf16(128, Nc);
            if (++ix == cwk.nx) { ix = 0; ++iy; }
            mbar_wait(t_empty + 8 * acc, ((g / P_NACC) & 1) ^ 1);
            fence_after();
            for (int kb = 0; kb < NKB; ++kb) {
              if (first) mbar_wait(a_full + 8 * kb, nu & 1);
              mbar_wait(b_full + 8 * st, ph);
              fence_after();
              const uint32_t al = desc_lo(sA + kb * A_BYTES), bl = desc_lo(sB + st * P_SLOT_BYTES);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_bf16(tmem_base + acc * P_SLOT_N, desc_make(al + 2 * k), desc_make(bl + 2 * k), IDESC, (kb | k) != 0);
              umma_commit(b_empty + 8 * st);
              if (last) umma_commit(a_empty + 8 * kb);      // this resident k-block is free once its MMAs are done
              if (++st == P_NSTAGE) { st = 0; ph ^= 1; }
            }
            umma_commit(t_full + 8 * acc);
            first = false;
            released |= last;
          }
        }
        if (!released) {
          // the last level had no chunk: hand the resident features back here
          for (int kb = 0; kb < NKB; ++kb) {
            if (first) mbar_wait(a_full + 8 * kb, nu & 1);
            umma_commit(a_empty + 8 * kb);
          }
        }
      }
    }
  } else {
    // ===================================================== epilogue: pick the window taps out of every chunk
    // Two warps per TMEM lane quarter (same 32 queries): warp `half` takes accumulator columns
    // [16*half, 16*half+16) of every target row and, afterwards, half of the output rows.
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    const uint32_t pair_bar = 1 + quarter;                 // named barrier of the two warps of a quarter
    const int ql = quarter * 32 + lane;                    // TMEM lane = row of the A box = qy * 16 + qx
    float *w = wins + ql * WS;
    uint32_t dcount = 0, g = 0;
    // this thread's query in a unit and its coordinates, fetched one unit ahead of their use
    int q_nx = -1, b_nx = 0;
    float cx_nx = 0.f, cy_nx = 0.f;
    auto fetch = [&](int u) {
      if (u >= pp.n_units) return;
      const int item = u % pp.n_items;
      b_nx = item / nblk;
      const int blk = item - b_nx * nblk;
      const int yq = (blk / p.nbx) * QB_H + ql / QB_W, xq = (blk % p.nbx) * QB_W + ql % QB_W;
      q_nx = (yq < p.H && xq < p.W) ? yq * p.W + xq : -1;
      if (q_nx >= 0) {
        cx_nx = __ldg(coords + ((size_t)b_nx * 2 + 0) * N1 + q_nx);
        cy_nx = __ldg(coords + ((size_t)b_nx * 2 + 1) * N1 + q_nx);
      }
    };
    fetch(blockIdx.x);
    // rows of the (2r+2)-row window this warp clears, and output rows it interpolates
    constexpr int ZSPLIT = T / 2, OSPLIT = (NW + 1) / 2;
    const int z_begin = half ? ZSPLIT : 0, z_end = half ? T : ZSPLIT;
    const int o_begin = half ? OSPLIT : 0, o_end = half ? NW : OSPLIT;
    for (int u = blockIdx.x; u < pp.n_units; u += gridDim.x) {
      const int lg = u / pp.n_items;                                  // fine levels (most chunks) first
      const int l_begin = lg * pp.lpu, l_end = min(p.num_levels, l_begin + pp.lpu);
      const int q = q_nx, b = b_nx;
      const float cxf = cx_nx, cyf = cy_nx;
      fetch(u + gridDim.x);
      for (int l = l_begin; l < l_end; ++l, ++dcount) {
        int x0 = 0, y0 = 0;
        float ax = 0.f, ay = 0.f;
        const bool relevant = q >= 0 && window_at_level<R>(cxf, cyf, l, p.Hl[l], p.Wl[l], x0, y0, ax, ay);
        named_bar_sync(pair_bar, 64);                      // the partner is done reading the previous window
        // every in-map tap of a relevant window lies in some chunk and is overwritten below; only windows that
        // leave the map (or miss it) keep zeros
        if (!relevant || x0 < 0 || y0 < 0 || x0 + T > p.Wl[l] || y0 + T > p.Hl[l]) {
          for (int k = z_begin * T; k < z_end * T; ++k) w[k] = 0.f;
        }
        named_bar_sync(pair_bar, 64);                      // cleared before either warp writes taps
        const uint32_t slot = dcount % P_NDESC;
        mbar_wait(d_full + 8 * slot, (dcount / P_NDESC) & 1);
        const LevelDesc d = descs[slot];
        __syncwarp();
        if (lane == 0) mbar_arrive(d_empty + 8 * slot);
        const ChunkWalk cwk(d);
        // wide chunk: this warp takes columns [16*half, +16) of every target row; narrow chunk (16 columns): every
        // other row.  Either way one tcgen05.ld of 16 columns per target row.
        const int coff = cwk.narrow ? 0 : half * HALF_W;
        const int rstep = cwk.narrow ? 2 : 1, rr0 = cwk.narrow ? half : 0;
        for (int iy = 0; iy < cwk.ny; ++iy) {
          const int rows = cwk.rows(iy), cy0 = d.by0 + iy * cwk.rows_max;
          for (int ix = 0; ix < cwk.nx; ++ix, ++g) {
            const uint32_t acc = g % P_NACC;
            const int hx0 = d.bx0 + ix * cwk.cw + coff;                          // x of this warp's 16 target columns
            const int i_lo = max(hx0 - x0, 0), i_hi = min(hx0 + HALF_W - x0, T); // window columns inside them
            const bool cols = relevant && i_lo < i_hi;
            // tap i of the window sits in accumulator column sft + i of these 16: the lane-dependent shift goes
            // into the store address, a 16-bit mask says which columns are taps of this lane's window at all
            const int sft = x0 - hx0;
            uint32_t mask = 0;
            if (cols) mask = (((1u << (i_hi - i_lo)) - 1u) << (sft + i_lo)) & 0xffffu;
            const bool any_cols = __any_sync(0xffffffffu, cols);
            mbar_wait(t_full + 8 * acc, (g / P_NACC) & 1);
            fence_after();
            const uint32_t taddr = tmem_base + acc * P_SLOT_N + coff + ((uint32_t)(quarter * 32) << 16);
            if (any_cols) {
#pragma unroll 1
              for (int rr = rr0; rr < rows; rr += 2 * rstep) {
                const int ra = rr, rb = rr + rstep;                              // two target rows per step
                const int j0 = cy0 + ra - y0, j1 = cy0 + rb - y0;                // their window rows
                const bool need0 = cols && j0 >= 0 && j0 < T, need1 = cols && rb < rows && j1 >= 0 && j1 < T;
                const bool any0 = __any_sync(0xffffffffu, need0), any1 = __any_sync(0xffffffffu, need1);
                if (!(any0 || any1)) continue;
                uint32_t v0[HALF_W], v1[HALF_W];
                if (any0) tmem_ld16(taddr + ra * cwk.cw, v0);
                if (any1) tmem_ld16(taddr + rb * cwk.cw, v1);
                tmem_ld_wait();
                if (need0) {
                  float *wr = w + j0 * T - sft;
#pragma unroll
                  for (int k = 0; k < HALF_W; ++k)
                    if (mask & (1u << k)) wr[k] = __uint_as_float(v0[k]);
                }
                if (need1) {
                  float *wr = w + j1 * T - sft;
#pragma unroll
                  for (int k = 0; k < HALF_W; ++k)
                    if (mask & (1u << k)) wr[k] = __uint_as_float(v1[k]);
                }
              }
            }
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(t_empty + 8 * acc);
          }
        }
        named_bar_sync(pair_bar, 64);                      // both warps' taps are in place

        // ---- bilinear interpolation of the finished window; channel = l*NW*NW + i*NW + j (x slow, y fast)
        if (q >= 0) {
          const int C = p.num_levels * NW * NW;
          float *op = out + ((size_t)b * C + (size_t)l * NW * NW) * N1 + q;
          float hprev[NW];
#pragma unroll
          for (int jj = 0; jj <= OSPLIT; ++jj) {
            const int j = o_begin + jj;                    // window row; output row j - 1
            if (j > o_end) break;
            float h[NW];
#pragma unroll
            for (int i = 0; i < NW; ++i) {
              const float v0 = w[j * T + i], v1 = w[j * T + i + 1];
              h[i] = fmaf(ax, v1 - v0, v0);
            }
            if (jj > 0) {
#pragma unroll
              for (int i = 0; i < NW; ++i)
                op[(size_t)(i * NW + (j - 1)) * N1] = fmaf(ay, h[i] - hprev[i], hprev[i]) * p.out_scale;
            }
#pragma unroll
            for (int i = 0; i < NW; ++i) hprev[i] = h[i];
          }
        }
      }
    }
  }

  fence_before();
  __syncthreads();
  fence_after();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
}

}  // namespace

namespace {

template <typename OpT>
__global__ void __launch_bounds__(256)
f32_to_16_kernel(const float *__restrict__ src, OpT *__restrict__ dst, long long n4) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4 *>(src) + i);
    reinterpret_cast<uint2 *>(dst)[i] = make_uint2(pack2<OpT>(v.x, v.y), pack2<OpT>(v.z, v.w));
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      ptr = nullptr;
    return (EncodeTiledFn)ptr;
  }();
  return fn;
}

// NHWC bf16 feature map as a 4-D tensor (D, W, H, B) with a (64, bw, bh, 1) box, 128-byte swizzle
int encode_nhwc(CUtensorMap *map, const void *base, int B, int D, int H, int W, int bw, int bh, int half_operands) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return RC_ECUDA;
  }
  cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t str[3] = {(cuuint64_t)D * 2, (cuuint64_t)W * D * 2, (cuuint64_t)H * W * D * 2};
  cuuint32_t box[4] = {(cuuint32_t)KB, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  cuuint32_t ones[4] = {1, 1, 1, 1};
  CUresult r = fn(map, half_operands ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(base), dims, str, box, ones,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(feature map %dx%dx%d) failed with CUresult %d", H, W, D, (int)r);
    return RC_ECUDA;
  }
  return RC_OK;
}

template <int R>
int alt_tc_launch(const AtcMaps &maps, const float *coords, float *out, const AtcParams &p, cudaStream_t s) {
  constexpr int T = 2 * R + 2;
  const size_t smem = 1024 + (size_t)NSTAGE * STAGE_BYTES + 4 * 32 * SCR_STRIDE * sizeof(float) +
                      128 * (T * T + 1) * sizeof(float) + 256;
  auto kern = alt_tc_fwd_kernel<R>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(p.nbx * p.nby, p.num_levels, p.B);
  kern<<<grid, ATC_THREADS, smem, s>>>(maps, coords, out, p);
  RC_LAUNCH_CHECK("alt_tc_fwd_kernel");
  return RC_OK;
}

// Work split of the persistent kernel: a unit is a query block with num_levels / nsplit levels.  One unit per
// block keeps the block's features resident longest; small batches are cut finer so that every SM has work.
template <int R>
int alt_tc_launch_persistent(const PersistMaps &maps, const float *coords, float *out, const AtcParams &p,
                             cudaStream_t s) {
  constexpr int T = 2 * R + 2;
  int dev = 0, sms = 0;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const long long items = (long long)p.B * p.nbx * p.nby;
  PersistParams pp{};
  pp.a = p;
  double best = 0.0;
  for (int ns = 1; ns <= p.num_levels; ns *= 2) {
    const int lpu = ceil_div(p.num_levels, ns);
    const long long units = items * ns;
    const double cost = (double)((units + sms - 1) / sms) * (2.25 * lpu + 0.5);   // chunks + feature load
    if (ns == 1 || cost < best) { best = cost; pp.nsplit = ns; pp.lpu = lpu; }
  }
  const long long units = items * pp.nsplit;
  if (units > 0x7fffffffLL) {
    set_error("alt lookup: %lld work units exceed the 32-bit range", units);
    return RC_EINVAL;
  }
  pp.n_units = (int)units;
  pp.n_items = (int)items;
  const size_t smem = 1024 + (size_t)P_MAX_KB * A_BYTES + (size_t)P_NSTAGE * P_SLOT_BYTES +
                      128 * (T * T + 1) * sizeof(float) + P_TAIL_BYTES;
  auto kern = alt_tc_persist_kernel<R>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = (int)(units < sms ? units : sms);
  if (grid == 0) return RC_OK;
  kern<<<grid, P_THREADS, smem, s>>>(maps, coords, out, pp);
  RC_LAUNCH_CHECK("alt_tc_persist_kernel");
  return RC_OK;
}

}  // namespace

int launch_alt_features_16(const float *src, void *dst, long long elems, int half_operands, cudaStream_t s) {
  const long long n4 = elems / 4;
  if (n4 == 0) return RC_OK;
  const int blocks = (int)((n4 + 255) / 256 < 148LL * 8 ? (n4 + 255) / 256 : 148LL * 8);
  if (half_operands) f32_to_16_kernel<__half><<<blocks, 256, 0, s>>>(src, (__half *)dst, n4);
  else f32_to_16_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>(src, (__nv_bfloat16 *)dst, n4);
  RC_LAUNCH_CHECK("f32_to_16_kernel");
  return RC_OK;
}

int launch_alt_lookup_tc(const void *f1_bf16, const void *f2_pyr_bf16, const float *coords, int B, int D, int H,
                         int W, int num_levels, int radius, const int64_t *off, int half_operands, float *out,
                         cudaStream_t s) {
  AtcParams p{};
  p.B = B; p.D = D; p.H = H; p.W = W; p.num_levels = num_levels;
  p.half_operands = half_operands;
  p.nbx = ceil_div(W, QB_W); p.nby = ceil_div(H, QB_H);
  p.out_scale = 1.0f / sqrtf((float)D);
  for (int l = 0; l < RC_MAX_LEVELS; ++l) {
    const int ll = l < num_levels ? l : 0;
    p.Hl[l] = H >> ll; p.Wl[l] = W >> ll;
  }
  auto level = [&](int l) { return (const __nv_bfloat16 *)f2_pyr_bf16 + off[l < num_levels ? l : 0]; };
  int r;
  if (D <= P_MAX_KB * KB) {
    // persistent kernel: query features resident, fmap2 in boxes of two target rows (32 or 16 targets wide)
    // The nine descriptors depend only on the buffers and the shape: a block is looked up once per refinement
    // iteration with the same ones, so the last few sets are kept per host thread (descriptors are plain data
    // holding addresses: a stale entry can only be hit by a call that passes the very same buffers and shape).
    struct Cached { const void *f1, *f2; int B, D, H, W, L, half; PersistMaps maps; };
    thread_local Cached cache[4];
    thread_local int next = 0;
    const PersistMaps *mp = nullptr;
    for (const Cached &c : cache)
      if (c.f1 == f1_bf16 && c.f2 == f2_pyr_bf16 && c.B == B && c.D == D && c.H == H && c.W == W && c.L == num_levels &&
          c.half == half_operands)
        mp = &c.maps;
    if (!mp) {
      Cached &c = cache[next];
      c.f1 = nullptr;                                     // invalid while being rewritten
      if ((r = encode_nhwc(&c.maps.f1, f1_bf16, B, D, H, W, QB_W, QB_H, half_operands))) return r;
      for (int l = 0; l < RC_MAX_LEVELS; ++l) {
        if ((r = encode_nhwc(&c.maps.f2w[l], level(l), B, D, p.Hl[l], p.Wl[l], CH_W, 2, half_operands))) return r;
        if ((r = encode_nhwc(&c.maps.f2n[l], level(l), B, D, p.Hl[l], p.Wl[l], HALF_W, 2, half_operands))) return r;
      }
      c.f1 = f1_bf16; c.f2 = f2_pyr_bf16; c.B = B; c.D = D; c.H = H; c.W = W; c.L = num_levels; c.half = half_operands;
      next = (next + 1) % 4;
      mp = &c.maps;
    }
    const PersistMaps &maps = *mp;
    switch (radius) {
      case 1: return alt_tc_launch_persistent<1>(maps, coords, out, p, s);
      case 2: return alt_tc_launch_persistent<2>(maps, coords, out, p, s);
      case 3: return alt_tc_launch_persistent<3>(maps, coords, out, p, s);
      case 4: return alt_tc_launch_persistent<4>(maps, coords, out, p, s);
      default: set_error("radius %d unsupported (1..4)", radius); return RC_EINVAL;
    }
  }
  AtcMaps maps;
  if ((r = encode_nhwc(&maps.f1, f1_bf16, B, D, H, W, QB_W, QB_H, half_operands))) return r;
  for (int l = 0; l < RC_MAX_LEVELS; ++l)
    if ((r = encode_nhwc(&maps.f2[l], level(l), B, D, p.Hl[l], p.Wl[l], CH_W, CH_H, half_operands))) return r;
  switch (radius) {
    case 1: return alt_tc_launch<1>(maps, coords, out, p, s);
    case 2: return alt_tc_launch<2>(maps, coords, out, p, s);
    case 3: return alt_tc_launch<3>(maps, coords, out, p, s);
    case 4: return alt_tc_launch<4>(maps, coords, out, p, s);
    default: set_error("radius %d unsupported (1..4)", radius); return RC_EINVAL;
  }
}

}  // namespace rc
