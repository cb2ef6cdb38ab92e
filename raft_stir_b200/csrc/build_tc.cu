// Tensor-core volume build with the pooled pyramid fused into the epilogue.
//
// Replaces CorrBlock.corr (core/corr.py:52-60: fp32 matmul + a separate
// full-volume / sqrt(D) pass) and the three avg_pool2d passes of
// CorrBlock.__init__ (core/corr.py:24-27) with ONE kernel:
//
//   C[b, q, t] = (1/sqrt(D)) * sum_d f1p[b][q][d] * f2p[b][t][d]
//
// as a tcgen05 GEMM (kind::f16, bf16 operands, fp32 accumulators in TMEM),
// 128 queries x 256 targets per tile, operands staged by TMA with the 128-byte
// swizzle.  A 256-target tile is one 8x32 spatial block of fmap2 (pack.cu), so a
// thread of the epilogue, which owns one query row of the accumulator, holds
// complete 2x2 / 4x4 / 8x8 pooling footprints in registers: levels 1..3 are
// in-register sums of the fp32 accumulators (no re-read of level 0, no extra
// rounding), written together with level 0.
//
// Warp roles (192 threads, 1 CTA per SM, persistent over a contiguous range of
// tiles):  warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner,
// warps 2..5 = epilogue (TMEM lane quarter = warp % 4).
// Pipelines: A (the 128-query operand) is stationary in smem for all tiles of
// one query block; B k-blocks stream through a ring; two 256-column TMEM
// accumulators let the epilogue of tile i overlap the MMAs of tile i+1.
// Epilogue output goes through swizzled smem staging and TMA tensor stores
// against 4-D maps (x, y, q, b) of each level, whose true extents (W>>l, H>>l,
// N) clip edge tiles -- which is exactly avg_pool2d's floor rule.
#include <cuda.h>

#include "common.cuh"

namespace rc {
namespace {

constexpr int BM = 128;          // queries per tile (UMMA M)
constexpr int BN = TB_N;         // 256 targets per tile (UMMA N)
constexpr int BK = 64;           // bf16 elements per k-block (128 bytes = swizzle span)
constexpr int UMMA_K = 16;
constexpr int A_KB_BYTES = BM * 128;   // 16 KB
constexpr int B_ST_BYTES = BN * 128;   // 32 KB
constexpr int TC_THREADS = 192;
constexpr int EPI_WARP0 = 2;

struct TcParams {
  int B, N, MT, NT, nbx, KB, num_levels, stages;
  int group_pairs, num_groups;      // pairs processed together (see TileWalk)
  float scale;                      // what is left of 1/sqrt(D) after the pack's exact power-of-two part
  // pyramid geometry (levels >= num_levels are unused)
  int Hl[4], Wl[4], tiles_y[4], tiles_x[4];
  long long slab[4];                // elements per query
  void *lvl[4];                     // base pointer of each level
};

// Order in which a CTA visits (pair b, query block m, target block n) tiles.
// Pairs are taken in groups of `group_pairs`; inside a group every CTA gets one
// contiguous range of the group's tiles in (b, m, n) order (balanced to within
// one tile; a range stays on one query block for NT tiles, so the A operand is
// reloaded rarely).  Grouping keeps the packed fmap2 of only a few pairs hot in
// L2 next to the output stream; one range over all pairs has every pair's
// operand in flight at once and misses.  Divisions happen once per range.
struct TileWalk {
  int grp, left;              // current group, tiles left in this CTA's range of it
  int b, m, n, by, bx;        // the next tile; (by, bx) = n's block coordinates
  __device__ __forceinline__ void init() { grp = -1; left = 0; b = m = n = by = bx = 0; }
  __device__ __forceinline__ bool next(const TcParams &p, int &ob, int &om, int &on, int &oby, int &obx) {
    while (left == 0) {
      ++grp;
      if (grp >= p.num_groups) return false;
      const int pair0 = grp * p.group_pairs;
      const int np = min(p.group_pairs, p.B - pair0);
      const int tpp = p.MT * p.NT;
      const long long tot = (long long)np * tpp;
      const int t0 = (int)(tot * blockIdx.x / gridDim.x);
      const int t1 = (int)(tot * (blockIdx.x + 1) / gridDim.x);
      left = t1 - t0;
      const int pb = t0 / tpp, r = t0 - pb * tpp;
      b = pair0 + pb;
      m = r / p.NT;
      n = r - m * p.NT;
      by = n / p.nbx;
      bx = n - by * p.nbx;
    }
    ob = b; om = m; on = n; oby = by; obx = bx;
    --left;
    if (++bx == p.nbx) { bx = 0; ++by; }
    if (++n == p.NT) {
      n = 0; by = 0; bx = 0;
      if (++m == p.MT) { m = 0; ++b; }
    }
    return true;
  }
  // after next(): does the following tile of this CTA use another A operand (or is there none)?
  __device__ __forceinline__ bool a_changes(int cb, int cm) const { return left == 0 || b != cb || m != cm; }
};

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Bounded wait: a protocol bug must surface as a launch error, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  unsigned long long t0 = 0;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
    if ((spin & 0x3ff) == 0x3ff) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 2000000000ull) __trap();   // 2 s
    }
  }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0,
                                            int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte-swizzled operand tile: rows of 128 bytes, 8-row groups 1024 bytes apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  const uint32_t lo = ((saddr >> 4) & 0x3fffu) | (1u << 16);            // start address, LBO = 1 (unused)
  const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);           // SBO, version 1, SWIZZLE_128B
  return ((uint64_t)hi << 32) | lo;
}
// kind::f16: D = F32 (bit 4), A = B = BF16 (bits 7, 10), both K-major, N>>3 at 17, M>>4 at 24
constexpr uint32_t IDESC_BF16 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                ((uint32_t)(BM >> 4) << 24);

// ---------------------------------------------------------------- epilogue staging
// An epilogue thread owns one query (TMEM lane) and sees the 8x32 target block
// two rows at a time.  Level 0 is written in STEPS of 256 bytes per query (bf16:
// 4 block rows = one row of 4 tiles, contiguous; fp32: 2 block rows = half of a
// row of 8 tiles, 8 pieces of 32 bytes): the thread parks its 16-byte units in a
// per-warp buffer, then the warp reads the buffer back query by query and issues
// coalesced 16-byte stores (a full 256-byte chunk per half warp).  Levels 1..3
// of the block (168 / 336 bytes per query) collect in a second buffer and leave
// once per tile the same way.
template <typename VolT>
struct Stg;
template <>
struct Stg<__nv_bfloat16> {
  static constexpr int TW = 8;            // tile width, elements
  static constexpr int L0_TILES = 4;      // level-0 tiles per block row of tiles
  static constexpr int L1_UNITS = 8, L2_UNITS = 2;      // 16-byte units per query
  static constexpr int NU = 11;           // L1 (8) + L2 (2) + L3 (1) units per query
};
template <>
struct Stg<float> {
  static constexpr int TW = 4;
  static constexpr int L0_TILES = 8;
  static constexpr int L1_UNITS = 16, L2_UNITS = 4;
  static constexpr int NU = 21;
};
constexpr int STEP_BYTES = 32 * 256;      // per-warp level-0 step buffer
template <typename VolT>
constexpr int stage_bytes_per_warp() { return STEP_BYTES + 32 * Stg<VolT>::NU * 16; }

__device__ __forceinline__ void sts16(uint32_t addr, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
}
__device__ __forceinline__ uint4 lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 p = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&p);
}
// 16-byte unit number `n` of a value row (8 bf16 or 4 fp32 elements) -> smem
template <typename VolT>
__device__ __forceinline__ void put_unit(uint32_t addr, const float *v) {
  if (sizeof(VolT) == 2)
    sts16(addr, pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  else
    sts16(addr, __float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3]));
}
// Step buffer: one 256-byte row per query, unit position XOR-rotated by the query
// (unit u of query q sits at q*256 + ((u ^ (q & 7)) << 4)), so that both the
// query-major writes and the unit-major read-back are conflict-free.  With a
// 256-byte-aligned buffer the XOR folds into the address: (row | rot) ^ (u << 4).

// ---------------------------------------------------------------- the kernel
// Read one level-0 step (256 bytes per query, 32 queries) back from the step buffer
// and store it: per instruction, 16 lanes cover the 256 bytes of one query.
//   ld_base : this lane's read address for iteration 0 (see the rotation above)
//   g       : this lane's global address for iteration 0; advances 2 queries per iteration
template <bool FULL>
__device__ __forceinline__ void flush_step(uint32_t ld_base, char *g, long long stride2, int hi, int nq,
                                           bool lane_ok) {
#pragma unroll 8
  for (int it = 0; it < 16; ++it) {
    const uint4 val = lds16((ld_base ^ (((2 * it) & 7) << 4)) + it * 512);
    if (FULL || (lane_ok && 2 * it + hi < nq)) *reinterpret_cast<uint4 *>(g) = val;
    g += stride2;
  }
}
// Same for a pooled level out of the per-tile buffer: UNITS 16-byte units per query.
template <int UNITS, bool FULL>
__device__ __forceinline__ void flush_pooled(uint32_t ld_base, int nu_bytes, char *g, long long stride_q, int sub,
                                             int nq, bool lane_ok) {
  constexpr int QPI = 32 / UNITS;       // queries per instruction
#pragma unroll
  for (int it = 0; it < 32 / QPI; ++it) {
    const uint4 val = lds16(ld_base + it * QPI * nu_bytes);
    if (FULL || (lane_ok && it * QPI + sub < nq)) *reinterpret_cast<uint4 *>(g) = val;
    g += QPI * stride_q;
  }
}

template <typename VolT, bool SCALE>
__global__ void __launch_bounds__(TC_THREADS, 1)
build_pyramid_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                        const TcParams p) {
  using S = Stg<VolT>;
  constexpr int ES = (int)sizeof(VolT);
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve (operand offsets are multiples of 1024)
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sA = smem_u32(smem);
  const uint32_t sB = sA + p.KB * A_KB_BYTES;
  const uint32_t sStage = sB + p.stages * B_ST_BYTES;
  const uint32_t sBar = sStage + 4 * stage_bytes_per_warp<VolT>();
  uint32_t *tmem_slot = (uint32_t *)(smem + (sBar - sA) + 128);
  // barrier table (8 bytes each)
  const uint32_t bar_a_full = sBar, bar_a_empty = sBar + 8;
  const uint32_t bar_b_full = sBar + 16;            // [stages]
  const uint32_t bar_b_empty = sBar + 16 + 8 * 4;   // [stages] (stages <= 4)
  const uint32_t bar_t_full = sBar + 16 + 8 * 8;    // [2]
  const uint32_t bar_t_empty = bar_t_full + 16;     // [2]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    mbar_init(bar_a_full, 1);
    mbar_init(bar_a_empty, 1);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(bar_b_full + 8 * s, 1);
      mbar_init(bar_b_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_t_full + 8 * s, 1);
      mbar_init(bar_t_empty + 8 * s, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      TileWalk walk;
      walk.init();
      int b, m, n, by, bx, cur_key = -1;
      uint32_t a_loads = 0, bs = 0;
      while (walk.next(p, b, m, n, by, bx)) {
        const int key = b * p.MT + m;
        if (key != cur_key) {
          if (a_loads > 0) mbar_wait(bar_a_empty, (a_loads - 1) & 1);
          mbar_expect_tx(bar_a_full, p.KB * A_KB_BYTES);
          for (int kb = 0; kb < p.KB; ++kb)
            tma_load_3d(sA + kb * A_KB_BYTES, &map_a, bar_a_full, kb * BK, m * BM, b);
          ++a_loads;
          cur_key = key;
        }
        for (int kb = 0; kb < p.KB; ++kb, ++bs) {
          const uint32_t st = bs % p.stages, ph = (bs / p.stages) & 1;
          mbar_wait(bar_b_empty + 8 * st, ph ^ 1);
          mbar_expect_tx(bar_b_full + 8 * st, B_ST_BYTES);
          tma_load_3d(sB + st * B_ST_BYTES, &map_b, bar_b_full + 8 * st, kb * BK, n * BN, b);
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      TileWalk walk;
      walk.init();
      int b, m, n, by, bx, cur_key = -1;
      uint32_t a_uses = 0, bs = 0, i = 0;
      while (walk.next(p, b, m, n, by, bx)) {
        const int key = b * p.MT + m;
        if (key != cur_key) {
          mbar_wait(bar_a_full, a_uses & 1);
          ++a_uses;
          cur_key = key;
        }
        const uint32_t acc = i & 1, use = i >> 1;
        ++i;
        mbar_wait(bar_t_empty + 8 * acc, (use & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < p.KB; ++kb, ++bs) {
          const uint32_t st = bs % p.stages, ph = (bs / p.stages) & 1;
          mbar_wait(bar_b_full + 8 * st, ph);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t ad = umma_desc_sw128(sA + kb * A_KB_BYTES + k * (UMMA_K * 2));
            const uint64_t bd = umma_desc_sw128(sB + st * B_ST_BYTES + k * (UMMA_K * 2));
            umma_bf16(d_tmem, ad, bd, IDESC_BF16, (kb | k) != 0);
          }
          umma_commit(bar_b_empty + 8 * st);   // frees the B stage when these MMAs retire
        }
        umma_commit(bar_t_full + 8 * acc);     // accumulator ready
        if (walk.a_changes(b, m)) umma_commit(bar_a_empty);   // last use of this A operand
      }
    }
  } else {
    // ===================================================== epilogue (4 warps)
    const int quarter = warp & 3;                 // TMEM lanes [32*quarter, +32)
    const uint32_t sbuf = sStage + (warp - EPI_WARP0) * stage_bytes_per_warp<VolT>();   // level-0 step buffer
    const uint32_t pbuf = sbuf + STEP_BYTES;                                            // levels 1..3
    const int L = p.num_levels;
    // per-lane constants of the staging buffers
    const uint32_t st_base = (sbuf + lane * 256) | ((lane & 7) << 4);    // write: unit u at st_base ^ (u << 4)
    const uint32_t pb_base = pbuf + lane * (S::NU * 16);                 // write: pooled unit u at pb_base + 16 u
    const int u16 = lane & 15, hi = lane >> 4;                           // read-back of a step: unit, query parity
    const uint32_t ld_base = (sbuf + hi * 256) | ((u16 ^ hi) << 4);
    const int l0_tile = ES == 2 ? (u16 >> 2) : (u16 >> 1);               // tile (inside the chunk) of unit u16
    constexpr int U1 = S::L1_UNITS, U2 = S::L2_UNITS;
    const int u1 = lane % U1, sub1 = lane / U1, u2 = lane % U2, sub2 = lane / U2;
    const uint32_t ld1 = pbuf + (sub1 * S::NU + u1) * 16;
    const int l2_tile = ES == 2 ? 0 : (u2 >> 1), l2_row = ES == 2 ? u2 : (u2 & 1);
    const uint32_t ld2 = pbuf + (sub2 * S::NU + U1 + (ES == 2 ? u2 : l2_row + 2 * l2_tile)) * 16;

    TileWalk walk;
    walk.init();
    int b, m, n, by, bx;
    uint32_t i = 0;
    while (walk.next(p, b, m, n, by, bx)) {
      const int x0 = bx * TB_W, y0 = by * TB_H;
      const int qw = m * BM + quarter * 32;            // first query of this warp
      const uint32_t acc = i & 1, use = i >> 1;
      ++i;
      const long long slab_q0 = (long long)b * p.N + qw;    // slab index of this warp's first query
      const int nq = min(32, p.N - qw);                     // valid queries of this warp (may be <= 0)

      // valid pooled cells of this block (floor rule of avg_pool2d; the rest is padding -> 0)
      const int nvx1 = p.Wl[1] - (x0 >> 1), nvy1 = p.Hl[1] - (y0 >> 1);
      const int nvx2 = p.Wl[2] - (x0 >> 2), nvy2 = p.Hl[2] - (y0 >> 2);
      const int nvx3 = p.Wl[3] - (x0 >> 3), nvy3 = p.Hl[3] - (y0 >> 3);
      const bool edge = (L > 1 && (nvx1 < 16 || nvy1 < 4)) || (L > 2 && (nvx2 < 8 || nvy2 < 2)) ||
                        (L > 3 && (nvx3 < 4 || nvy3 < 1));
      // level-0 chunk of this block: tiles tx0 .. tx0 + L0_TILES - 1 of tile rows (y0 >> 2), (y0 >> 2) + 1
      const int tx0 = x0 / S::TW;
      const int nt0 = min(S::L0_TILES, p.tiles_x[0] - tx0);
      const bool full0 = nq == 32 && nt0 == S::L0_TILES;
      const long long slab0B = p.slab[0] * ES;
      char *g0 = (char *)p.lvl[0] + ((long long)(y0 >> 2) * p.tiles_x[0] + tx0) * 64 + (slab_q0 + hi) * slab0B +
                 (ES == 2 ? u16 * 16 : (u16 >> 1) * 64 + (u16 & 1) * 16);
      const long long tile_row0B = (long long)p.tiles_x[0] * 64;

      mbar_wait(bar_t_full + 8 * acc, use & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quarter * 32) << 16);

      float p1prev[16], p2prev[8];
#pragma unroll
      for (int rp = 0; rp < 4; ++rp) {
        uint32_t r0[32], r1[32];
        tmem_ld32(taddr + (2 * rp) * 32, r0);
        tmem_ld32(taddr + (2 * rp + 1) * 32, r1);
        tmem_ld_wait();
        if (rp == 3) {
          // all TMEM reads of this tile are done: hand the accumulator back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_t_empty + 8 * acc);
        }
        float v0[32], v1[32], p1[16];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          v0[k] = __uint_as_float(r0[k]);
          v1[k] = __uint_as_float(r1[k]);
          if (SCALE) { v0[k] *= p.scale; v1[k] *= p.scale; }
        }
        // ---- level 0 units of block rows 2rp, 2rp+1
        if (ES == 2) {
          const int r = (2 * rp) & 3;                    // row inside the tile
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            put_unit<VolT>(st_base ^ ((k * 4 + r) << 4), v0 + 8 * k);
            put_unit<VolT>(st_base ^ ((k * 4 + r + 1) << 4), v1 + 8 * k);
          }
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            put_unit<VolT>(st_base ^ ((k * 2) << 4), v0 + 4 * k);
            put_unit<VolT>(st_base ^ ((k * 2 + 1) << 4), v1 + 4 * k);
          }
        }
        // ---- pooled levels, from the fp32 values
#pragma unroll
        for (int k = 0; k < 16; ++k) p1[k] = 0.25f * ((v0[2 * k] + v0[2 * k + 1]) + (v1[2 * k] + v1[2 * k + 1]));
        if (edge) {
#pragma unroll
          for (int k = 0; k < 16; ++k) p1[k] = (k < nvx1 && rp < nvy1) ? p1[k] : 0.f;
        }
        if (L > 1) {
#pragma unroll
          for (int k = 0; k < 16 / S::TW; ++k) put_unit<VolT>(pb_base + (k * 4 + rp) * 16, p1 + S::TW * k);
        }
        if (rp & 1) {
          float p2[8];
#pragma unroll
          for (int k = 0; k < 8; ++k)
            p2[k] = 0.25f * ((p1prev[2 * k] + p1prev[2 * k + 1]) + (p1[2 * k] + p1[2 * k + 1]));
          if (edge) {
#pragma unroll
            for (int k = 0; k < 8; ++k) p2[k] = (k < nvx2 && (rp >> 1) < nvy2) ? p2[k] : 0.f;
          }
          if (L > 2) {
            put_unit<VolT>(pb_base + (U1 + (rp >> 1)) * 16, p2);
            if (ES == 4) put_unit<VolT>(pb_base + (U1 + 2 + (rp >> 1)) * 16, p2 + 4);
          }
          if (rp == 3) {
            float p3[8];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float a = 0.25f * ((p2prev[2 * k] + p2prev[2 * k + 1]) + (p2[2 * k] + p2[2 * k + 1]));
              p3[k] = (!edge || (k < nvx3 && 0 < nvy3)) ? a : 0.f;
              p3[4 + k] = 0.f;
            }
            if (L > 3) put_unit<VolT>(pb_base + (U1 + U2) * 16, p3);
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) p2prev[k] = p2[k];
          }
        } else {
#pragma unroll
          for (int k = 0; k < 16; ++k) p1prev[k] = p1[k];
        }

        // ---- a level-0 step is complete: read it back query by query, coalesced stores
        if (ES == 4 || (rp & 1)) {
          __syncwarp();
          // bf16: step = one row of tiles; fp32: step = two of the four rows of a row of tiles
          const int trow = ES == 2 ? (rp >> 1) : (rp >> 1);
          const int half = ES == 2 ? 0 : (rp & 1) * 32;
          if ((y0 >> 2) + trow < p.tiles_y[0] && nt0 > 0) {
            char *g = g0 + trow * tile_row0B + half;
            if (full0) flush_step<true>(ld_base, g, 2 * slab0B, hi, nq, true);
            else flush_step<false>(ld_base, g, 2 * slab0B, hi, nq, l0_tile < nt0);
          }
          __syncwarp();
        }
      }

      // ---- levels 1..3 of this block
      __syncwarp();
      if (L > 1) {
        const int ty = y0 >> 3, tx1 = (x0 >> 1) / S::TW;
        const int nt = min(U1 / 4, p.tiles_x[1] - tx1);
        if (ty < p.tiles_y[1] && nt > 0) {
          const long long slabB = p.slab[1] * ES;
          char *g = (char *)p.lvl[1] + ((long long)ty * p.tiles_x[1] + tx1) * 64 + (slab_q0 + sub1) * slabB + u1 * 16;
          if (nq == 32 && nt == U1 / 4) flush_pooled<U1, true>(ld1, S::NU * 16, g, slabB, sub1, nq, true);
          else flush_pooled<U1, false>(ld1, S::NU * 16, g, slabB, sub1, nq, (u1 >> 2) < nt);
        }
      }
      if (L > 2) {
        const int ty = y0 >> 4, tx2 = (x0 >> 2) / S::TW;
        const int row0 = (y0 >> 2) & 3;                       // first of the block's two rows inside the tile
        if (ty < p.tiles_y[2]) {
          const long long slabB = p.slab[2] * ES;
          char *g = (char *)p.lvl[2] + ((long long)ty * p.tiles_x[2] + tx2) * 64 + (slab_q0 + sub2) * slabB +
                    l2_tile * 64 + (row0 + l2_row) * 16;
          flush_pooled<U2, false>(ld2, S::NU * 16, g, slabB, sub2, nq, tx2 + l2_tile < p.tiles_x[2]);
        }
      }
      if (L > 3) {
        const int ty = y0 >> 5, tx = (x0 >> 3) / S::TW;
        const int row = (y0 >> 3) & 3;
        if (ty < p.tiles_y[3] && tx < p.tiles_x[3] && lane < nq) {
          char *dst = (char *)p.lvl[3] + ((long long)ty * p.tiles_x[3] + tx) * 64 + (slab_q0 + lane) * p.slab[3] * ES +
                      row * 16;
          const uint4 val = lds16(pb_base + (U1 + U2) * 16);
          if (ES == 4) {
            *reinterpret_cast<uint4 *>(dst) = val;
          } else {
            // 4 cells = 8 bytes: the left or right half of a tile row.  The left writer also
            // zeroes the right half when no block to its right exists (padding columns).
            const bool right = (bx & 1) != 0;
            if (right) *reinterpret_cast<uint2 *>(dst + 8) = make_uint2(val.x, val.y);
            else if (bx + 1 < p.nbx) *reinterpret_cast<uint2 *>(dst) = make_uint2(val.x, val.y);
            else *reinterpret_cast<uint4 *>(dst) = val;
          }
        }
      }
      __syncwarp();
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
  }
}

// ---------------------------------------------------------------- host side
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

int encode(CUtensorMap *map, CUtensorMapDataType dt, int rank, const void *base, const cuuint64_t *dims,
           const cuuint64_t *strides_bytes, const cuuint32_t *box, CUtensorMapSwizzle sw, const char *what) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return RC_ECUDA;
  }
  cuuint32_t ones[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, dt, (cuuint32_t)rank, const_cast<void *>(base), dims, strides_bytes, box, ones,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(%s) failed with CUresult %d", what, (int)r);
    return RC_ECUDA;
  }
  return RC_OK;
}

template <typename VolT>
int build_tc_t(const void *f1p, const void *f2p, int B, int D, int H, int W, const rc_layout &lay,
               VolT *pyr, cudaStream_t s) {
  const int N = H * W;
  const int L = lay.num_levels;
  if (L > 4) {
    set_error("tensor-core build supports at most 4 pyramid levels (got %d)", L);
    return RC_EINVAL;
  }
  if (D % 8 != 0) {
    set_error("tensor-core build needs D %% 8 == 0 (got %d)", D);
    return RC_EINVAL;
  }
  TcParams p{};
  p.B = B;
  p.N = N;
  p.MT = ceil_div(N, BM);
  p.nbx = ceil_div(W, TB_W);
  p.NT = p.nbx * ceil_div(H, TB_H);
  p.KB = ceil_div(D, BK);
  p.num_levels = L;
  float pre_scale = 1.f;
  split_scale(D, &pre_scale, &p.scale);      // pre_scale was applied to fmap1 by the pack
  for (int l = 0; l < 4; ++l) {
    const bool on = l < L;
    p.Hl[l] = on ? lay.H[l] : 0;
    p.Wl[l] = on ? lay.W[l] : 0;
    p.tiles_y[l] = on ? lay.tiles_y[l] : 0;
    p.tiles_x[l] = on ? lay.tiles_x[l] : 0;
    p.slab[l] = on ? lay.slab[l] : 0;
    p.lvl[l] = on ? (void *)(pyr + lay.offset[l]) : nullptr;
  }
  const long long rows2 = (long long)p.NT * TB_N;

  int dev = 0, sms = 0, max_smem = 0;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  RC_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const size_t fixed = (size_t)p.KB * A_KB_BYTES + 4 * (size_t)stage_bytes_per_warp<VolT>() + 256 + 1024;
  p.stages = 0;
  for (int st = 4; st >= 2; --st)
    if (fixed + (size_t)st * B_ST_BYTES <= (size_t)max_smem) {
      p.stages = st;
      break;
    }
  if (p.stages == 0) {
    set_error("D = %d does not fit the tensor-core build's shared memory budget", D);
    return RC_EINVAL;
  }
  const size_t smem = fixed + (size_t)p.stages * B_ST_BYTES;

  // pairs per group: about two query blocks' worth of tiles per CTA and group
  const long long tpp = (long long)p.MT * p.NT;
  long long gp = (2LL * p.NT * sms + tpp - 1) / tpp;
  if (gp < 1) gp = 1;
  if (gp > B) gp = B;
  p.group_pairs = (int)gp;
  p.num_groups = ceil_div(B, p.group_pairs);

  CUtensorMap ma, mb;
  {
    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t str[2] = {(cuuint64_t)D * 2, (cuuint64_t)N * D * 2};
    cuuint32_t box[3] = {BK, BM, 1};
    int r = encode(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, f1p, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B, "fmap1");
    if (r) return r;
  }
  {
    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)rows2, (cuuint64_t)B};
    cuuint64_t str[2] = {(cuuint64_t)D * 2, (cuuint64_t)rows2 * D * 2};
    cuuint32_t box[3] = {BK, BN, 1};
    int r = encode(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, f2p, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B, "fmap2");
    if (r) return r;
  }

  auto kern = p.scale == 1.f ? build_pyramid_tc_kernel<VolT, false> : build_pyramid_tc_kernel<VolT, true>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long group_tiles = (long long)p.group_pairs * tpp;
  const int grid = group_tiles < sms ? (int)group_tiles : sms;
  kern<<<grid, TC_THREADS, smem, s>>>(ma, mb, p);
  RC_LAUNCH_CHECK("build_pyramid_tc_kernel");
  return RC_OK;
}

}  // namespace

int launch_build_tc(const void *f1p, const void *f2p, int B, int D, int H, int W, const rc_layout &lay,
                    int vol_dtype, int math, void *pyramid, cudaStream_t s) {
  if (math != RC_MATH_BF16) {
    set_error("tensor-core build: only RC_MATH_BF16 is implemented (math=%d)", math);
    return RC_EINVAL;
  }
  return vol_dtype == RC_VOL_BF16
             ? build_tc_t<__nv_bfloat16>(f1p, f2p, B, D, H, W, lay, (__nv_bfloat16 *)pyramid, s)
             : build_tc_t<float>(f1p, f2p, B, D, H, W, lay, (float *)pyramid, s);
}

}  // namespace rc
