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
// swizzle.  A 256-target tile is one 8x32 spatial block of fmap2 (pack.cu), so
// complete 2x2 / 4x4 / 8x8 pooling footprints sit in one accumulator row:
// levels 1..3 are in-register sums of the fp32 accumulators (no re-read of
// level 0, no extra rounding), written together with level 0, all in the
// 64-byte-tile layout of rc_layout.
//
// Warp roles (320 threads, 1 CTA per SM, persistent):  warp 0 = TMA producer,
// warp 1 = MMA issuer + TMEM owner, warps 2..9 = epilogue in two groups of four
// (TMEM lane quarter = warp % 4).  Group g takes block columns 16g..16g+15 of
// all 8 rows of every tile: a complete 8x16 sub-block, so all pooled levels stay
// thread-local.  An epilogue warp copies its 128 accumulator values to
// registers with eight tcgen05.ld and hands the accumulator back at once; the
// conversion, pooling and stores then overlap the MMAs of the next tiles.
// Pipelines: A (the 128-query operand) is stationary in smem for all tiles of
// one query block; B k-blocks (32 KB) stream through a ring; two 256-column
// TMEM accumulators alternate.  The single-thread producer and MMA loops keep
// stage index and phase incrementally (no divisions): their per-stage overhead
// is on the critical path of the tensor pipe.
//
// Default form (template parameter PAIR, see the kernel's comment): two CTAs of a
// cluster issue ONE tcgen05 cta_group::2 MMA of 256 queries x 256 targets; each
// keeps its own query operand and accumulators and loads only half of every
// target k-block (16 KB ring slots).  With 2-byte storage the epilogue of that
// form hands levels 0 and 1 to TMA tensor stores out of double-buffered staging
// (TMAST), so the output stream drains while the next rows are converted.  The
// single-CTA form above remains behind RC_BUILD_PAIR=0 and must give the same bits.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace {

constexpr int BM = 128;          // queries per tile (UMMA M)
constexpr int BN = TB_N;         // 256 targets per tile (UMMA N)
constexpr int BK = 64;           // bf16 elements per k-block (128 bytes = swizzle span); kind::tf32: 32 fp32
constexpr int UMMA_K = 16;       // bf16 elements per MMA (32 bytes); kind::tf32: 8 fp32, the same 32 bytes
constexpr int A_KB_BYTES = BM * 128;   // 16 KB
constexpr int B_ST_BYTES = BN * 128;   // 32 KB
constexpr int MAX_STAGES = 8;
constexpr int EPI_WARP0 = 2;
constexpr int EPI_WARPS = 8;           // two groups of four
constexpr int TC_THREADS = 32 * (EPI_WARP0 + EPI_WARPS);
constexpr int GW = TB_W / 2;           // block columns per epilogue group (16)

struct TcParams {
  int B, N, MT, NT, nbx, KB, num_levels, stages;
  int MTW;                          // query-block units the tile walk counts per pair: MT, or ceil(MT / 2) for CTA pairs
  int group_pairs, num_groups;      // pairs processed together (see TileWalk)
  int debug;                        // attribution switches; only read by builds with -DRC_TC_ATTRIB (tools/power_probe.py)
  int b_nhwc;                       // fmap2 operand is plain NHWC [B][H][W][D]: its 8x32 blocks are cut by a 4-D TMA box
  uint32_t idesc;                   // instruction descriptor (operand format: bf16, f16 or tf32)
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
// L2 next to the output stream.  Divisions happen once per range.
struct TileWalk {
  int grp, left;              // current group, tiles left in this CTA's range of it
  int b, m, n, by, bx;        // the next tile; (by, bx) = n's block coordinates
  int wid, nw;                // this walker and the number of walkers (CTAs, or CTA pairs: both CTAs of a pair walk alike)
  __device__ __forceinline__ void init(int walker, int walkers) {
    grp = -1; left = 0; b = m = n = by = bx = 0;
    wid = walker; nw = walkers;
  }
  __device__ __forceinline__ bool next(const TcParams &p, int &ob, int &om, int &on, int &oby, int &obx) {
    while (left == 0) {
      ++grp;
      if (grp >= p.num_groups) return false;
      const int pair0 = grp * p.group_pairs;
      const int np = min(p.group_pairs, p.B - pair0);
      const int tpp = p.MTW * p.NT;
      const long long tot = (long long)np * tpp;
      const int t0 = (int)(tot * wid / nw);
      const int t1 = (int)(tot * (wid + 1) / nw);
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
      if (++m == p.MTW) { m = 0; ++b; }
    }
    return true;
  }
  // after next(): does the following tile of this CTA use another A operand (or is there none)?
  __device__ __forceinline__ bool a_changes(int cb, int cm) const { return left == 0 || b != cb || m != cm; }
};

// PTX wrappers (mbarrier, TMA, tcgen05, L2 policies): tc_ptx.cuh
using namespace tc;

// Attribution builds (-DRC_TC_ATTRIB, RC_TC_DEBUG & 32): cycles a role spends in each of its waits, printed per CTA at the
// end of the kernel.  The shipped library compiles RC_TIMED(x, stmt) to stmt.
#ifdef RC_TC_ATTRIB
#define RC_TIMED(acc, stmt) do { const long long t0_ = clock64(); stmt; (acc) += clock64() - t0_; } while (0)
#else
#define RC_TIMED(acc, stmt) do { stmt; } while (0)
#endif

// Instruction descriptor: D = F32 (bit 4), operand format of A and B at bits 7 and 10 (kind::f16: 0 = IEEE half,
// 1 = bf16; kind::tf32: 2 = tf32), both K-major, N >> 3 at bit 17, M >> 4 at bit 24 (M = 256 for a CTA pair).
constexpr uint32_t make_idesc(uint32_t fmt, int M) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---------------------------------------------------------------- epilogue staging
// An epilogue thread owns one query (TMEM lane) and an 8 x 16 sub-block of the
// tile's targets.  Level 0 leaves in STEPS of 128 bytes per query (bf16: 4 block
// rows = 2 tiles, contiguous; fp32: 2 block rows = half of 4 tiles, 4 pieces of
// 32 bytes): the thread parks its 16-byte units in a per-warp buffer, then the
// warp reads the buffer back query by query and issues coalesced 32-byte stores
// (st.global.v8: 4 lanes cover one query's 128 bytes, every lane a full sector).
// The level-1 cells of the sub-block (64 / 128 bytes per query) take the same
// route once per tile.  Levels 2 and 3: with fp32 storage the owning thread
// stores its 16 / 8 bytes directly; with 2-byte storage a thread's share is only
// 8 / 4 bytes, so the two column groups merge theirs through shared memory and
// one of them stores whole sectors (end of the epilogue loop).
template <typename VolT>
struct Stg;
struct Stg16 {                            // 2-byte elements (bf16, fp16)
  static constexpr int TW = 8;            // tile width, elements
  static constexpr int L0_GT = 2;         // level-0 tiles per group and row of tiles
  static constexpr int L1_GT = 1;         // level-1 tiles per group
};
template <>
struct Stg<__nv_bfloat16> : Stg16 {};
template <>
struct Stg<__half> : Stg16 {};
template <>
struct Stg<float> {
  static constexpr int TW = 4;
  static constexpr int L0_GT = 4;
  static constexpr int L1_GT = 2;
};
constexpr int STEP_BYTES = 32 * 128;      // per-warp level-0 step buffer
constexpr int L1BUF_BYTES = 32 * 128;     // per-warp level-1 buffer (bf16 uses half)
constexpr int STAGE_BYTES_PER_WARP = STEP_BYTES + L1BUF_BYTES;
// The same for the epilogue that stores with TMA (CTA pairs, 2-byte storage): two level-0 step buffers (a step's bulk
// store drains while the next step is computed) and one level-1 buffer per warp, then 1 KB per TMEM quarter for the
// level-2/3 exchange behind all warps' buffers.  Every KB saved here is ring depth: the operand loads share L2 and the
// TMA unit with the output stream (three stages: the epilogue waited 32 % of its time for an accumulator, four: 22 %).
constexpr int TS_L1_OFF = 2 * STEP_BYTES, TS_L1_BYTES = 2048;
constexpr int TS_BYTES_PER_WARP = TS_L1_OFF + TS_L1_BYTES;  // 10 KB
constexpr int TS_XQ_OFF = EPI_WARPS * TS_BYTES_PER_WARP, TS_XQ_QUARTER = 32 * 16 + 32 * 4;   // per lane 16 B + 4 B
constexpr int TS_BYTES = TS_XQ_OFF + 4 * TS_XQ_QUARTER;      // 82.5 KB: with the 64 KB operand, FIVE 16 KB ring stages fit

__device__ __forceinline__ void sts16(uint32_t addr, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
}
__device__ __forceinline__ uint4 lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
// one 16-byte unit (8 bf16 or 4 fp32 values) -> shared memory
template <typename VolT>
__device__ __forceinline__ void put_unit(uint32_t addr, const float *v) {
  if (sizeof(VolT) == 2)
    sts16(addr, pack2<VolT>(v[0], v[1]), pack2<VolT>(v[2], v[3]), pack2<VolT>(v[4], v[5]), pack2<VolT>(v[6], v[7]));
  else
    sts16(addr, __float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3]));
}
// Buffers with 128-byte query rows: unit u of query q sits at q*128 + ((u ^ (q & 7)) << 4),
// so both the query-major writes and the unit-major read-back are conflict-free.  With a
// 128-byte-aligned row the XOR folds into the address: (row | rot) ^ (u << 4).
// Read-back of such a buffer and coalesced store, one full 32-byte sector per lane (st.global.v8): 4 lanes cover one
// query's 128 bytes, 8 queries per instruction.  `sbuf`: the buffer's base; lane = (query within 8) * 4 + unit pair;
// g: this lane's global address for its first query.  (With 16-byte stores the kernel was 4 % slower.)
// an empty volatile statement that "rewrites" 16 registers: later uses cannot be scheduled above it
__device__ __forceinline__ void pin_regs(uint32_t (&r)[16]) {
  asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                    "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]));
}
__device__ __forceinline__ void named_bar_arrive(uint32_t id, uint32_t count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
// one full 32-byte sector per lane
__device__ __forceinline__ void stg32_hint(void *dst, const uint4 &a, const uint4 &b, uint64_t policy) {
  asm volatile("st.global.L2::cache_hint.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8}, %9;" ::"l"(dst), "r"(a.x), "r"(a.y),
               "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w), "l"(policy)
               : "memory");
}

template <bool FULL>
__device__ __forceinline__ void flush128_v8(uint32_t sbuf, int lane, char *g, long long stride8, int nq, bool lane_ok,
                                            uint64_t pol) {
  const int up = lane & 3, sub8 = lane >> 2;
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int q = 8 * it + sub8;
    const uint32_t row = sbuf + q * 128;
    const uint4 a = lds16(row + (((2 * up) ^ (q & 7)) << 4));
    const uint4 b = lds16(row + (((2 * up + 1) ^ (q & 7)) << 4));
    if (FULL || (lane_ok && q < nq)) stg32_hint(g, a, b, pol);
    g += stride8;
  }
}

// the same for 64-byte query rows (2-byte level 1): 2 lanes per query, 16 queries per instruction
template <bool FULL>
__device__ __forceinline__ void flush64_v8(uint32_t lbuf, int lane, char *g, long long stride16, int nq, uint64_t pol) {
  const int up = lane & 1, sub16 = lane >> 1;
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int q = 16 * it + sub16;
    const uint32_t row = lbuf + q * 64;
    const uint4 a = lds16(row + (((2 * up) ^ ((q >> 1) & 3)) << 4));
    const uint4 b = lds16(row + (((2 * up + 1) ^ ((q >> 1) & 3)) << 4));
    if (FULL || q < nq) stg32_hint(g, a, b, pol);
    g += stride16;
  }
}

// ---------------------------------------------------------------- the kernel
// TF32 = false: bf16 operands (kind::f16), the query operand stationary in shared memory.
// TF32 = true:  fp32 operands read as tf32 (kind::tf32, 32 channels per 128-byte k-block).  Twice the operand bytes
//               do not fit next to the epilogue staging as a stationary operand, so the 16 KB query k-block rides
//               in the ring slot with its target k-block (one expect_tx, two TMA loads per slot).
// PAIR = true:  two CTAs of a cluster (one TPC) work as a pair on ONE MMA of 256 queries x 256 targets
//               (tcgen05 cta_group::2): CTA r of the pair owns query block 2 m + r (its own A operand and its own
//               128 x 256 accumulator: the epilogue is the single-CTA one) and loads only HALF of every target k-block
//               (block rows 4 r .. 4 r + 3: 16 KB instead of 32 KB).  The operand stream out of L2, the larger part of
//               what this kernel moves through L2, is halved.  The leader (rank 0) issues the MMAs; loads of both
//               CTAs count on the leader's "full" barriers; the leader's commits are multicast to both CTAs' "empty" and
//               "accumulator full" barriers; both epilogues arrive on the leader's "accumulator empty" barrier.
template <typename VolT, bool SCALE, bool TF32, bool PAIR>
__global__ void __launch_bounds__(TC_THREADS, 1)
build_pyramid_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                        const __grid_constant__ CUtensorMap map_l0, const __grid_constant__ CUtensorMap map_l1,
                        const TcParams p) {
  using S = Stg<VolT>;
  constexpr int ES = (int)sizeof(VolT);
  constexpr bool TMAST = PAIR && ES == 2;            // levels 0 and 1 leave through TMA tensor stores (map_l0, map_l1)
  constexpr uint32_t WSTAGE = TMAST ? TS_BYTES_PER_WARP : STAGE_BYTES_PER_WARP;
  constexpr int BKE = TF32 ? BK / 2 : BK;                                   // operand elements per k-block
  constexpr uint32_t B_SLOT = PAIR ? B_ST_BYTES / 2 : B_ST_BYTES;           // target bytes per ring slot of this CTA
  constexpr uint32_t SLOT = TF32 ? A_KB_BYTES + B_SLOT : B_SLOT;            // bytes per ring slot
  constexpr uint32_t CTAS = PAIR ? 2 : 1;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve (operand offsets are multiples of 1024)
  // The TMA-store variant needs every byte (5 ring stages leave 256 bytes) and has no slack to align itself with: the
  // dynamic shared memory of a kernel without static shared memory starts 1024-aligned, as declared; checked, not assumed.
  if (TMAST && (smem_u32(smem_raw) & 1023u)) __trap();
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sA = smem_u32(smem);
  const uint32_t sB = TF32 ? sA : sA + p.KB * A_KB_BYTES;
  const uint32_t sStage = sB + p.stages * SLOT;
  const uint32_t sBar = sStage + (TMAST ? TS_BYTES : EPI_WARPS * STAGE_BYTES_PER_WARP);
  uint32_t *tmem_slot = (uint32_t *)(smem + (sBar - sA) + 240);
  // barrier table (8 bytes each)
  const uint32_t bar_a_full = sBar, bar_a_empty = sBar + 8;
  const uint32_t bar_b_full = sBar + 16;                        // [stages]
  const uint32_t bar_b_empty = sBar + 16 + 8 * MAX_STAGES;      // [stages]
  const uint32_t bar_t_full = sBar + 16 + 8 * 2 * MAX_STAGES;   // [2]
  const uint32_t bar_t_empty = bar_t_full + 16;                 // [2]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;          // 0 = the pair's leader
  const int walker = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int walkers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    mbar_init(bar_a_full, 1);
    mbar_init(bar_a_empty, 1);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(bar_b_full + 8 * s, 1);
      mbar_init(bar_b_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_t_full + 8 * s, 1);
      mbar_init(bar_t_empty + 8 * s, CTAS * EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
  }
  fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();      // the peer's barriers exist before anything is sent to them
  fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      const uint64_t pol_in = l2_policy_evict_last();      // target blocks: every query block of the pair re-reads them
      // the stationary query operand is read ONCE (by this CTA): it must not compete with the target blocks for L2
      const uint64_t pol_a = TF32 ? pol_in : l2_policy_evict_first();
      // where this CTA's loads report: its own "full" barriers, or the leader's
      const uint32_t full_a = PAIR ? mapa_shared(bar_a_full, 0) : bar_a_full;
      const uint32_t full_b = PAIR ? mapa_shared(bar_b_full, 0) : bar_b_full;
      TileWalk walk;
      walk.init(walker, walkers);
      int b, m, n, by, bx, cur_key = -1;
      uint32_t a_par = 0;                    // parity of the A-empty phase to wait for next
      bool a_loaded = false;
      uint32_t st = 0, ph = 0;               // B ring: stage, parity of its current round
      while (walk.next(p, b, m, n, by, bx)) {
        const int key = b * p.MTW + m;
        const int mq = (PAIR ? 2 * m + (int)rank : m) * BM;      // first query of this CTA's block (may lie past N: zeros)
        if (!TF32 && key != cur_key) {
          if (a_loaded) { mbar_wait(bar_a_empty, a_par); a_par ^= 1; }
          if (rank == 0) mbar_expect_tx(bar_a_full, CTAS * p.KB * A_KB_BYTES);
          for (int kb = 0; kb < p.KB; ++kb) {
            if (PAIR) tma_load_3d_2sm(sA + kb * A_KB_BYTES, &map_a, full_a, kb * BKE, mq, b, pol_a);
            else tma_load_3d(sA + kb * A_KB_BYTES, &map_a, bar_a_full, kb * BKE, mq, b, pol_a);
          }
          a_loaded = true;
          cur_key = key;
        }
        // The query operand of the NEXT query block is read by nobody else, so its load would come from DRAM while the
        // MMAs wait: three tiles ahead, ask L2 to fetch it (if this walker's range reaches that block).
        if (!TF32 && n == (p.NT > 3 ? p.NT - 3 : 0) && walk.left >= p.NT - n) {
          int nb = b, nm = m + 1;
          if (nm == p.MTW) { nm = 0; ++nb; }
          for (int kb = 0; kb < p.KB; ++kb) tma_prefetch_3d(&map_a, kb * BKE, (PAIR ? 2 * nm + (int)rank : nm) * BM, nb);
        }
        for (int kb = 0; kb < p.KB; ++kb) {
          mbar_wait(bar_b_empty + 8 * st, ph ^ 1);
          if (rank == 0) mbar_expect_tx(bar_b_full + 8 * st, CTAS * SLOT);
          const uint32_t dst = sB + st * SLOT;
          if (PAIR) {
            const uint32_t fb = full_b + 8 * st;
            if (TF32) tma_load_3d_2sm(dst, &map_a, fb, kb * BKE, mq, b, pol_in);
            if (!TF32 && p.b_nhwc)   // box (64 d, 32 x, 4 y): this CTA's four rows of the 8x32 block
              tma_load_4d_2sm(dst, &map_b, fb, kb * BKE, bx * TB_W, by * TB_H + (int)rank * (TB_H / 2), b, pol_in);
            else
              tma_load_3d_2sm(dst + (TF32 ? A_KB_BYTES : 0), &map_b, fb, kb * BKE, n * BN + (int)rank * (BN / 2), b, pol_in);
          } else {
            if (TF32) tma_load_3d(dst, &map_a, bar_b_full + 8 * st, kb * BKE, mq, b, pol_in);
            if (!TF32 && p.b_nhwc)     // box (64 d, 32 x, 8 y): lands in the block order the pack would have written; outside the map: zeros
              tma_load_4d_hint(dst, &map_b, bar_b_full + 8 * st, kb * BKE, bx * TB_W, by * TB_H, b, pol_in);
            else
              tma_load_3d(dst + (TF32 ? A_KB_BYTES : 0), &map_b, bar_b_full + 8 * st, kb * BKE, n * BN, b, pol_in);
          }
          if (++st == (uint32_t)p.stages) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (of a pair: the leader's)
    if (lane == 0 && rank == 0) {
      TileWalk walk;
      walk.init(walker, walkers);
      int b, m, n, by, bx, cur_key = -1;
      uint32_t a_par = 0;                    // parity of the A-full phase to wait for next
      uint32_t st = 0, ph = 0;
      uint32_t i = 0;                        // tile counter: accumulator i & 1, its use number i >> 1
      const uint32_t a_lo = desc_lo(sA), b_lo = desc_lo(sB);
#ifdef RC_TC_ATTRIB
      long long w_afull = 0, w_tempty = 0, w_bfull = 0;
      const long long t_start = clock64();
#endif
      while (walk.next(p, b, m, n, by, bx)) {
        const int key = b * p.MTW + m;
        if (!TF32 && key != cur_key) {
          RC_TIMED(w_afull, mbar_wait(bar_a_full, a_par));
          a_par ^= 1;
          cur_key = key;
        }
        const uint32_t acc = i & 1;
        RC_TIMED(w_tempty, mbar_wait(bar_t_empty + 8 * acc, ((i >> 1) & 1) ^ 1));
        ++i;
        fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < p.KB; ++kb) {
          RC_TIMED(w_bfull, mbar_wait(bar_b_full + 8 * st, ph));
          fence_after();
          const uint32_t al = TF32 ? b_lo + ((st * SLOT) >> 4) : a_lo + ((kb * A_KB_BYTES) >> 4);
          const uint32_t bl = b_lo + ((st * SLOT + (TF32 ? A_KB_BYTES : 0)) >> 4);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {      // four MMAs of 32 operand bytes per k-block, either kind
            const uint64_t ad = desc_make(al + k * ((UMMA_K * 2) >> 4)), bd = desc_make(bl + k * ((UMMA_K * 2) >> 4));
            const uint32_t accum = (kb | k) != 0;
            if (PAIR) {
              if (TF32) umma_tf32_2sm(d_tmem, ad, bd, p.idesc, accum);
              else umma_bf16_2sm(d_tmem, ad, bd, p.idesc, accum);
            } else {
              if (TF32) umma_tf32(d_tmem, ad, bd, p.idesc, accum);
              else umma_bf16(d_tmem, ad, bd, p.idesc, accum);      // bf16 or IEEE-half operands: the descriptor's format fields
            }
          }
          if (PAIR) umma_commit_2sm(bar_b_empty + 8 * st);   // frees the stage (in both CTAs) when these MMAs retire
          else umma_commit(bar_b_empty + 8 * st);
          if (++st == (uint32_t)p.stages) { st = 0; ph ^= 1; }
        }
        if (PAIR) umma_commit_2sm(bar_t_full + 8 * acc);     // accumulator ready
        else umma_commit(bar_t_full + 8 * acc);
        if (!TF32 && walk.a_changes(b, m)) {                 // last use of this A operand
          if (PAIR) umma_commit_2sm(bar_a_empty);
          else umma_commit(bar_a_empty);
        }
      }
#ifdef RC_TC_ATTRIB
      if ((p.debug & 32) && blockIdx.x < 8)
        printf("mma cta %d: tiles %u total %lld wait a_full %lld t_empty %lld b_full %lld\n", blockIdx.x, i,
               clock64() - t_start, w_afull, w_tempty, w_bfull);
#endif
    }
  } else {
    // ===================================================== epilogue (2 groups x 4 warps)
    const int quarter = warp & 3;                 // TMEM lanes [32*quarter, +32)
    const int group = (warp - EPI_WARP0) >> 2;    // block columns 16*group .. +15
    const uint32_t sbuf = sStage + (warp - EPI_WARP0) * WSTAGE;                 // level-0 step buffer(s)
    const uint32_t lbuf = sbuf + (TMAST ? TS_L1_OFF : STEP_BYTES);              // level-1 buffer(s)
    const int L = p.num_levels;
    // per-lane constants of the staging buffers (see flush128_v8 / flush64_v8)
    const uint32_t st_base = (sbuf + lane * 128) | ((lane & 7) << 4);           // write: unit u at st_base ^ (u << 4)
    // level 1: bf16 = 4 units (one tile) per query, 64-byte rows; fp32 = 8 units (two tiles), 128-byte rows
    const uint32_t l1_st = ES == 2 ? ((lbuf + lane * 64) | (((lane >> 1) & 3) << 4)) : ((lbuf + lane * 128) | ((lane & 7) << 4));
    const uint64_t pol_out = l2_policy_evict_first();

    const uint32_t t_empty = PAIR ? mapa_shared(bar_t_empty, 0) : bar_t_empty;   // the MMA issuer's barrier (pair: the leader's)

    TileWalk walk;
    walk.init(walker, walkers);
    int b, m, n, by, bx;
    uint32_t i = 0;
#ifdef RC_TC_ATTRIB
    long long w_tfull = 0, w_store = 0;
    const long long e_start = clock64();
#endif
    while (walk.next(p, b, m, n, by, bx)) {
      const uint32_t acc = i & 1, t_par = (i >> 1) & 1;
      const bool first_tile = i == 0;
      ++i;
      const int x0 = bx * TB_W, y0 = by * TB_H;
      const int xg = x0 + group * GW;                  // first map column of this group's sub-block
      const int qw = (PAIR ? 2 * m + (int)rank : m) * BM + quarter * 32;   // first query of this warp (past N: nothing is stored)
      const long long slab_q0 = (long long)b * p.N + qw;    // slab index of this warp's first query
      const int nq = min(32, p.N - qw);                     // valid queries of this warp (may be <= 0)

      // valid pooled cells of this sub-block (floor rule of avg_pool2d; the rest is padding -> 0)
      const int nvx1 = p.Wl[1] - (xg >> 1), nvy1 = p.Hl[1] - (y0 >> 1);
      const int nvx2 = p.Wl[2] - (xg >> 2), nvy2 = p.Hl[2] - (y0 >> 2);
      const int nvx3 = p.Wl[3] - (xg >> 3), nvy3 = p.Hl[3] - (y0 >> 3);
      const bool edge = (L > 1 && (nvx1 < 8 || nvy1 < 4)) || (L > 2 && (nvx2 < 4 || nvy2 < 2)) ||
                        (L > 3 && (nvx3 < 2 || nvy3 < 1));
      // level-0 chunk of this group: tiles txg .. txg + L0_GT - 1 of tile rows (y0 >> 2), (y0 >> 2) + 1
      const int txg = xg / S::TW;
      const int nt0 = min(S::L0_GT, p.tiles_x[0] - txg);
      const bool full0 = nq == 32 && nt0 == S::L0_GT;
      const long long slab0B = p.slab[0] * ES;
      // level-0 flush: lane = (query within 8) * 4 + pair of units; a pair is 32 contiguous bytes of one tile
      const int up4 = lane & 3;
      char *g0 = (char *)p.lvl[0] + ((long long)(y0 >> 2) * p.tiles_x[0] + txg) * 64 + (slab_q0 + (lane >> 2)) * slab0B +
                 (ES == 2 ? up4 * 32 : up4 * 64);
      const bool pair_ok = (ES == 2 ? (up4 >> 1) : up4) < nt0;
      const long long tile_row0B = (long long)p.tiles_x[0] * 64;

      // ---- accumulator -> registers, then hand it back at once
      RC_TIMED(w_tfull, mbar_wait(bar_t_full + 8 * acc, t_par));
      fence_after();
      const uint32_t taddr = tmem_base + acc * BN + group * GW + ((uint32_t)(quarter * 32) << 16);
      // TMAST: block rows 0..3 now, rows 4..7 after the first pair of rows has been consumed (96 instead of 128
      // accumulator registers live at the peak: with all 128 the kernel spilled, and a spill reload sat right behind
      // this wait); the accumulator is handed back when the second half has landed.  Otherwise all 8 rows at once.
      uint32_t raw[8][GW];
#pragma unroll
      for (int r = 0; r < (TMAST ? 4 : 8); ++r) tmem_ld16(taddr + r * TB_W, raw[r]);
      tmem_ld_wait();
      if (!TMAST) {
        fence_before();
        __syncwarp();
        if (lane == 0) {
          if (PAIR) mbar_arrive_cluster(t_empty + 8 * acc);
          else mbar_arrive(bar_t_empty + 8 * acc);
        }
      }
      // Attribution builds only (python -m raft_stir_b200.build --attrib -> _build/attrib/libraftcorr.so, RC_TC_DEBUG in
      // the environment): 1 = no global stores of levels 0/1 and no level-2/3 stores, 2 = the same, 4 = no staging
      // writes either, 8 = the epilogue stops after its TMEM reads.  Results are wrong by construction; the shipped
      // library compiles with dbg = 0 and none of this.
#ifdef RC_TC_ATTRIB
      const int dbg = p.debug;
      if (dbg & 8) {
        if (TMAST) {
#pragma unroll
          for (int r = 4; r < 8; ++r) tmem_ld16(taddr + r * TB_W, raw[r]);
          tmem_ld_wait();
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(t_empty + 8 * acc);
        }
        asm volatile("" ::"r"(raw[0][0]), "r"(raw[7][15]));
        continue;
      }
#else
      constexpr int dbg = 0;
#endif

      float p1prev[8], p2prev[4];
      uint32_t l1keep[2][4];                            // TMA-store epilogue: level-1 rows 0 and 1, packed
      uint32_t l2w[4] = {0u, 0u, 0u, 0u}, l3w = 0u;     // 2-byte storage: this thread's level-2 rows and level-3 cells, packed
#pragma unroll
      for (int rp = 0; rp < 4; ++rp) {
        float v0[GW], v1[GW], p1[8];
#pragma unroll
        for (int k = 0; k < GW; ++k) {
          v0[k] = __uint_as_float(raw[2 * rp][k]);
          v1[k] = __uint_as_float(raw[2 * rp + 1][k]);
          if (SCALE) { v0[k] *= p.scale; v1[k] *= p.scale; }
        }
        // ---- level 0 units of block rows 2rp, 2rp+1
        if (TMAST && rp == 1) {
#pragma unroll
          for (int r = 4; r < 8; ++r) tmem_ld16(taddr + r * TB_W, raw[r]);      // in flight under this pair of rows
        }
        if (TMAST && rp == 2) {
          tmem_ld_wait();
          pin_regs(raw[4]); pin_regs(raw[5]); pin_regs(raw[6]); pin_regs(raw[7]);   // no use of these is scheduled above the wait
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(t_empty + 8 * acc);
        }
        if (TMAST && !(rp & 1)) {
          // the bulk store that last read this step buffer (and, at rp == 2, the level-1 buffer: the previous tile's
          // last group) has to be done with shared memory: all but the newest group
          if (lane == 0) RC_TIMED(w_store, bulk_wait_read<1>());
          __syncwarp();
        }
        if (ES == 2 && !(dbg & 4)) {
          const int r = (2 * rp) & 3;                    // row inside the tile
          const uint32_t sb = TMAST ? st_base + (rp >> 1) * STEP_BYTES : st_base;
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            put_unit<VolT>(sb ^ ((k * 4 + r) << 4), v0 + 8 * k);
            put_unit<VolT>(sb ^ ((k * 4 + r + 1) << 4), v1 + 8 * k);
          }
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            put_unit<VolT>(st_base ^ ((k * 2) << 4), v0 + 4 * k);
            put_unit<VolT>(st_base ^ ((k * 2 + 1) << 4), v1 + 4 * k);
          }
        }
        // ---- pooled levels, from the fp32 values
#pragma unroll
        for (int k = 0; k < 8; ++k) p1[k] = 0.25f * ((v0[2 * k] + v0[2 * k + 1]) + (v1[2 * k] + v1[2 * k + 1]));
        if (edge) {
#pragma unroll
          for (int k = 0; k < 8; ++k) p1[k] = (k < nvx1 && rp < nvy1) ? p1[k] : 0.f;
        }
        if (L > 1) {
          // level-1 row rp of the sub-block: bf16 one unit (row rp of its tile), fp32 two (tiles 0, 1)
          if (TMAST) {
            // one level-1 buffer: the store of the previous tile's cells is known to be done with it only after the wait
            // at rp == 2, so rows 0 and 1 stay in registers (packed) until then
            if (rp < 2) {
#pragma unroll
              for (int k = 0; k < 4; ++k) l1keep[rp][k] = pack2<VolT>(p1[2 * k], p1[2 * k + 1]);
            } else {
              if (rp == 2) {
                sts16(l1_st, l1keep[0][0], l1keep[0][1], l1keep[0][2], l1keep[0][3]);
                sts16(l1_st ^ (1 << 4), l1keep[1][0], l1keep[1][1], l1keep[1][2], l1keep[1][3]);
              }
              put_unit<VolT>(l1_st ^ (rp << 4), p1);
            }
          } else if (ES == 2) {
            put_unit<VolT>(l1_st ^ (rp << 4), p1);
          } else {
            put_unit<VolT>(l1_st ^ (rp << 4), p1);
            put_unit<VolT>(l1_st ^ ((4 + rp) << 4), p1 + 4);
          }
        }
        if (rp & 1) {
          // ---- level-2 row rp >> 1: 4 cells, straight from this thread's registers
          float p2[4];
#pragma unroll
          for (int k = 0; k < 4; ++k)
            p2[k] = 0.25f * ((p1prev[2 * k] + p1prev[2 * k + 1]) + (p1[2 * k] + p1[2 * k + 1]));
          if (edge) {
#pragma unroll
            for (int k = 0; k < 4; ++k) p2[k] = (k < nvx2 && (rp >> 1) < nvy2) ? p2[k] : 0.f;
          }
          const int x2 = xg >> 2;                        // first level-2 column of the sub-block
          if (L > 2 && lane < nq && (y0 >> 4) < p.tiles_y[2] && x2 / S::TW < p.tiles_x[2]) {
            char *g2 = (char *)p.lvl[2] + ((long long)(y0 >> 4) * p.tiles_x[2] + x2 / S::TW) * 64 +
                       (slab_q0 + lane) * p.slab[2] * ES + (((y0 >> 2) & 3) + (rp >> 1)) * 16 + (x2 % S::TW) * ES;
            if (ES == 4)
              stg16_hint(g2, make_uint4(__float_as_uint(p2[0]), __float_as_uint(p2[1]), __float_as_uint(p2[2]),
                                        __float_as_uint(p2[3])), pol_out);
          }
          if (ES == 2) {          // stored after the tile, merged with the other group's half (see below)
            l2w[(rp >> 1) * 2] = pack2<VolT>(p2[0], p2[1]);
            l2w[(rp >> 1) * 2 + 1] = pack2<VolT>(p2[2], p2[3]);
          }
          if (rp == 3) {
            // ---- level 3: 2 cells
            float p3[2];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const float a = 0.25f * ((p2prev[2 * k] + p2prev[2 * k + 1]) + (p2[2 * k] + p2[2 * k + 1]));
              p3[k] = (!edge || (k < nvx3 && 0 < nvy3)) ? a : 0.f;
            }
            const int x3 = xg >> 3;
            if (L > 3 && lane < nq && (y0 >> 5) < p.tiles_y[3] && x3 / S::TW < p.tiles_x[3]) {
              char *g3 = (char *)p.lvl[3] + ((long long)(y0 >> 5) * p.tiles_x[3] + x3 / S::TW) * 64 +
                         (slab_q0 + lane) * p.slab[3] * ES + ((y0 >> 3) & 3) * 16 + (x3 % S::TW) * ES;
              if (ES == 4) stg8_hint(g3, __float_as_uint(p3[0]), __float_as_uint(p3[1]), pol_out);
            }
            if (ES == 2) l3w = pack2<VolT>(p3[0], p3[1]);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) p2prev[k] = p2[k];
          }
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) p1prev[k] = p1[k];
        }

        // ---- a level-0 step is complete: read it back query by query, coalesced stores
        if (TMAST) {
          if (rp & 1) {
            // this step's rows of tiles [32 queries][2 tiles x 64 B] sit in the step buffer in the 128-byte swizzle:
            // one tensor store (box clipped at the end of the map's row of tiles, its last tile row and query), asynchronous
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              if (!(dbg & 3)) tma_store_4d(&map_l0, sbuf + (rp >> 1) * STEP_BYTES, txg * 32, (y0 >> 2) + (rp >> 1), qw, b, pol_out);
              if (rp == 3 && L > 1 && !(dbg & 3))     // and the tile's level-1 cells: [32 queries][64 B], 64-byte swizzle
                tma_store_4d(&map_l1, lbuf, ((xg >> 1) / S::TW) * 32, y0 >> 3, qw, b, pol_out);
              bulk_commit();
            }
          }
        } else if (ES == 4 || (rp & 1)) {
          __syncwarp();
          // bf16: step = one row of tiles; fp32: step = two of the four rows of a row of tiles
          if ((y0 >> 2) + (rp >> 1) < p.tiles_y[0] && nt0 > 0 && !(dbg & 2)) {
            char *g = g0 + (rp >> 1) * tile_row0B + (ES == 2 ? 0 : (rp & 1) * 32);
            if (dbg & 1) flush128_v8<false>(sbuf, lane, g, 8 * slab0B, 0, pair_ok, pol_out);
            else if (full0) flush128_v8<true>(sbuf, lane, g, 8 * slab0B, nq, true, pol_out);
            else flush128_v8<false>(sbuf, lane, g, 8 * slab0B, nq, pair_ok, pol_out);
          }
          __syncwarp();
        }
      }

      // ---- level-1 cells of the sub-block
      if (!TMAST && L > 1) {
        const int tx1 = (xg >> 1) / S::TW;
        const int nt1 = min(S::L1_GT, p.tiles_x[1] - tx1);
        if ((y0 >> 3) < p.tiles_y[1] && nt1 > 0 && !(dbg & 3)) {
          const long long slab1B = p.slab[1] * ES;
          char *g1 = (char *)p.lvl[1] + ((long long)(y0 >> 3) * p.tiles_x[1] + tx1) * 64;
          if (ES == 2) {
            g1 += (slab_q0 + (lane >> 1)) * slab1B + (lane & 1) * 32;
            if (nq == 32) flush64_v8<true>(lbuf, lane, g1, 16 * slab1B, nq, pol_out);
            else flush64_v8<false>(lbuf, lane, g1, 16 * slab1B, nq, pol_out);
          } else {
            g1 += (slab_q0 + (lane >> 2)) * slab1B + (lane & 3) * 32;     // pair of units = rows 2k, 2k+1 of tile (lane & 3) >> 1
            if (nq == 32 && nt1 == 2) flush128_v8<true>(lbuf, lane, g1, 8 * slab1B, nq, true, pol_out);
            else flush128_v8<false>(lbuf, lane, g1, 8 * slab1B, nq, ((lane & 3) >> 1) < nt1, pol_out);
          }
        }
        __syncwarp();
      }

      // ---- levels 2 and 3 with 2-byte storage.  A block's level-2 cells are ONE 32-byte sector per query (2 rows of
      // 8 cells), its level-3 cells 8 bytes; each column group holds half of either.  Written separately they are
      // four 8-byte and two 4-byte scattered stores per query, each occupying a whole sector on the way to L2 (a quarter
      // more write traffic than the kernel has output).  Group 1 hands its halves to group 0 through the unused half
      // of its level-1 buffer ("full" = named barrier 1 + quarter, "empty" = 5 + quarter, 64 threads each: the
      // arriving warp and the waiting one); group 0 stores whole sectors.
      if (ES == 2 && L > 2 && !(dbg & 3)) {
        const uint32_t xq = TMAST ? sStage + TS_XQ_OFF + quarter * TS_XQ_QUARTER + lane * 16      // behind the warps' buffers
                                  : sStage + (((warp - EPI_WARP0) & 3) + 4) * WSTAGE + STEP_BYTES + 2048 + lane * 32;   // in the group-1 warp's buffer
        const uint32_t xq3 = TMAST ? sStage + TS_XQ_OFF + quarter * TS_XQ_QUARTER + 512 + lane * 4 : xq + 16;     // its level-3 word
        if (group == 1) {
          if (!first_tile) named_bar_sync(5 + quarter, 64);
          sts16(xq, l2w[0], l2w[1], l2w[2], l2w[3]);
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(xq3), "r"(l3w) : "memory");
          named_bar_arrive(1 + quarter, 64);
        } else {
          named_bar_sync(1 + quarter, 64);
          const uint4 o = lds16(xq);
          uint32_t o3;
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(o3) : "r"(xq3) : "memory");
          named_bar_arrive(5 + quarter, 64);
          const int x2 = x0 >> 2, x3 = x0 >> 3;
          if (lane < nq && (y0 >> 4) < p.tiles_y[2] && x2 / S::TW < p.tiles_x[2]) {
            char *g2 = (char *)p.lvl[2] + ((long long)(y0 >> 4) * p.tiles_x[2] + x2 / S::TW) * 64 +
                       (slab_q0 + lane) * p.slab[2] * ES + ((y0 >> 2) & 3) * 16;
            stg32_hint(g2, make_uint4(l2w[0], l2w[1], o.x, o.y), make_uint4(l2w[2], l2w[3], o.z, o.w), pol_out);
          }
          if (L > 3 && lane < nq && (y0 >> 5) < p.tiles_y[3] && x3 / S::TW < p.tiles_x[3]) {
            char *g3 = (char *)p.lvl[3] + ((long long)(y0 >> 5) * p.tiles_x[3] + x3 / S::TW) * 64 +
                       (slab_q0 + lane) * p.slab[3] * ES + ((y0 >> 3) & 3) * 16 + (x3 % S::TW) * ES;
            // a tile row holds the level-3 cells of two blocks; with no block to the right, its cells 4..7 are
            // padding columns and must read as zero
            if (!(bx & 1) && bx + 1 >= p.nbx) stg16_hint(g3, make_uint4(l3w, o3, 0u, 0u), pol_out);
            else stg8_hint(g3, l3w, o3, pol_out);
          }
        }
      }
    }
    if (TMAST) {
      if (lane == 0) bulk_wait_all();      // shared memory stays until the last stores have read it
      __syncwarp();
    }
#ifdef RC_TC_ATTRIB
    if ((p.debug & 32) && blockIdx.x < 8 && lane == 0 && (warp == 2 || warp == 7))
      printf("epi cta %d warp %d: tiles %u total %lld wait t_full %lld store-buffer %lld\n", blockIdx.x, warp, i,
             clock64() - e_start, w_tfull, w_store);
#endif
  }

  fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();      // neither CTA leaves while the other may still signal its barriers or use its operands
  fence_after();
  if (warp == 1) {
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
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

// RC_BUILD_PAIR=0 in the environment selects the single-CTA kernel (kept for A/B timing and for devices or contexts that
// cannot co-schedule two CTAs of a cluster); read once.
bool pairs_enabled() {
  static const bool on = [] {
    const char *e = getenv("RC_BUILD_PAIR");
    return !(e && e[0] == '0');
  }();
  return on;
}

template <typename VolT, bool TF32, bool PAIR>
int build_tc_t(const void *f1p, const void *f2p, int B, int D, int H, int W, const rc_layout &lay,
               VolT *pyr, bool half_operands, bool nhwc, cudaStream_t s) {
  constexpr int BKE = TF32 ? BK / 2 : BK;
  constexpr int OPB = TF32 ? 4 : 2;                 // operand bytes per element
  constexpr size_t B_SLOT = PAIR ? B_ST_BYTES / 2 : B_ST_BYTES;
  constexpr size_t SLOT = TF32 ? A_KB_BYTES + B_SLOT : B_SLOT;
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
  p.MTW = PAIR ? ceil_div(p.MT, 2) : p.MT;
  p.nbx = ceil_div(W, TB_W);
  p.NT = p.nbx * ceil_div(H, TB_H);
  p.KB = ceil_div(D, BKE);
  p.idesc = make_idesc(TF32 ? 2u : half_operands ? 0u : 1u, PAIR ? 2 * BM : BM);
  const CUtensorMapDataType op_type = TF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                      : half_operands ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  p.num_levels = L;
  float pre_scale = 1.f;
  split_scale(D, &pre_scale, &p.scale);      // pre_scale was applied to fmap1 by the pack
  if (nhwc) p.scale = 1.0f / sqrtf((float)D);   // caller's features: unscaled, the whole factor goes to the epilogue
  p.b_nhwc = nhwc ? 1 : 0;
#ifdef RC_TC_ATTRIB
  { const char *e = getenv("RC_TC_DEBUG"); p.debug = e ? atoi(e) : 0; }
#endif
  if (nhwc && (TF32 || D % BKE != 0)) {
    set_error("NHWC feature input needs 16-bit operands and D %% 64 == 0 (got D = %d)", D);
    return RC_EINVAL;
  }
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
  constexpr bool TMAST = PAIR && sizeof(VolT) == 2;
  const size_t fixed = (TF32 ? 0 : (size_t)p.KB * A_KB_BYTES) +
                       (TMAST ? (size_t)TS_BYTES : (size_t)EPI_WARPS * STAGE_BYTES_PER_WARP) + 256 + (TMAST ? 0 : 1024);
  p.stages = 0;
  for (int st = MAX_STAGES; st >= 2; --st)
    if (fixed + (size_t)st * SLOT <= (size_t)max_smem) {
      p.stages = st;
      break;
    }
  if (p.stages == 0) {
    set_error("D = %d does not fit the tensor-core build's shared memory budget", D);
    return RC_EINVAL;
  }
  const size_t smem = fixed + (size_t)p.stages * SLOT;

  auto kern = p.scale == 1.f ? build_pyramid_tc_kernel<VolT, false, TF32, PAIR> : build_pyramid_tc_kernel<VolT, true, TF32, PAIR>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

  // walkers of the tile list: CTAs, or CTA pairs (as many as the device co-schedules, at most one per two SMs)
  int walkers = sms;
  if (PAIR) {
    cudaLaunchConfig_t q{};
    q.gridDim = dim3((unsigned)(sms & ~1));
    q.blockDim = dim3(TC_THREADS);
    q.dynamicSmemBytes = smem;
    cudaLaunchAttribute qa{};
    qa.id = cudaLaunchAttributeClusterDimension;
    qa.val.clusterDim.x = 2;
    qa.val.clusterDim.y = 1;
    qa.val.clusterDim.z = 1;
    q.attrs = &qa;
    q.numAttrs = 1;
    // asked once per kernel instantiation, device and shared-memory size (the query costs more than the tensor maps)
    static std::atomic<long long> cached{-1};                  // (device << 48) | (smem << 16) | clusters
    const long long key = ((long long)dev << 48) | ((long long)smem << 16);
    long long c = cached.load(std::memory_order_relaxed);
    int clusters = 0;
    if (c >= 0 && (c & ~0xffffLL) == key) {
      clusters = (int)(c & 0xffff);
    } else {
      RC_CUDA(cudaOccupancyMaxActiveClusters(&clusters, kern, &q));
      cached.store(key | (clusters & 0xffff), std::memory_order_relaxed);
    }
    walkers = clusters < sms / 2 ? clusters : sms / 2;
    if (walkers < 1) {
      set_error("the device cannot co-schedule a CTA pair of the tensor-core build (set RC_BUILD_PAIR=0)");
      return RC_ECUDA;
    }
  }

  // Pairs per group.  All walkers enter a group together and miss L2 on its operands together, so groups should be
  // few; but a group's operands have to stay in L2 next to the output stream.  Measured at the Sintel shape, 64 pairs
  // (7.2 MB of 16-bit operands per pair, `tools/gpu_group.sh`): 3 pairs per group 2.37 ms, 4: 2.24-2.39, 8: 2.26, 11 or
  // more: 2.35 with twice the DRAM reads (79 MB no longer stay).  So: as many pairs as fit 60 MB, and at least about
  // one query block's worth of tiles per walker.
  const long long tpp = (long long)p.MTW * p.NT;
  long long gp = ((long long)p.NT * walkers + tpp - 1) / tpp;
  const long long pair_operand_bytes = ((long long)N + rows2) * D * OPB;
  if (TF32) gp = (gp + 1) / 2;            // fp32 operands: twice the bytes to keep hot in L2 per pair
  // (with 4-byte storage the output stream is twice as wide and the larger groups were slower: 5.71 against 5.49 ms)
  if (TMAST && 60000000LL / pair_operand_bytes > gp) gp = 60000000LL / pair_operand_bytes;
#ifdef RC_TC_ATTRIB
  { const char *e = getenv("RC_TC_GROUP"); if (e) gp = atoi(e); }
#endif
  if (gp < 1) gp = 1;
  if (gp > B) gp = B;
  p.group_pairs = (int)gp;
  p.num_groups = ceil_div(B, p.group_pairs);

  CUtensorMap ma, mb;
  {
    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t str[2] = {(cuuint64_t)D * OPB, (cuuint64_t)N * D * OPB};
    cuuint32_t box[3] = {BKE, BM, 1};
    int r = encode(&ma, op_type, 3, f1p, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B, "fmap1");
    if (r) return r;
  }
  // a CTA of a pair loads half of a target block: block rows 4 r .. 4 r + 3 = 128 consecutive packed rows
  if (nhwc) {
    cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t str[3] = {(cuuint64_t)D * OPB, (cuuint64_t)W * D * OPB, (cuuint64_t)H * W * D * OPB};
    cuuint32_t box[4] = {BKE, TB_W, PAIR ? TB_H / 2 : TB_H, 1};
    int r = encode(&mb, op_type, 4, f2p, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B, "fmap2 (NHWC)");
    if (r) return r;
  } else {
    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)rows2, (cuuint64_t)B};
    cuuint64_t str[2] = {(cuuint64_t)D * OPB, (cuuint64_t)rows2 * D * OPB};
    cuuint32_t box[3] = {BKE, PAIR ? BN / 2 : BN, 1};
    int r = encode(&mb, op_type, 3, f2p, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B, "fmap2");
    if (r) return r;
  }

  // TMA stores of levels 0 and 1: a level as [B][N][tiles_y][tiles_x * 32 elements] (a row of tiles is contiguous, 64 bytes
  // per tile); the box is one warp's 32 queries x its tiles of one tile row (2 at level 0 = 128 bytes, 1 at level 1),
  // clipped at the level's edges.  The box's inner extent equals the swizzle span (a 64-byte inner extent under the
  // 128-byte swizzle is laid out differently in shared memory).
  CUtensorMap ml0 = ma, ml1 = ma;
  if (TMAST) {
    for (int l = 0; l < (L > 1 ? 2 : 1); ++l) {
      cuuint64_t dims[4] = {(cuuint64_t)lay.tiles_x[l] * 32, (cuuint64_t)lay.tiles_y[l], (cuuint64_t)N, (cuuint64_t)B};
      cuuint64_t str[3] = {(cuuint64_t)lay.tiles_x[l] * 64, (cuuint64_t)lay.slab[l] * 2, (cuuint64_t)N * lay.slab[l] * 2};
      cuuint32_t box[4] = {l == 0 ? 64u : 32u, 1, 32, 1};
      int r = encode(l == 0 ? &ml0 : &ml1, CU_TENSOR_MAP_DATA_TYPE_UINT16, 4, p.lvl[l], dims, str, box,
                     l == 0 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, l == 0 ? "pyramid level 0" : "pyramid level 1");
      if (r) return r;
    }
  }

  const long long group_tiles = (long long)p.group_pairs * tpp;
  const int used = group_tiles < walkers ? (int)group_tiles : walkers;
  if (PAIR) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2u * (unsigned)used);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at{};
    at.id = cudaLaunchAttributeClusterDimension;
    at.val.clusterDim.x = 2;
    at.val.clusterDim.y = 1;
    at.val.clusterDim.z = 1;
    cfg.attrs = &at;
    cfg.numAttrs = 1;
    RC_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, ml0, ml1, p));
  } else {
    kern<<<used, TC_THREADS, smem, s>>>(ma, mb, ml0, ml1, p);
  }
  RC_LAUNCH_CHECK("build_pyramid_tc_kernel");
  return RC_OK;
}

template <typename VolT, bool TF32>
int build_tc_sel(const void *f1p, const void *f2p, int B, int D, int H, int W, const rc_layout &lay, VolT *pyr,
                 bool half_operands, bool nhwc, cudaStream_t s) {
  return pairs_enabled() ? build_tc_t<VolT, TF32, true>(f1p, f2p, B, D, H, W, lay, pyr, half_operands, nhwc, s)
                         : build_tc_t<VolT, TF32, false>(f1p, f2p, B, D, H, W, lay, pyr, half_operands, nhwc, s);
}

}  // namespace

// shared with build_bwd.cu (declared in tc_ptx.cuh)
int encode_tiled(CUtensorMap *map, CUtensorMapDataType dt, int rank, const void *base, const cuuint64_t *dims,
                 const cuuint64_t *strides_bytes, const cuuint32_t *box, CUtensorMapSwizzle sw, const char *what) {
  return encode(map, dt, rank, base, dims, strides_bytes, box, sw, what);
}

int launch_build_tc(const void *f1p, const void *f2p, int B, int D, int H, int W, const rc_layout &lay,
                    int vol_dtype, int math, int nhwc, void *pyramid, cudaStream_t s) {
  const bool nh = nhwc != 0;
  if (math == RC_MATH_TF32)
    return vol_dtype == RC_VOL_BF16  ? build_tc_sel<__nv_bfloat16, true>(f1p, f2p, B, D, H, W, lay, (__nv_bfloat16 *)pyramid, false, nh, s)
           : vol_dtype == RC_VOL_F16 ? build_tc_sel<__half, true>(f1p, f2p, B, D, H, W, lay, (__half *)pyramid, false, nh, s)
                                     : build_tc_sel<float, true>(f1p, f2p, B, D, H, W, lay, (float *)pyramid, false, nh, s);
  if (math != RC_MATH_BF16 && math != RC_MATH_F16) {
    set_error("tensor-core build: math=%d is not a tensor-core mode", math);
    return RC_EINVAL;
  }
  const bool half = math == RC_MATH_F16;
  return vol_dtype == RC_VOL_BF16  ? build_tc_sel<__nv_bfloat16, false>(f1p, f2p, B, D, H, W, lay, (__nv_bfloat16 *)pyramid, half, nh, s)
         : vol_dtype == RC_VOL_F16 ? build_tc_sel<__half, false>(f1p, f2p, B, D, H, W, lay, (__half *)pyramid, half, nh, s)
                                   : build_tc_sel<float, false>(f1p, f2p, B, D, H, W, lay, (float *)pyramid, half, nh, s);
}

}  // namespace rc
