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
  int N, MT, NT, nbx, KB, num_levels, stages;
  int total_tiles;
  float scale;
  // level 3 of a bf16 volume is 8-byte rows: below TMA's 16-byte minimum, stored directly
  void *l3_ptr;
  long long l3_slab;
  int l3_pitch, l3_H, l3_W;
  // level extents for skipping fully out-of-range stores
  int Hl[4], Wl[4];
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
__device__ __forceinline__ void tma_store_4d(const CUtensorMap *map, uint32_t src, int c0, int c1, int c2,
                                             int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
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
template <typename VolT>
struct Stg;
template <>
struct Stg<__nv_bfloat16> {
  static constexpr int ES = 2;
  static constexpr int OFF_L0 = 0, OFF_L1 = 4096, OFF_L2 = 5120, OFF_L3 = 5632, BYTES = 6144;
  static constexpr bool L3_TMA = false;
};
template <>
struct Stg<float> {
  static constexpr int ES = 4;
  static constexpr int OFF_L0 = 0, OFF_L1 = 8192, OFF_L2 = 10240, OFF_L3 = 11264, BYTES = 12288;
  static constexpr bool L3_TMA = true;
};

// Write NV fp32 values (one row segment of this thread's query) into the staging
// row `lane` of a [32 rows x RB bytes] box laid out with the TMA swizzle of span RB.
template <typename VolT, int NV>
__device__ __forceinline__ void stage_row(uint32_t base, int lane, const float *v) {
  constexpr int RB = NV * (int)sizeof(VolT);
  constexpr int CH = RB / 16;
  const uint32_t row = base + lane * RB;
  const int sw = ((lane * RB) >> 7) & (CH - 1);
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    uint32_t w0, w1, w2, w3;
    if (sizeof(VolT) == 4) {
      w0 = __float_as_uint(v[4 * j + 0]); w1 = __float_as_uint(v[4 * j + 1]);
      w2 = __float_as_uint(v[4 * j + 2]); w3 = __float_as_uint(v[4 * j + 3]);
    } else {
      __nv_bfloat162 p0 = __floats2bfloat162_rn(v[8 * j + 0], v[8 * j + 1]);
      __nv_bfloat162 p1 = __floats2bfloat162_rn(v[8 * j + 2], v[8 * j + 3]);
      __nv_bfloat162 p2 = __floats2bfloat162_rn(v[8 * j + 4], v[8 * j + 5]);
      __nv_bfloat162 p3 = __floats2bfloat162_rn(v[8 * j + 6], v[8 * j + 7]);
      w0 = *reinterpret_cast<uint32_t *>(&p0); w1 = *reinterpret_cast<uint32_t *>(&p1);
      w2 = *reinterpret_cast<uint32_t *>(&p2); w3 = *reinterpret_cast<uint32_t *>(&p3);
    }
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((j ^ sw) << 4)), "r"(w0), "r"(w1),
                 "r"(w2), "r"(w3)
                 : "memory");
  }
}

// ---------------------------------------------------------------- the kernel
template <typename VolT>
__global__ void __launch_bounds__(TC_THREADS, 1)
build_pyramid_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                        const __grid_constant__ CUtensorMap map_l0, const __grid_constant__ CUtensorMap map_l1,
                        const __grid_constant__ CUtensorMap map_l2, const __grid_constant__ CUtensorMap map_l3,
                        const TcParams p) {
  using S = Stg<VolT>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve (all offsets multiples of 1024)
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sA = smem_u32(smem);
  const uint32_t sB = sA + p.KB * A_KB_BYTES;
  const uint32_t sStage = sB + p.stages * B_ST_BYTES;
  const uint32_t sBar = sStage + 4 * 2 * S::BYTES;
  uint32_t *tmem_slot = (uint32_t *)(smem + (sBar - sA) + 128);
  // barrier table (8 bytes each)
  const uint32_t bar_a_full = sBar, bar_a_empty = sBar + 8;
  const uint32_t bar_b_full = sBar + 16;            // [stages]
  const uint32_t bar_b_empty = sBar + 16 + 8 * 4;   // [stages] (stages <= 4)
  const uint32_t bar_t_full = sBar + 16 + 8 * 8;    // [2]
  const uint32_t bar_t_empty = bar_t_full + 16;     // [2]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // contiguous tile range of this CTA
  const long long t_begin = (long long)p.total_tiles * blockIdx.x / gridDim.x;
  const long long t_end = (long long)p.total_tiles * (blockIdx.x + 1) / gridDim.x;

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
      long long cur_key = -1;
      uint32_t a_loads = 0, bs = 0;
      for (long long t = t_begin; t < t_end; ++t) {
        const long long key = t / p.NT;          // (b, m)
        const int n = (int)(t - key * p.NT);
        const int b = (int)(key / p.MT), m = (int)(key - (long long)b * p.MT);
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
      long long cur_key = -1;
      uint32_t a_uses = 0, bs = 0, i = 0;
      for (long long t = t_begin; t < t_end; ++t, ++i) {
        const long long key = t / p.NT;
        if (key != cur_key) {
          mbar_wait(bar_a_full, a_uses & 1);
          ++a_uses;
          cur_key = key;
        }
        const uint32_t acc = i & 1, use = i >> 1;
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
        const bool last_of_key = (t + 1 >= t_end) || ((t + 1) / p.NT != key);
        if (last_of_key) umma_commit(bar_a_empty);
      }
    }
  } else {
    // ===================================================== epilogue (4 warps)
    const int quarter = warp & 3;                 // TMEM lanes [32*quarter, +32)
    const uint32_t stg0 = sStage + (warp - EPI_WARP0) * 2 * S::BYTES;
    const int L = p.num_levels;
    uint32_t i = 0, rpc = 0;
    for (long long t = t_begin; t < t_end; ++t, ++i) {
      const long long key = t / p.NT;
      const int n = (int)(t - key * p.NT);
      const int b = (int)(key / p.MT), m = (int)(key - (long long)b * p.MT);
      const int ty = n / p.nbx, tx = n - ty * p.nbx;
      const int x0 = tx * TB_W, y0 = ty * TB_H;
      const int qw = m * BM + quarter * 32;       // first query of this warp
      const uint32_t acc = i & 1, use = i >> 1;
      mbar_wait(bar_t_full + 8 * acc, use & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quarter * 32) << 16);

      float p1prev[16], p2prev[8];
#pragma unroll
      for (int rp = 0; rp < 4; ++rp, ++rpc) {
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
        const uint32_t buf = stg0 + (rpc & 1) * S::BYTES;
        if (lane == 0) tma_store_wait_read<1>();   // the buffer used two row-pairs ago is drained
        __syncwarp();

        float v0[32], v1[32], p1[16];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          v0[k] = __uint_as_float(r0[k]) * p.scale;
          v1[k] = __uint_as_float(r1[k]) * p.scale;
        }
        stage_row<VolT, 32>(buf + S::OFF_L0, lane, v0);
        stage_row<VolT, 32>(buf + S::OFF_L0 + 32 * 32 * S::ES, lane, v1);
#pragma unroll
        for (int k = 0; k < 16; ++k) p1[k] = 0.25f * ((v0[2 * k] + v0[2 * k + 1]) + (v1[2 * k] + v1[2 * k + 1]));
        if (L > 1) stage_row<VolT, 16>(buf + S::OFF_L1, lane, p1);
        float p2[8], p3[4];
        if (rp & 1) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            p2[k] = 0.25f * ((p1prev[2 * k] + p1prev[2 * k + 1]) + (p1[2 * k] + p1[2 * k + 1]));
          if (L > 2) stage_row<VolT, 8>(buf + S::OFF_L2, lane, p2);
          if (rp == 3) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              p3[k] = 0.25f * ((p2prev[2 * k] + p2prev[2 * k + 1]) + (p2[2 * k] + p2[2 * k + 1]));
            if (L > 3) {
              if (S::L3_TMA) {
                stage_row<VolT, 4>(buf + S::OFF_L3, lane, p3);
              } else {
                const int q = qw + lane, y3 = y0 >> 3, x3 = x0 >> 3;
                if (q < p.N && y3 < p.l3_H) {
                  VolT *dst = (VolT *)p.l3_ptr + ((size_t)b * p.N + q) * p.l3_slab + (size_t)y3 * p.l3_pitch + x3;
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    if (x3 + k < p.l3_W) dst[k] = __float2bfloat16_rn(p3[k]);
                }
              }
            }
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) p2prev[k] = p2[k];
          }
        } else {
#pragma unroll
          for (int k = 0; k < 16; ++k) p1prev[k] = p1[k];
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0 && qw < p.N) {
          const int ya = y0 + 2 * rp;
          if (x0 < p.Wl[0]) {
            if (ya < p.Hl[0]) tma_store_4d(&map_l0, buf + S::OFF_L0, x0, ya, qw, b);
            if (ya + 1 < p.Hl[0]) tma_store_4d(&map_l0, buf + S::OFF_L0 + 32 * 32 * S::ES, x0, ya + 1, qw, b);
          }
          if (L > 1 && (x0 >> 1) < p.Wl[1] && (ya >> 1) < p.Hl[1])
            tma_store_4d(&map_l1, buf + S::OFF_L1, x0 >> 1, ya >> 1, qw, b);
          if ((rp & 1) && L > 2 && (x0 >> 2) < p.Wl[2] && (ya >> 2) < p.Hl[2])
            tma_store_4d(&map_l2, buf + S::OFF_L2, x0 >> 2, ya >> 2, qw, b);
          if (rp == 3 && L > 3 && S::L3_TMA && (x0 >> 3) < p.Wl[3] && (ya >> 3) < p.Hl[3])
            tma_store_4d(&map_l3, buf + S::OFF_L3, x0 >> 3, ya >> 3, qw, b);
        }
        if (lane == 0) tma_store_commit();
      }
    }
    if (lane == 0) tma_store_wait_all();
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

CUtensorMapSwizzle swizzle_for(int row_bytes) {
  return row_bytes >= 128 ? CU_TENSOR_MAP_SWIZZLE_128B
         : row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
         : row_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                           : CU_TENSOR_MAP_SWIZZLE_NONE;
}

template <typename VolT>
int build_tc_t(const void *f1p, const void *f2p, int B, int D, int H, int W, const rc_layout &lay,
               VolT *pyr, cudaStream_t s) {
  using S = Stg<VolT>;
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
  p.N = N;
  p.MT = ceil_div(N, BM);
  p.nbx = ceil_div(W, TB_W);
  p.NT = p.nbx * ceil_div(H, TB_H);
  p.KB = ceil_div(D, BK);
  p.num_levels = L;
  p.scale = 1.0f / sqrtf((float)D);
  p.total_tiles = B * p.MT * p.NT;
  for (int l = 0; l < 4; ++l) {
    p.Hl[l] = l < L ? lay.H[l] : 0;
    p.Wl[l] = l < L ? lay.W[l] : 0;
  }
  if (L > 3) {
    p.l3_ptr = pyr + lay.offset[3];
    p.l3_slab = lay.slab[3];
    p.l3_pitch = lay.pitch[3];
    p.l3_H = lay.H[3];
    p.l3_W = lay.W[3];
  }
  const long long rows2 = (long long)p.NT * TB_N;

  int dev = 0, sms = 0, max_smem = 0;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  RC_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const size_t fixed = (size_t)p.KB * A_KB_BYTES + 4 * 2 * S::BYTES + 256 + 1024;
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

  CUtensorMap ma, mb, ml[4];
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
  const CUtensorMapDataType odt = sizeof(VolT) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  for (int l = 0; l < 4; ++l) {
    const int ll = l < L ? l : 0;   // unused maps alias level 0 (never dereferenced)
    const bool tma_level = (l < L) && (l < 3 || S::L3_TMA);
    const int lv = tma_level ? l : 0;
    (void)ll;
    const int bw = 32 >> lv;
    cuuint64_t dims[4] = {(cuuint64_t)lay.W[lv], (cuuint64_t)lay.H[lv], (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t str[3] = {(cuuint64_t)lay.pitch[lv] * S::ES, (cuuint64_t)lay.slab[lv] * S::ES,
                         (cuuint64_t)N * lay.slab[lv] * S::ES};
    cuuint32_t box[4] = {(cuuint32_t)bw, 1, 32, 1};
    int r = encode(&ml[l], odt, 4, pyr + lay.offset[lv], dims, str, box, swizzle_for(bw * S::ES), "pyramid level");
    if (r) return r;
  }

  auto kern = build_pyramid_tc_kernel<VolT>;
  RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = p.total_tiles < sms ? p.total_tiles : sms;
  kern<<<grid, TC_THREADS, smem, s>>>(ma, mb, ml[0], ml[1], ml[2], ml[3], p);
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
