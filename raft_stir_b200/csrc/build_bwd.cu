// Backward of the volume build + pyramid (what autograd runs for
// core/corr.py:13-27 and :52-60 during train.py:175), on the tensor cores.
//
// Pooling is linear, so level l of the pyramid is fmap1^T . pool_l(fmap2) and
// the gradient pyramid can be consumed level by level without ever un-pooling
// it to volume size (s = 1/sqrt(D)):
//   dfmap1[b,d,q]   = s * sum_l sum_k  dC_l[b,q,k] * pool_l(fmap2)[b,d,k]          (GEMM 1)
//   dpool_l[b,d,k]  = s * sum_q        dC_l[b,q,k] * fmap1[b,d,q]                  (GEMM 2)
//   dfmap2[b,d,y,x] = sum_l dpool_l[b,d,y>>l,x>>l] / 4^l   (inside level l's footprint)
// A slab of the gradient pyramid is the query's map in tile order (rc_layout),
// i.e. a fixed permutation k of the targets plus zero padding; pool_l(fmap2) is
// laid out with the same permutation, so both contractions run over whole slabs.
//
// Three launches, no library call, nothing allocated:
//   1. bwd_prep_kernel   f1r = s*fmap1 and f2pad = s*pool_l(fmap2) in slab order, zero padded
//   2. build_bwd_tc_kernel   ONE persistent tcgen05 kind::tf32 kernel for both GEMMs.  A work item is an
//        accumulator of 128 rows x D columns in TMEM (two of them alternate):
//          kind 1: rows = 128 queries of a pair, contraction over the slabs of all levels
//                  A = dC_l   [128 q x 32 k]  K-major  (one TMA box of the gradient pyramid)
//                  B = f2pad  [D    x 32 k]  K-major
//          kind 2: rows = 128 slab positions of one level, contraction over the pair's queries
//                  A = dC_l   [128 k x 32 q]  MN-major (four TMA boxes of 32 k x 32 q in the 32-byte-atom
//                             swizzle MN-major tf32 operands need: the gradient pyramid is read in place,
//                             never transposed in memory)
//                  B = f1r    [D    x 32 q]  K-major
//        so both kinds share stage shape (16 KB + 32 KB), MMA shape (128 x D x 8, four per k-block)
//        and epilogue (lane = row -> 128-byte coalesced stores into out[d][row]).  Items of one pair are
//        adjacent in the work list, so the second read of its gradient slabs tends to hit L2.
//   3. bwd_fold_kernel   dfmap2 from the dpool levels
//
// Precision.  kind::tf32 reads the upper 19 bits of its fp32 operands.
//   fast    : features rounded to nearest tf32 by the prep, gradient read as is (truncated to tf32 by the
//             tensor core): about 5e-4 relative on the result.
//   precise : 3xTF32.  Four extra warps split every operand tile in shared memory into hi (what the tensor core
//             sees of the raw value) and lo = value - hi, and the MMA warp accumulates hi*hi + lo*hi + hi*lo:
//             fp32-grade results (the reference's autograd runs fp32 SGEMMs) at three times the MMA work.
#include <cuda.h>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace {

using namespace tc;

constexpr int BM = 128;                 // accumulator rows per work item (UMMA M)
constexpr int KBE = 32;                 // fp32 elements of the contraction per k-block (128 bytes)
constexpr int MAXDT = 256;              // feature channels per work item (UMMA N), at most
constexpr int A_BYTES = BM * 128;       // 16 KB
constexpr int B_BYTES = MAXDT * 128;    // 32 KB
constexpr int SLOT = A_BYTES + B_BYTES;
constexpr int EPI_WARP0 = 2, EPI_WARPS = 4;
constexpr int CONV_WARP0 = EPI_WARP0 + EPI_WARPS, CONV_WARPS = 4;
constexpr int ML = RC_MAX_LEVELS;

struct BwdMaps {
  CUtensorMap dc_k[ML];     // gradient level l, box 32 k x 128 q   (kind 1: K-major A)
  CUtensorMap dc_mn[ML];    // gradient level l, box 32 k x 32 q    (kind 2: MN-major A)
  CUtensorMap f2pad;        // [B][Dp][Kp], box 32 x DT
  CUtensorMap f1r;          // [B][Dp][Npad] (q extent N), box 32 x DT
};

struct BwdParams {
  int B, N, D, L;
  int DT, n_dt;             // channels per item, items along D
  int qt;                   // kind-1 items per pair and D-tile: ceil(N / 128)
  int kb1[ML];              // kind 1: k-blocks of level l = ceil(slab_l / 32)
  int koffp[ML];            // level l's offset on f2pad's k axis (multiple of 32)
  int t2[ML];               // kind 2: items of level l = ceil(slab_l / 128)
  int koffq[ML];            // level l's offset on dpool's k axis (multiple of 128)
  int kb2;                  // kind 2: k-blocks = ceil(N / 32)
  int nkb1;                 // sum of kb1
  int items_per_pair;       // (qt + sum t2) * n_dt
  long long total_items;
  int Kq;                   // dpool row pitch
  int seg;                  // k-blocks accumulated in TMEM before the epilogue takes the partial sum
  uint32_t idesc1, idesc2;
  float *dfmap1;            // [B][D][N]
  float *dpool;             // [B][D][Kq]
};

struct Item {
  int kind, b, mt, level, dt, nkb;
};

__device__ __forceinline__ Item decode_item(const BwdParams &p, long long idx) {
  Item it;
  it.b = (int)(idx / p.items_per_pair);
  int r = (int)(idx - (long long)it.b * p.items_per_pair);
  it.dt = r % p.n_dt;
  r /= p.n_dt;
  it.level = 0;
  if (r < p.qt) {
    it.kind = 1;
    it.mt = r;
    it.nkb = p.nkb1;
  } else {
    it.kind = 2;
    r -= p.qt;
    int l = 0;
    while (l < p.L - 1 && r >= p.t2[l]) { r -= p.t2[l]; ++l; }
    it.level = l;
    it.mt = r;
    it.nkb = p.kb2;
  }
  return it;
}

// kind::tf32: D = F32 (bit 4), A = B = TF32 (format 2 at bits 7 and 10), A MN-major = bit 15, N>>3 at 17, M>>4 at 24
constexpr uint32_t idesc_tf32(int M, int N, bool a_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (a_mn ? (1u << 15) : 0u) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

// PRECISE = false: 192 threads: warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2..5 epilogue.
// PRECISE = true : 320 threads: four more warps (6..9) write the lo parts of every landed stage into one lo slot.
template <bool PRECISE>
__global__ void __launch_bounds__(PRECISE ? 320 : 192, 1)
build_bwd_tc_kernel(const __grid_constant__ BwdMaps maps, const BwdParams p) {
  constexpr int STAGES = PRECISE ? 3 : 4;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sRing = smem_u32(smem);
  const uint32_t sLo = sRing + STAGES * SLOT;                       // PRECISE: lo parts of one stage (A then B)
  const uint32_t sBar = sLo + (PRECISE ? SLOT : 0);
  uint32_t *tmem_slot = (uint32_t *)(smem + (sBar - sRing) + 192);
  const uint32_t bar_full = sBar;                    // [STAGES]
  const uint32_t bar_empty = sBar + 8 * 4;           // [STAGES]
  const uint32_t bar_t_full = sBar + 8 * 8;          // [2]
  const uint32_t bar_t_empty = sBar + 8 * 10;        // [2]
  const uint32_t bar_lo_full = sBar + 8 * 12;        // PRECISE
  const uint32_t bar_lo_empty = sBar + 8 * 13;       // PRECISE

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t stage_bytes = A_BYTES + (uint32_t)p.DT * 128u;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_t_full + 8 * s, 1);
      mbar_init(bar_t_empty + 8 * s, EPI_WARPS);
    }
    mbar_init(bar_lo_full, CONV_WARPS);
    mbar_init(bar_lo_empty, 1);
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
    // ===================================================== TMA producer
    if (lane == 0) {
      const uint64_t pol_keep = l2_policy_evict_last();
      uint32_t st = 0, ph = 0;
      for (long long idx = blockIdx.x; idx < p.total_items; idx += gridDim.x) {
        const Item it = decode_item(p, idx);
        if (it.kind == 1) {
          for (int l = 0; l < p.L; ++l) {
            for (int kb = 0; kb < p.kb1[l]; ++kb) {
              mbar_wait(bar_empty + 8 * st, ph ^ 1);
              const uint32_t full = bar_full + 8 * st, dst = sRing + st * SLOT;
              mbar_expect_tx(full, stage_bytes);
              tma_load_3d_plain(dst, &maps.dc_k[l], full, kb * KBE, it.mt * BM, it.b);
              tma_load_3d(dst + A_BYTES, &maps.f2pad, full, p.koffp[l] + kb * KBE, it.dt * p.DT, it.b, pol_keep);
              if (++st == STAGES) { st = 0; ph ^= 1; }
            }
          }
        } else {
          const CUtensorMap *mp = &maps.dc_mn[it.level];
          for (int qb = 0; qb < p.kb2; ++qb) {
            mbar_wait(bar_empty + 8 * st, ph ^ 1);
            const uint32_t full = bar_full + 8 * st, dst = sRing + st * SLOT;
            mbar_expect_tx(full, stage_bytes);
#pragma unroll
            for (int c = 0; c < 4; ++c)
              tma_load_3d_plain(dst + c * 4096, mp, full, it.mt * BM + c * KBE, qb * KBE, it.b);
            tma_load_3d(dst + A_BYTES, &maps.f1r, full, qb * KBE, it.dt * p.DT, it.b, pol_keep);
            if (++st == STAGES) { st = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      uint32_t st = 0, ph = 0, i = 0, lph = 0;
      for (long long idx = blockIdx.x; idx < p.total_items; idx += gridDim.x) {
        const Item it = decode_item(p, idx);
        const bool mn = it.kind == 2;
        const uint32_t idesc = mn ? p.idesc2 : p.idesc1;
        for (int kb0 = 0; kb0 < it.nkb; kb0 += p.seg) {
          // one segment of the contraction per accumulator: the tensor core's fp32 accumulator truncates on every
          // add, an error that grows with the number of adds; partial sums are combined by the epilogue in fp32
          const uint32_t acc = i & 1;
          mbar_wait(bar_t_empty + 8 * acc, ((i >> 1) & 1) ^ 1);
          ++i;
          fence_after();
          const uint32_t d_tmem = tmem_base + acc * MAXDT;
          const int kend = min(it.nkb, kb0 + p.seg);
          for (int kb = kb0; kb < kend; ++kb) {
            mbar_wait(bar_full + 8 * st, ph);
            if (PRECISE) { mbar_wait(bar_lo_full, lph); }
            fence_after();
            const uint32_t a_addr = sRing + st * SLOT, b_addr = a_addr + A_BYTES;
#pragma unroll
            for (int pass = 0; pass < (PRECISE ? 3 : 1); ++pass) {
              // pass 0: hi * hi   pass 1: lo * hi   pass 2: hi * lo
              const uint32_t aa = pass == 1 ? sLo : a_addr;
              const uint32_t bb = pass == 2 ? sLo + A_BYTES : b_addr;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint64_t adesc = mn ? desc_make_mn32(aa + k * 1024, 4096) : desc_make(desc_lo(aa + k * 32));
                const uint64_t bdesc = desc_make(desc_lo(bb + k * 32));
                umma_tf32(d_tmem, adesc, bdesc, idesc, ((kb - kb0) | k | pass) != 0);
              }
            }
            umma_commit(bar_empty + 8 * st);         // raw stage free once these MMAs retire
            if (PRECISE) { umma_commit(bar_lo_empty); lph ^= 1; }
            if (++st == STAGES) { st = 0; ph ^= 1; }
          }
          umma_commit(bar_t_full + 8 * acc);
        }
      }
    }
  } else if (warp < CONV_WARP0) {
    // ===================================================== epilogue: accumulator rows -> out[d][row]
    const int quarter = warp & 3;                  // TMEM lanes [32*quarter, +32)
    uint32_t i = 0;
    for (long long idx = blockIdx.x; idx < p.total_items; idx += gridDim.x) {
      const Item it = decode_item(p, idx);
      const int row = it.mt * BM + quarter * 32 + lane;
      float *out;
      long long pitch;
      bool ok;
      if (it.kind == 1) {
        pitch = p.N;
        out = p.dfmap1 + (long long)it.b * p.D * pitch + row;
        ok = row < p.N;
      } else {
        pitch = p.Kq;
        out = p.dpool + (long long)it.b * p.D * pitch + p.koffq[it.level] + row;
        ok = true;                                 // every level's span on the k axis is padded to whole items
      }
      const int d0 = it.dt * p.DT;
      for (int kb0 = 0; kb0 < it.nkb; kb0 += p.seg) {
        const uint32_t acc = i & 1, t_par = (i >> 1) & 1;
        ++i;
        mbar_wait(bar_t_full + 8 * acc, t_par);
        fence_after();
        const uint32_t taddr = tmem_base + acc * MAXDT + ((uint32_t)(quarter * 32) << 16);
        for (int c = 0; c < p.DT; c += 16) {
          uint32_t v[16];
          tmem_ld16(taddr + c, v);
          tmem_ld_wait();
          if (kb0 == 0) {
#pragma unroll
            for (int r = 0; r < 16; ++r) {
              const int d = d0 + c + r;
              if (ok && d < p.D) out[(long long)d * pitch] = __uint_as_float(v[r]);
            }
          } else {          // later segments: this thread wrote the same addresses before
#pragma unroll
            for (int r = 0; r < 16; ++r) {
              const int d = d0 + c + r;
              if (ok && d < p.D) out[(long long)d * pitch] += __uint_as_float(v[r]);
            }
          }
        }
        fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_t_empty + 8 * acc);
      }
    }
  } else if (PRECISE) {
    // ===================================================== splitters: lo = value - (upper 19 bits of value)
    const int t = threadIdx.x - CONV_WARP0 * 32;   // 0..127
    const uint32_t chunks = stage_bytes >> 4;      // 16-byte chunks of a stage (A then B: same offsets in the lo slot)
    uint32_t st = 0, ph = 0, lph = 0;
    for (long long idx = blockIdx.x; idx < p.total_items; idx += gridDim.x) {
      const Item it = decode_item(p, idx);
      for (int kb = 0; kb < it.nkb; ++kb) {
        mbar_wait(bar_full + 8 * st, ph);
        mbar_wait(bar_lo_empty, lph ^ 1);
        const uint32_t src = sRing + st * SLOT;
        for (uint32_t c = t; c < chunks; c += CONV_WARPS * 32) {
          uint4 v;
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(src + c * 16) : "memory");
          const float lx = __uint_as_float(v.x) - __uint_as_float(v.x & 0xffffe000u);
          const float ly = __uint_as_float(v.y) - __uint_as_float(v.y & 0xffffe000u);
          const float lz = __uint_as_float(v.z) - __uint_as_float(v.z & 0xffffe000u);
          const float lw = __uint_as_float(v.w) - __uint_as_float(v.w & 0xffffe000u);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sLo + c * 16), "r"(__float_as_uint(lx)), "r"(__float_as_uint(ly)),
                       "r"(__float_as_uint(lz)), "r"(__float_as_uint(lw)) : "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_lo_full);
        lph ^= 1;
        if (++st == STAGES) { st = 0; ph ^= 1; }
      }
    }
  }

  fence_before();
  __syncthreads();
  fence_after();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
  }
}

// ---------------------------------------------------------------- operands of the two GEMMs
struct PrepLevels {
  int L, H, W;
  int Hl[ML], Wl[ML], tiles_y[ML], tiles_x[ML];
  int koffp[ML], kseg[ML];       // level l's span on f2pad's k axis: offset, length (slab rounded up to 32)
  int soff[ML];                  // offset of level l's logical map in shared memory (levels >= 1)
  int koffq[ML];                 // level l's offset on dpool's k axis
};

__device__ __forceinline__ float round_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

// One block per (pair, channel) row: blockIdx.x = b * Dp + d.
//   f1r  [b][d][q]                       = s * fmap1[b][d][q]                               (0 for d >= D, q >= N)
//   f2pad[b][d][koffp_l + tile_offset]   = s * level l of the 2x2 mean pyramid of fmap2[b][d] (0 in the padding)
// Levels >= 1 are pooled level by level in shared memory (the floor-truncating avg_pool2d chain of
// core/corr.py:24-27); every level is then written in slab order, one row of 4 x 4 tiles at a time, so that
// stores are contiguous and no index needs a division by a tile count.
template <bool ROUND>
__global__ void __launch_bounds__(256)
bwd_prep_kernel(const float *__restrict__ fmap1, const float *__restrict__ fmap2, float *__restrict__ f1r,
                float *__restrict__ f2pad, int D, int Dp, int N, int Npad, int Kp, float s, PrepLevels f) {
  extern __shared__ float pooled[];             // logical maps of levels 1 .. L-1
  const int bd = blockIdx.x, tid = threadIdx.x;
  const int b = bd / Dp, d = bd - b * Dp;
  float *o1 = f1r + (size_t)bd * Npad;
  float *o2 = f2pad + (size_t)bd * Kp;
  if (d >= D) {                                 // channel padding up to a multiple of 16: zero rows
    for (int i = tid; i < Npad; i += 256) o1[i] = 0.f;
    for (int i = tid; i < Kp; i += 256) o2[i] = 0.f;
    return;
  }
  const float *in1 = fmap1 + ((size_t)b * D + d) * N;
  const float *in2 = fmap2 + ((size_t)b * D + d) * N;
  for (int q = tid; q < Npad; q += 256) {
    const float v = q < N ? s * __ldg(in1 + q) : 0.f;
    o1[q] = ROUND ? round_tf32(v) : v;
  }
  // pooled levels, logical layout, unscaled
  for (int l = 1; l < f.L; ++l) {
    const int Wl = f.Wl[l], n = f.Hl[l] * Wl;
    float *dst = pooled + f.soff[l];
    if (l == 1) {
      for (int i = tid; i < n; i += 256) {
        const int y = i / Wl, x = i - y * Wl;
        const float *p0 = in2 + (size_t)(2 * y) * f.W + 2 * x;
        dst[i] = 0.25f * ((__ldg(p0) + __ldg(p0 + 1)) + (__ldg(p0 + f.W) + __ldg(p0 + f.W + 1)));
      }
    } else {
      const float *src = pooled + f.soff[l - 1];
      const int Ws = f.Wl[l - 1];
      for (int i = tid; i < n; i += 256) {
        const int y = i / Wl, x = i - y * Wl;
        const float *p0 = src + (2 * y) * Ws + 2 * x;
        dst[i] = 0.25f * ((p0[0] + p0[1]) + (p0[Ws] + p0[Ws + 1]));
      }
    }
    __syncthreads();
  }
  // slab order
  for (int l = 0; l < f.L; ++l) {
    const int Hl = f.Hl[l], Wl = f.Wl[l];
    const int row_elems = f.tiles_x[l] * 16;            // one row of tiles, contiguous in the slab
    const float *src = l == 0 ? nullptr : pooled + f.soff[l];
    float *dst = o2 + f.koffp[l];
    for (int ty = 0; ty < f.tiles_y[l]; ++ty) {
      for (int i = tid; i < row_elems; i += 256) {
        const int y = ty * TILE_H + ((i >> 2) & 3), x = (i >> 4) * 4 + (i & 3);
        float v = 0.f;
        if (y < Hl && x < Wl) v = s * (l == 0 ? __ldg(in2 + (size_t)y * f.W + x) : src[y * Wl + x]);
        dst[ty * row_elems + i] = ROUND ? round_tf32(v) : v;
      }
    }
    for (int i = f.tiles_y[l] * row_elems + tid; i < f.kseg[l]; i += 256) dst[i] = 0.f;
  }
}

// One block per (pair, channel) row of dfmap2: blockIdx.x = b * D + d.
//   dfmap2[b][d][y][x] = sum_l dpool_l[b][d][y >> l][x >> l] / 4^l   (inside level l's footprint)
// The dpool spans of levels >= 1 (a quarter of the row at most) are staged in shared memory in slab order; level 0
// is read in place, 16 bytes of a tile row at a time.
__global__ void __launch_bounds__(256)
bwd_fold_kernel(const float *__restrict__ dpool, float *__restrict__ dfmap2, int N, int Kq, int stage_elems,
                PrepLevels f) {
  extern __shared__ float coarse[];             // dpool[koffq_1 ...), slab order
  const int tid = threadIdx.x;
  const float *row = dpool + (size_t)blockIdx.x * Kq;
  const int base1 = f.L > 1 ? f.koffq[1] : 0;
  for (int i = tid; i < stage_elems; i += 256) coarse[i] = __ldg(row + base1 + i);
  __syncthreads();
  float *out = dfmap2 + (size_t)blockIdx.x * N;
  const int W = f.W;
  for (int r = tid; r < N; r += 256) {
    const int y = r / W, x = r - y * W;
    float v = __ldg(row + tile_offset<float>(y, x, f.tiles_x[0]));
    float w = 0.25f;
    for (int l = 1; l < f.L; ++l, w *= 0.25f) {
      const int yl = y >> l, xl = x >> l;
      if (yl < f.Hl[l] && xl < f.Wl[l])
        v += w * coarse[f.koffq[l] - base1 + (int)tile_offset<float>(yl, xl, f.tiles_x[l])];
    }
    out[r] = v;
  }
}

struct Geometry {
  int Dp, DT, n_dt, Npad, Kp, Kq;
  int koffp[ML], koffq[ML];
  int64_t f1r_bytes, f2pad_bytes, dpool_bytes;
};

Geometry geometry(int B, int D, int N, const rc_layout &g) {
  Geometry q{};
  q.Dp = (D + 15) / 16 * 16;
  q.DT = q.Dp < MAXDT ? q.Dp : MAXDT;
  q.n_dt = ceil_div(q.Dp, q.DT);
  q.Npad = (N + 3) / 4 * 4;
  for (int l = 0; l < g.num_levels; ++l) {
    q.koffp[l] = q.Kp;
    q.koffq[l] = q.Kq;
    q.Kp += (int)round_up64(g.slab[l], KBE);
    q.Kq += (int)round_up64(g.slab[l], BM);
  }
  q.f1r_bytes = round_up64((int64_t)B * q.Dp * q.Npad * 4, 1024);
  q.f2pad_bytes = round_up64((int64_t)B * q.Dp * q.Kp * 4, 1024);
  q.dpool_bytes = round_up64((int64_t)B * D * q.Kq * 4, 1024);
  return q;
}

}  // namespace

int64_t build_bwd_workspace_bytes(int B, int D, int H, int W, int num_levels) {
  rc_layout g;
  if (B < 1 || D < 1) return RC_EINVAL;
  if (rc_pyramid_layout(B, H, W, num_levels, RC_VOL_F32, &g) != RC_OK) return RC_EINVAL;
  const Geometry q = geometry(B, D, H * W, g);
  return q.f1r_bytes + q.f2pad_bytes + q.dpool_bytes;
}

int launch_build_bwd(const float *gpyr, const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                     const rc_layout &g, int precise, float *dfmap1, float *dfmap2, void *workspace,
                     cudaStream_t s) {
  const int N = H * W, L = g.num_levels;
  const Geometry q = geometry(B, D, N, g);
  RC_REQUIRE(((uintptr_t)gpyr & 15) == 0 && ((uintptr_t)workspace & 255) == 0,
             "rc_build_bwd: grad_pyramid must be 16-byte aligned and the workspace 256-byte aligned");
  RC_REQUIRE((long long)B * q.Dp < (1ll << 31), "rc_build_bwd: B * D too large");
  float *f1r = (float *)workspace;
  float *f2pad = (float *)((char *)workspace + q.f1r_bytes);
  float *dpool = (float *)((char *)workspace + q.f1r_bytes + q.f2pad_bytes);
  const float scale = 1.0f / sqrtf((float)D);

  PrepLevels f{};
  f.L = L; f.H = H; f.W = W;
  int pooled_elems = 0;
  for (int l = 0; l < L; ++l) {
    f.Hl[l] = g.H[l]; f.Wl[l] = g.W[l]; f.tiles_y[l] = g.tiles_y[l]; f.tiles_x[l] = g.tiles_x[l];
    f.koffp[l] = q.koffp[l];
    f.kseg[l] = (int)round_up64(g.slab[l], KBE);
    f.koffq[l] = q.koffq[l];
    f.soff[l] = pooled_elems;
    if (l >= 1) pooled_elems += g.H[l] * g.W[l];
  }
  int max_optin = 0, dev = 0;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));

  // ---- 1. operands
  {
    const size_t smem = (size_t)pooled_elems * sizeof(float);
    RC_REQUIRE(smem <= (size_t)max_optin, "rc_build_bwd: feature map %dx%d too large", H, W);
    auto kern = precise ? bwd_prep_kernel<false> : bwd_prep_kernel<true>;
    if (smem > 48 * 1024) RC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)(B * q.Dp), 256, smem, s>>>(fmap1, fmap2, f1r, f2pad, D, q.Dp, N, q.Npad, q.Kp, scale, f);
    RC_LAUNCH_CHECK("bwd_prep_kernel");
  }

  // ---- 2. both GEMMs
  {
    BwdMaps maps;
    BwdParams p{};
    p.B = B; p.N = N; p.D = D; p.L = L;
    p.DT = q.DT; p.n_dt = q.n_dt;
    p.qt = ceil_div(N, BM);
    p.kb2 = ceil_div(N, KBE);
    int items = p.qt;
    for (int l = 0; l < L; ++l) {
      p.kb1[l] = (int)((g.slab[l] + KBE - 1) / KBE);
      p.nkb1 += p.kb1[l];
      p.koffp[l] = q.koffp[l];
      p.t2[l] = (int)((g.slab[l] + BM - 1) / BM);
      p.koffq[l] = q.koffq[l];
      items += p.t2[l];
    }
    p.items_per_pair = items * p.n_dt;
    p.total_items = (long long)B * p.items_per_pair;
    p.Kq = q.Kq;
    p.seg = precise ? 48 : (1 << 30);
    p.idesc1 = idesc_tf32(BM, q.DT, false);
    p.idesc2 = idesc_tf32(BM, q.DT, true);
    p.dfmap1 = dfmap1;
    p.dpool = dpool;
    int r;
    for (int l = 0; l < L; ++l) {
      cuuint64_t dims[3] = {(cuuint64_t)g.slab[l], (cuuint64_t)N, (cuuint64_t)B};
      cuuint64_t str[2] = {(cuuint64_t)g.slab[l] * 4, (cuuint64_t)N * g.slab[l] * 4};
      cuuint32_t box1[3] = {KBE, BM, 1}, box2[3] = {KBE, KBE, 1};
      const float *base = gpyr + g.offset[l];
      if ((r = encode_tiled(&maps.dc_k[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, str, box1,
                            CU_TENSOR_MAP_SWIZZLE_128B, "gradient level (rows = queries)"))) return r;
      if ((r = encode_tiled(&maps.dc_mn[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, str, box2,
                            CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, "gradient level (rows = slab positions)"))) return r;
    }
    for (int l = L; l < ML; ++l) { maps.dc_k[l] = maps.dc_k[0]; maps.dc_mn[l] = maps.dc_mn[0]; }
    {
      cuuint64_t dims[3] = {(cuuint64_t)q.Kp, (cuuint64_t)q.Dp, (cuuint64_t)B};
      cuuint64_t str[2] = {(cuuint64_t)q.Kp * 4, (cuuint64_t)q.Dp * q.Kp * 4};
      cuuint32_t box[3] = {KBE, (cuuint32_t)q.DT, 1};
      if ((r = encode_tiled(&maps.f2pad, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, f2pad, dims, str, box,
                            CU_TENSOR_MAP_SWIZZLE_128B, "pooled fmap2"))) return r;
    }
    {
      cuuint64_t dims[3] = {(cuuint64_t)N, (cuuint64_t)q.Dp, (cuuint64_t)B};
      cuuint64_t str[2] = {(cuuint64_t)q.Npad * 4, (cuuint64_t)q.Dp * q.Npad * 4};
      cuuint32_t box[3] = {KBE, (cuuint32_t)q.DT, 1};
      if ((r = encode_tiled(&maps.f1r, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, f1r, dims, str, box,
                            CU_TENSOR_MAP_SWIZZLE_128B, "fmap1"))) return r;
    }
    int sms = 0;
    RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = p.total_items < sms ? (int)p.total_items : sms;
    if (precise) {
      const size_t smem = 3 * SLOT + SLOT + 256 + 1024;
      RC_CUDA(cudaFuncSetAttribute(build_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      build_bwd_tc_kernel<true><<<grid, 320, smem, s>>>(maps, p);
    } else {
      const size_t smem = 4 * SLOT + 256 + 1024;
      RC_CUDA(cudaFuncSetAttribute(build_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      build_bwd_tc_kernel<false><<<grid, 192, smem, s>>>(maps, p);
    }
    RC_LAUNCH_CHECK("build_bwd_tc_kernel");
  }

  // ---- 3. dfmap2
  {
    const int stage_elems = L > 1 ? q.Kq - q.koffq[1] : 0;
    const size_t smem = (size_t)stage_elems * sizeof(float);
    RC_REQUIRE(smem <= (size_t)max_optin, "rc_build_bwd: feature map %dx%d too large", H, W);
    if (smem > 48 * 1024)
      RC_CUDA(cudaFuncSetAttribute(bwd_fold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    bwd_fold_kernel<<<(unsigned)(B * D), 256, smem, s>>>(dpool, dfmap2, N, q.Kq, stage_elems, f);
    RC_LAUNCH_CHECK("bwd_fold_kernel");
  }
  return RC_OK;
}

}  // namespace rc
