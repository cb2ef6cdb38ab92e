// The motion encoder's first layer on the tensor cores, fed straight by the lookup.
//
// Reference: BasicMotionEncoder / SmallMotionEncoder.forward, core/update.py:70-71 and :88-90:
//     cor = F.relu(self.convc1(corr))            convc1 = Conv2d(L*(2r+1)^2, O, 1)
// i.e. per query pixel a [C] x [C, O] product on the C = 324 (RAFT-small: 196) samples the lookup just produced.  The
// reference writes those samples as fp32 [B, C, H, W] (1 296 bytes per query), reads them back in the convolution
// and runs it in fp32.  Here the lookup leaves them as IEEE half in the same channel-major layout (648 bytes per
// query, rc_lookup_fwd_f16's kernel) and this kernel consumes them as the A operand of a tcgen05 GEMM:
//
//     out[b, o, q] = relu(bias[o] + sum_c  S[b, c, q] * Wt[o, c])        M = 128 queries, N = O, K = C
//
//   A = S   [128 q x 64 c] per k-block, MN-major (queries contiguous): two TMA boxes of 64 q x 64 c, 128-byte swizzle;
//           channels past C arrive as zeros (TMA out-of-bounds fill), so C needs no padding in memory
//   B = Wt  [O x C] fp16, K-major, RESIDENT in shared memory for the whole kernel: full 64-channel k-blocks in the
//           128-byte swizzle plus the 16-channel tail steps (C = 324 = 5 * 64 + 4) as 32-byte-swizzled tiles of
//           O x 32 bytes, so that 324 channels cost 168 KB instead of 192
//   D       fp32 in TMEM, two accumulators of O columns; epilogue: + bias, ReLU, lane = query -> 128-byte stores
//
// One persistent CTA per SM: warp 0 TMA producer (3-slot ring of 16 KB), warp 1 MMA issuer, warps 2..9 epilogue
// (two per TMEM lane quarter, alternate 16-channel slices).
// HBM-bound: per query it reads C * 2 bytes and writes O * 4 (fp32 out) or O * 2.
#include <cuda.h>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rc {
namespace {

using namespace tc;

constexpr int BM = 128;                       // queries per tile
constexpr int KB = 64;                        // channels per k-block
constexpr int A_CHUNK = 64 * 128;             // one box: 64 channels x 64 queries (128 bytes) = 8 KB
constexpr int A_STAGE = 2 * A_CHUNK;          // 16 KB
constexpr int STAGES = 3;
constexpr int MAX_O = 256;
constexpr int C1_EPI_WARPS = 8;               // two per TMEM lane quarter: each takes every other 16-channel slice
constexpr int C1_THREADS = 32 * (2 + C1_EPI_WARPS);

struct C1Params {
  int B, N, C, O;
  int nfull, tsteps;                          // full 64-channel k-blocks; 16-channel steps of the tail
  int mt;                                     // tiles per pair
  long long total;
  uint32_t idesc;
  uint32_t w_bytes;                           // resident weight bytes
  const float *bias;
  void *out;                                  // [B][O][N]
  int out_f16;
};

struct C1Maps {
  CUtensorMap a;          // samples  [B][C][Npad] fp16 (q extent N): box 64 q x 64 c
  CUtensorMap w_full;     // weights  [O][Cp] fp16: box 64 c x O, 128-byte swizzle
  CUtensorMap w_tail;     // the same tensor: box 16 c x O, 32-byte swizzle
};

// kind::f16, A = B = F16 (format 0), D = F32 (bit 4), A MN-major (bit 15), N>>3 at 17, M>>4 at 24
constexpr uint32_t idesc_f16_amn(int M, int N) {
  return (1u << 4) | (1u << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// K-major operand in the 32-byte swizzle: rows of 32 bytes (one K = 16 step of 16-bit elements), 8-row groups 256 bytes apart
constexpr uint32_t DESC_HI_SW32 = (256u >> 4) | (1u << 14) | (6u << 29);
__device__ __forceinline__ uint64_t desc_make_sw32(uint32_t saddr) {
  return ((uint64_t)DESC_HI_SW32 << 32) | ((saddr >> 4) & 0x3fffu) | (1u << 16);
}
// MN-major operand of 16-bit elements in the 128-byte swizzle: 128 contiguous bytes (64 elements) along M per K row,
// K rows 128 bytes apart, groups of 8 K rows SBO = 1024 bytes apart, successive 64-element spans along M LBO apart
__device__ __forceinline__ uint64_t desc_make_mn128(uint32_t saddr, uint32_t lbo_bytes) {
  return ((uint64_t)DESC_HI << 32) | ((saddr >> 4) & 0x3fffu) | (((lbo_bytes >> 4) & 0x3fffu) << 16);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

__global__ void __launch_bounds__(C1_THREADS, 1)
convc1_tc_kernel(const __grid_constant__ C1Maps maps, const C1Params p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sW = smem_u32(smem);                              // resident weights: full blocks, then tail steps
  const uint32_t sWtail = sW + p.nfull * (uint32_t)p.O * 128u;
  const uint32_t sRing = (sW + p.w_bytes + 1023u) & ~1023u;
  const uint32_t sBar = sRing + STAGES * A_STAGE;
  uint32_t *tmem_slot = (uint32_t *)(smem + (sBar - sW) + 96);
  const uint32_t bar_full = sBar, bar_empty = sBar + 8 * STAGES;   // [STAGES] each
  const uint32_t bar_t_full = sBar + 16 * STAGES;                  // [2]
  const uint32_t bar_t_empty = bar_t_full + 16;                    // [2]
  const uint32_t bar_w = bar_t_empty + 16;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = p.nfull + (p.tsteps > 0 ? 1 : 0);

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_t_full + 8 * s, 1);
      mbar_init(bar_t_empty + 8 * s, C1_EPI_WARPS);
    }
    mbar_init(bar_w, 1);
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
      // the weights, once
      mbar_expect_tx(bar_w, p.w_bytes);
      for (int kb = 0; kb < p.nfull; ++kb) tma_load_2d(sW + kb * p.O * 128, &maps.w_full, bar_w, kb * KB, 0);
      for (int t = 0; t < p.tsteps; ++t) tma_load_2d(sWtail + t * p.O * 32, &maps.w_tail, bar_w, p.nfull * KB + t * 16, 0);
      uint32_t st = 0, ph = 0;
      for (long long idx = blockIdx.x; idx < p.total; idx += gridDim.x) {
        const int b = (int)(idx / p.mt), m = (int)(idx - (long long)b * p.mt);
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(bar_empty + 8 * st, ph ^ 1);
          const uint32_t full = bar_full + 8 * st, dst = sRing + st * A_STAGE;
          mbar_expect_tx(full, A_STAGE);
          tma_load_3d_plain(dst, &maps.a, full, m * BM, kb * KB, b);
          tma_load_3d_plain(dst + A_CHUNK, &maps.a, full, m * BM + 64, kb * KB, b);
          if (++st == STAGES) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      mbar_wait(bar_w, 0);
      uint32_t st = 0, ph = 0, i = 0;
      for (long long idx = blockIdx.x; idx < p.total; idx += gridDim.x) {
        const uint32_t acc = i & 1;
        mbar_wait(bar_t_empty + 8 * acc, ((i >> 1) & 1) ^ 1);
        ++i;
        fence_after();
        const uint32_t d_tmem = tmem_base + acc * MAX_O;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(bar_full + 8 * st, ph);
          fence_after();
          const uint32_t a_addr = sRing + st * A_STAGE;
          const bool tail = kb >= p.nfull;
          const int steps = tail ? p.tsteps : 4;
          for (int k = 0; k < steps; ++k) {
            const uint64_t adesc = desc_make_mn128(a_addr + k * 2048, A_CHUNK);     // 16 channel rows per step
            const uint64_t bdesc = tail ? desc_make_sw32(sWtail + k * p.O * 32)
                                        : desc_make(desc_lo(sW + kb * p.O * 128 + k * 32));
            umma_bf16(d_tmem, adesc, bdesc, p.idesc, (kb | k) != 0);
          }
          umma_commit(bar_empty + 8 * st);
          if (++st == STAGES) { st = 0; ph ^= 1; }
        }
        umma_commit(bar_t_full + 8 * acc);
      }
    }
  } else {
    // ===================================================== epilogue: + bias, ReLU, out[b][o][q]
    // The epilogue is the long pole of this kernel (MMAs of a tile: 1.4 us; 256 scattered 128-byte stores per query
    // quarter behind 16 TMEM round trips: 7 us with one warp per quarter), so two warps share each quarter.
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    uint32_t i = 0;
    for (long long idx = blockIdx.x; idx < p.total; idx += gridDim.x) {
      const int b = (int)(idx / p.mt), m = (int)(idx - (long long)b * p.mt);
      const uint32_t acc = i & 1, t_par = (i >> 1) & 1;
      ++i;
      const int q = m * BM + quarter * 32 + lane;
      const bool ok = q < p.N;
      mbar_wait(bar_t_full + 8 * acc, t_par);
      fence_after();
      const uint32_t taddr = tmem_base + acc * MAX_O + ((uint32_t)(quarter * 32) << 16);
      for (int c = 16 * half; c < p.O; c += 32) {
        uint32_t v[16];
        tmem_ld16(taddr + c, v);
        tmem_ld_wait();
        if (ok) {
          if (p.out_f16) {
            __half *o = (__half *)p.out + ((long long)b * p.O + c) * p.N + q;
#pragma unroll
            for (int r = 0; r < 16; ++r)
              o[(long long)r * p.N] = __float2half_rn(fmaxf(__uint_as_float(v[r]) + __ldg(p.bias + c + r), 0.f));
          } else {
            float *o = (float *)p.out + ((long long)b * p.O + c) * p.N + q;
#pragma unroll
            for (int r = 0; r < 16; ++r) o[(long long)r * p.N] = fmaxf(__uint_as_float(v[r]) + __ldg(p.bias + c + r), 0.f);
          }
        }
      }
      fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_t_empty + 8 * acc);
    }
  }

  fence_before();
  __syncthreads();
  fence_after();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
  }
}

// weights fp32 [O][C] -> fp16 [O][Cp] (Cp = C rounded up to 8: TMA needs 16-byte row pitches), saturating
__global__ void __launch_bounds__(256)
convc1_pack_weights_kernel(const float *__restrict__ w, __half *__restrict__ out, int O, int C, int Cp) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= O * Cp) return;
  const int o = i / Cp, c = i - o * Cp;
  out[i] = c < C ? __float2half_rn(clamp_half(__ldg(w + (size_t)o * C + c))) : __float2half_rn(0.f);
}

}  // namespace

int convc1_cp(int C) { return (C + 7) / 8 * 8; }
int convc1_npad(int N) { return (N + 7) / 8 * 8; }

int launch_convc1_pack_weights(const float *w, int O, int C, void *w16, cudaStream_t s) {
  const int Cp = convc1_cp(C);
  convc1_pack_weights_kernel<<<ceil_div(O * Cp, 256), 256, 0, s>>>(w, (__half *)w16, O, C, Cp);
  RC_LAUNCH_CHECK("convc1_pack_weights_kernel");
  return RC_OK;
}

// samples: fp16 [B][C][Npad] (written by the lookup with row pitch Npad); out: [B][O][N] fp32 or fp16
int launch_convc1(const void *samples, const void *w16, const float *bias, int B, int N, int C, int O, void *out,
                  int out_f16, cudaStream_t s) {
  const int Cp = convc1_cp(C), Npad = convc1_npad(N);
  C1Params p{};
  p.B = B; p.N = N; p.C = C; p.O = O;
  p.nfull = C / KB;
  p.tsteps = ceil_div(C % KB, 16);
  p.mt = ceil_div(N, BM);
  p.total = (long long)B * p.mt;
  p.idesc = idesc_f16_amn(BM, O);
  p.w_bytes = (uint32_t)(p.nfull * O * 128 + p.tsteps * O * 32);
  p.bias = bias;
  p.out = out;
  p.out_f16 = out_f16;
  C1Maps maps;
  int r;
  {
    cuuint64_t dims[3] = {(cuuint64_t)N, (cuuint64_t)C, (cuuint64_t)B};
    cuuint64_t str[2] = {(cuuint64_t)Npad * 2, (cuuint64_t)C * Npad * 2};
    cuuint32_t box[3] = {64, KB, 1};
    if ((r = encode_tiled(&maps.a, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, samples, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B,
                          "lookup samples"))) return r;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)C, (cuuint64_t)O};
    cuuint64_t str[1] = {(cuuint64_t)Cp * 2};
    cuuint32_t box[2] = {KB, (cuuint32_t)O}, boxt[2] = {16, (cuuint32_t)O};
    if ((r = encode_tiled(&maps.w_full, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w16, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B,
                          "convc1 weights"))) return r;
    if ((r = encode_tiled(&maps.w_tail, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w16, dims, str, boxt, CU_TENSOR_MAP_SWIZZLE_32B,
                          "convc1 weights (tail)"))) return r;
  }
  int dev = 0, sms = 0, max_smem = 0;
  RC_CUDA(cudaGetDevice(&dev));
  RC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  RC_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const size_t smem = 1024 + round_up64(p.w_bytes, 1024) + (size_t)STAGES * A_STAGE + 256;
  RC_REQUIRE(smem <= (size_t)max_smem, "rc_lookup_convc1_fwd: %d x %d weights do not fit in shared memory", O, C);
  RC_CUDA(cudaFuncSetAttribute(convc1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = p.total < sms ? (int)p.total : sms;
  convc1_tc_kernel<<<grid, C1_THREADS, smem, s>>>(maps, p);
  RC_LAUNCH_CHECK("convc1_tc_kernel");
  return RC_OK;
}

}  // namespace rc
