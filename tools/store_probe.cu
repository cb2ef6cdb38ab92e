// Write-bandwidth probe for the volume build's roofline (B200, sm_100a).  Standalone:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/_build/store_probe tools/store_probe.cu
//   tools/_build/store_probe [GiB]          one JSON line per variant (CUDA events, best of 5 after a warm-up)
//
// The build writes 16x more than it reads, so the copy bandwidth in MEASURED_PEAKS.json (read + write) is not its
// ceiling; the question is what a pure write stream can reach here, and how that depends on the shape of the stream.
// Variants (all hand-written, 148 x k persistent CTAs):
//   seq_st16[_ef]     contiguous slice per CTA, 16-byte st.global per thread (ef = L2 evict_first hint)
//   seq_bulk          the same stream as cp.async.bulk shared -> global, 8 KB per copy
//   chunk<C>[_ef]     the build's pattern: many rows (one per query, 14 336 bytes apart like a Sintel bf16 level-0
//                     slab); at any time the grid writes the same C-byte piece of every row, then moves on to the
//                     next piece -- C contiguous bytes per row per visit
//   copy_ld16_st16    read + write of the same total bytes, for comparison with the copy peak
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void st16(void *dst, uint4 v) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st16_hint(void *dst, uint4 v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.b32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
}

template <bool EF>
__global__ void __launch_bounds__(512) seq_st16(uint4 *buf, size_t n16) {
  const uint64_t pol = policy_evict_first();
  const size_t per = (n16 + gridDim.x - 1) / gridDim.x;
  const size_t lo = per * blockIdx.x, hi = lo + per < n16 ? lo + per : n16;
  const uint4 v = make_uint4(blockIdx.x, threadIdx.x, 3, 4);
  for (size_t i = lo + threadIdx.x; i < hi; i += 4 * 512) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * 512 < hi) { if (EF) st16_hint(buf + i + u * 512, v, pol); else st16(buf + i + u * 512, v); }
  }
}

__global__ void __launch_bounds__(512) copy_ld16_st16(const uint4 *__restrict__ src, uint4 *dst, size_t n16) {
  const size_t per = (n16 + gridDim.x - 1) / gridDim.x;
  const size_t lo = per * blockIdx.x, hi = lo + per < n16 ? lo + per : n16;
  for (size_t i = lo + threadIdx.x; i < hi; i += 4 * 512) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) if (i + u * 512 < hi) v[u] = __ldg(src + i + u * 512);
#pragma unroll
    for (int u = 0; u < 4; ++u) if (i + u * 512 < hi) st16(dst + i + u * 512, v[u]);
  }
}

// shared -> global bulk copies of 8 KB; one elected thread per warp issues, 4 copies in flight per warp
__global__ void __launch_bounds__(256) seq_bulk(uint8_t *buf, size_t bytes) {
  extern __shared__ __align__(128) uint8_t sm[];
  for (int i = threadIdx.x; i < 8192 / 16; i += blockDim.x) ((uint4 *)sm)[i] = make_uint4(i, 2, 3, 4);
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const size_t per = ((bytes / 8192 + gridDim.x - 1) / gridDim.x) * 8192;
  const size_t lo = per * blockIdx.x, hi = lo + per < bytes ? lo + per : bytes;
  const int warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(sm);
    int k = 0;
    for (size_t off = lo + (size_t)warp * 8192; off + 8192 <= hi; off += (size_t)nw * 8192) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 8192;" ::"l"(buf + off), "r"(s) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (++k >= 4) asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

// rows of `row_bytes`, visited piece by piece: piece p (C bytes) of every row before piece p + 1 of any
template <int C, bool EF>
__global__ void __launch_bounds__(256) chunk_scatter(uint8_t *buf, int rows, int row_bytes) {
  const uint64_t pol = policy_evict_first();
  constexpr int LPC = C / 16;                          // lanes per chunk
  constexpr int RPW = 32 / (LPC > 32 ? 32 : LPC);      // rows per warp instruction (C <= 512)
  const int lane = threadIdx.x & 31, warp_g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int pieces = row_bytes / C;
  const uint4 v = make_uint4(lane, 2, 3, 4);
  for (int p = 0; p < pieces; ++p) {
    if (C <= 512) {
      for (int r0 = warp_g * RPW; r0 < rows; r0 += nwarps * RPW) {
        const int r = r0 + lane / LPC;
        if (r < rows) {
          uint8_t *dst = buf + (size_t)r * row_bytes + (size_t)p * C + (lane % LPC) * 16;
          if (EF) st16_hint(dst, v, pol); else st16(dst, v);
        }
      }
    } else {
      for (int r = warp_g; r < rows; r += nwarps) {
#pragma unroll
        for (int u = 0; u < C / 512; ++u) {
          uint8_t *dst = buf + (size_t)r * row_bytes + (size_t)p * C + u * 512 + lane * 16;
          if (EF) st16_hint(dst, v, pol); else st16(dst, v);
        }
      }
    }
  }
}

template <typename F>
static void run(const char *name, size_t bytes, double rw_factor, F launch) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int it = 0; it < 6; ++it) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (it > 0 && ms < best) best = ms;
  }
  printf("{\"variant\": \"%s\", \"GiB\": %.2f, \"ms\": %.4f, \"GB/s\": %.1f}\n", name, bytes / 1073741824.0, best,
         rw_factor * bytes / best / 1e6);
  fflush(stdout);
}

int main(int argc, char **argv) {
  const double gib = argc > 1 ? atof(argv[1]) : 4.0;
  const int row_bytes = 14336;
  const int rows = (int)(gib * 1073741824.0 / row_bytes);
  const size_t bytes = (size_t)rows * row_bytes;
  uint8_t *buf, *src;
  CK(cudaMalloc(&buf, bytes));
  CK(cudaMalloc(&src, bytes));
  CK(cudaMemset(src, 1, bytes));
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t n16 = bytes / 16;
  run("memset", bytes, 1, [&] { CK(cudaMemsetAsync(buf, 0, bytes)); });
  run("seq_st16", bytes, 1, [&] { seq_st16<false><<<sms * 4, 512>>>((uint4 *)buf, n16); });
  run("seq_st16_ef", bytes, 1, [&] { seq_st16<true><<<sms * 4, 512>>>((uint4 *)buf, n16); });
  CK(cudaFuncSetAttribute(seq_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192));
  run("seq_bulk", bytes, 1, [&] { seq_bulk<<<sms * 4, 256, 8192>>>(buf, bytes); });
  run("chunk128", bytes, 1, [&] { chunk_scatter<128, false><<<sms * 8, 256>>>(buf, rows, row_bytes); });
  run("chunk256", bytes, 1, [&] { chunk_scatter<256, false><<<sms * 8, 256>>>(buf, rows, row_bytes); });
  run("chunk256_ef", bytes, 1, [&] { chunk_scatter<256, true><<<sms * 8, 256>>>(buf, rows, row_bytes); });
  run("chunk512", bytes, 1, [&] { chunk_scatter<512, false><<<sms * 8, 256>>>(buf, rows, row_bytes); });
  run("chunk1024", bytes, 1, [&] { chunk_scatter<1024, false><<<sms * 8, 256>>>(buf, rows, row_bytes); });
  run("chunk1024_ef", bytes, 1, [&] { chunk_scatter<1024, true><<<sms * 8, 256>>>(buf, rows, row_bytes); });
  run("chunk2048", bytes, 1, [&] { chunk_scatter<2048, false><<<sms * 8, 256>>>(buf, rows, row_bytes); });
  run("copy_ld16_st16", bytes, 2, [&] { copy_ld16_st16<<<sms * 4, 512>>>((const uint4 *)src, (uint4 *)buf, n16); });
  return 0;
}
