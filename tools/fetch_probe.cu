// Does B200's L2 fetch 32-byte sectors from HBM on their own, or always 64-byte pairs?  (The lookup's layout question:
// 64-byte pyramid tiles assume the pair.)  Standalone:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/fetch_probe tools/fetch_probe.cu
//   ncu --metrics dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum tools/_build/fetch_probe
// Each thread reads ONE 16-byte word from the first 32-byte sector of a distinct, pseudo-randomly chosen 128-byte
// line of a 4 GiB buffer (so the neighbouring sectors are never wanted), once per setting of
// cudaLimitMaxL2FetchGranularity (32, 64, 128).  DRAM bytes per request = dram__bytes_read.sum / requests.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void sparse_read(const uint4 *buf, size_t lines, uint4 *sink, unsigned mul) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t line = (i * (size_t)mul) % lines;        // odd multiplier, power-of-two line count: a permutation
  const uint4 v = __ldcg(buf + line * 8);               // first 16 bytes of the 128-byte line
  if (v.x == 0x12345678u) sink[0] = v;                  // never true; keeps the load
}

int main() {
  const size_t bytes = 4ull << 30, lines = bytes / 128;
  uint4 *buf, *sink;
  CK(cudaMalloc(&buf, bytes));
  CK(cudaMalloc(&sink, 64));
  CK(cudaMemset(buf, 0, bytes));
  const size_t requests = 1ull << 24;                   // 16 M requests = 512 MB if 32 bytes each
  for (size_t gran : {32, 64, 128}) {
    cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran);
    size_t got = 0;
    cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaMemset(buf, 0, 1 << 20));                    // separates the launches in the ncu list
    CK(cudaEventRecord(e0));
    sparse_read<<<(unsigned)(requests / 256), 256>>>(buf, lines, sink, 2654435761u);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("{\"limit_requested\": %zu, \"set_limit_rc\": %d, \"limit_reported\": %zu, \"requests\": %zu, \"ms\": %.3f}\n", gran,
           (int)e, got, requests, ms);
  }
  return 0;
}
