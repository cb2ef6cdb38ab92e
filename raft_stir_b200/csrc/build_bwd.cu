// Backward of the volume build + pyramid (what autograd runs for
// core/corr.py:13-27 and :52-60 during train.py:175).
//
// Pooling is linear, so level l of the pyramid is fmap1^T . pool_l(fmap2) and
// the gradient pyramid can be consumed level by level without ever un-pooling
// it to volume size:
//   dfmap1[b,d,q]      = s * sum_l sum_t  dC_l[b,q,t] * pool_l(fmap2)[b,d,t]
//   dpool_l[b,d,t]     = s * sum_q        dC_l[b,q,t] * fmap1[b,d,q]
//   dfmap2[b,d,y,x]    = sum_l dpool_l[b,d,y>>l,x>>l] / 4^l      (inside level l's footprint)
// with s = 1/sqrt(D).  A slab of the gradient pyramid is the query's map in tile
// order (rc_layout), i.e. a fixed permutation of the targets plus zero padding;
// pool_l(fmap2) is laid out with the same permutation, so the two contractions
// are plain fp32 GEMMs over whole slabs and go to cuBLAS
// SGEMM, the same arithmetic the reference's autograd uses; the pooled/padded
// copy of fmap2 and the fold back to fmap2's shape are small kernels here.
#include <cublas_v2.h>

#include "common.cuh"

namespace rc {
namespace {

// f2pad_l[b][d][tile_offset(y, x)] = mean of the 2^l x 2^l block of fmap2[b][d], 0 in the padding
__global__ void __launch_bounds__(256)
pool_pad_kernel(const float *__restrict__ fmap2, float *__restrict__ dst, int BD, int H, int W, int l,
                int Hl, int Wl, int tiles_x, long long per) {
  const long long total = (long long)BD * per;
  const int k = 1 << l;
  const float inv = 1.0f / (float)(k * k);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long bd = i / per;
    const int r = (int)(i - bd * per);
    const int tile = r / Tile<float>::ELEMS, in = r - tile * Tile<float>::ELEMS;
    const int y = (tile / tiles_x) * TILE_H + in / Tile<float>::W;
    const int x = (tile % tiles_x) * Tile<float>::W + in % Tile<float>::W;
    float v = 0.f;
    if (x < Wl && y < Hl) {
      const float *src = fmap2 + (size_t)bd * H * W + (size_t)(y * k) * W + x * k;
      for (int dy = 0; dy < k; ++dy)
        for (int dx = 0; dx < k; ++dx) v += src[(size_t)dy * W + dx];
      v *= inv;
    }
    dst[i] = v;
  }
}

struct FoldLevels {
  int L;
  int Hl[RC_MAX_LEVELS], Wl[RC_MAX_LEVELS], tiles_x[RC_MAX_LEVELS];
  long long off[RC_MAX_LEVELS];    // element offset of level l in the dpool buffer
  long long slab[RC_MAX_LEVELS];
};

__global__ void __launch_bounds__(256)
fold_kernel(const float *__restrict__ dpool, float *__restrict__ dfmap2, int BD, int H, int W, FoldLevels f) {
  const long long total = (long long)BD * H * W;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long bd = i / ((long long)H * W);
    const int r = (int)(i - bd * H * W);
    const int y = r / W, x = r - y * W;
    float v = 0.f;
    float w = 1.f;
    for (int l = 0; l < f.L; ++l, w *= 0.25f) {
      const int yl = y >> l, xl = x >> l;
      if (yl < f.Hl[l] && xl < f.Wl[l])
        v += w * dpool[f.off[l] + bd * f.slab[l] + tile_offset<float>(yl, xl, f.tiles_x[l])];
    }
    dfmap2[i] = v;
  }
}

cublasHandle_t handle_for_device(int dev) {
  static thread_local cublasHandle_t handles[64] = {};
  if (dev < 0 || dev >= 64) return nullptr;
  if (!handles[dev]) {
    if (cublasCreate(&handles[dev]) != CUBLAS_STATUS_SUCCESS) handles[dev] = nullptr;
  }
  return handles[dev];
}

int grid_for(long long total) {
  long long b = (total + 255) / 256;
  return (int)(b < 148 * 16 ? (b < 1 ? 1 : b) : 148 * 16);
}

}  // namespace

int64_t build_bwd_workspace_bytes(int B, int D, int H, int W, int num_levels) {
  rc_layout g;
  if (rc_pyramid_layout(B, H, W, num_levels, RC_VOL_F32, &g) != RC_OK) return RC_EINVAL;
  int64_t slabs = 0;
  for (int l = 0; l < num_levels; ++l) slabs += g.slab[l];
  return 2 * round_up64((int64_t)B * D * slabs * (int64_t)sizeof(float), 256);
}

int launch_build_bwd(const float *gpyr, const float *fmap1, const float *fmap2, int B, int D, int H, int W,
                     const rc_layout &g, float *dfmap1, float *dfmap2, void *workspace,
                     size_t workspace_bytes, cudaStream_t s) {
  (void)workspace_bytes;
  const int N = H * W, L = g.num_levels;
  int64_t slabs = 0;
  FoldLevels f{};
  f.L = L;
  for (int l = 0; l < L; ++l) {
    f.Hl[l] = g.H[l]; f.Wl[l] = g.W[l]; f.tiles_x[l] = g.tiles_x[l];
    f.slab[l] = g.slab[l];
    f.off[l] = (long long)B * D * slabs;   // level l block: [B*D][slab_l]
    slabs += g.slab[l];
  }
  float *f2pad = (float *)workspace;
  float *dpool = (float *)((char *)workspace + round_up64((int64_t)B * D * slabs * (int64_t)sizeof(float), 256));

  int dev = 0;
  RC_CUDA(cudaGetDevice(&dev));
  cublasHandle_t h = handle_for_device(dev);
  if (!h) {
    set_error("cublasCreate failed");
    return RC_ECUDA;
  }
  if (cublasSetStream(h, s) != CUBLAS_STATUS_SUCCESS) {
    set_error("cublasSetStream failed");
    return RC_ECUDA;
  }
  const float scale = 1.0f / sqrtf((float)D), zero = 0.f, one = 1.f;
  for (int l = 0; l < L; ++l) {
    float *f2l = f2pad + f.off[l];
    float *dpl = dpool + f.off[l];
    const long long total = (long long)B * D * g.slab[l];
    pool_pad_kernel<<<grid_for(total), 256, 0, s>>>(fmap2, f2l, B * D, H, W, l, g.H[l], g.W[l], g.tiles_x[l],
                                                    g.slab[l]);
    RC_LAUNCH_CHECK("pool_pad_kernel");
    const float *dC = gpyr + g.offset[l];
    const int K = (int)g.slab[l];
    // dfmap1[b] (row-major [D][N]) (+)= scale * f2pad_l[b] ([D][K]) . dC_l[b]^T ([K][N])
    cublasStatus_t st = cublasSgemmStridedBatched(
        h, CUBLAS_OP_T, CUBLAS_OP_N, N, D, K, &scale, dC, K, (long long)N * K, f2l, K, (long long)D * K,
        l == 0 ? &zero : &one, dfmap1, N, (long long)D * N, B);
    if (st != CUBLAS_STATUS_SUCCESS) {
      set_error("cublasSgemmStridedBatched (dfmap1, level %d) failed: %d", l, (int)st);
      return RC_ECUDA;
    }
    // dpool_l[b] (row-major [D][K]) = scale * fmap1[b] ([D][N]) . dC_l[b] ([N][K])
    st = cublasSgemmStridedBatched(h, CUBLAS_OP_N, CUBLAS_OP_N, K, D, N, &scale, dC, K, (long long)N * K,
                                   fmap1, N, (long long)D * N, &zero, dpl, K, (long long)D * K, B);
    if (st != CUBLAS_STATUS_SUCCESS) {
      set_error("cublasSgemmStridedBatched (dpool, level %d) failed: %d", l, (int)st);
      return RC_ECUDA;
    }
  }
  fold_kernel<<<grid_for((long long)B * D * N), 256, 0, s>>>(dpool, dfmap2, B * D, H, W, f);
  RC_LAUNCH_CHECK("fold_kernel");
  return RC_OK;
}

}  // namespace rc
