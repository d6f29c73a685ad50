// fp32 CUDA-core GEMM with the fused epilogue (exact-fp32 fallback and numerics oracle for
// the tcgen05 path).  64x64 output tile, BK = 16, 256 threads, 4x4 register micro-tile.
#include "common.cuh"

#define SBM 64
#define SBN 64
#define SBK 16

__global__ void __launch_bounds__(256) k_gemm_simt(const ftg_gemm_args p) {
  __shared__ float As[SBK][SBM + 4];
  __shared__ float Bs[SBK][SBN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * SBM, n0 = blockIdx.x * SBN;
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < p.K; k0 += SBK) {
    // ---- stage A tile (SBM x SBK) : pick the thread->element map that walks the contiguous axis
#pragma unroll
    for (int it = 0; it < (SBM * SBK) / 256; ++it) {
      int e = tid + it * 256;
      int mm, kk;
      if (p.a_t) { mm = e % SBM; kk = e / SBM; } else { kk = e % SBK; mm = e / SBK; }
      int gm = m0 + mm, gk = k0 + kk;
      float v = 0.f;
      if (gm < p.M && gk < p.K) v = p.a_t ? p.a[(size_t)gk * p.lda + gm] : p.a[(size_t)gm * p.lda + gk];
      As[kk][mm] = v;
    }
#pragma unroll
    for (int it = 0; it < (SBN * SBK) / 256; ++it) {
      int e = tid + it * 256;
      int nn, kk;
      if (p.b_t) { nn = e % SBN; kk = e / SBN; } else { kk = e % SBK; nn = e / SBK; }
      int gn = n0 + nn, gk = k0 + kk;
      float v = 0.f;
      if (gn < p.N && gk < p.K) v = p.b_t ? p.b[(size_t)gk * p.ldb + gn] : p.b[(size_t)gn * p.ldb + gk];
      Bs[kk][nn] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SBK; ++kk) {
      float a4[4], b4[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a4[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b4[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a4[i], b4[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n < p.N) ftg_epilogue_store(p, m, n, acc[i][j]);
    }
  }
}

extern "C" int ftg_gemm_simt(const ftg_gemm_args* args, ftg_stream_t stream) {
  const ftg_gemm_args& p = *args;
  if (p.M <= 0 || p.N <= 0) return 0;
  dim3 grid(ftg_cdiv(p.N, SBN), ftg_cdiv(p.M, SBM));
  k_gemm_simt<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
  FTG_LAUNCH_CHECK();
  return 0;
}
