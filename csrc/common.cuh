// Shared device helpers for the fedtgan_b200 kernel library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <atomic>

#define FTG_ERR_UNSUPPORTED 1000
#define FTG_ERR_SCRATCH_TOO_SMALL 1001

extern std::atomic<long long> g_ftg_launch_count;   // defined in elementwise.cu (RPC handler threads launch concurrently)

#define FTG_LAUNCH_CHECK()                         \
  do {                                             \
    ++g_ftg_launch_count;                          \
    cudaError_t e__ = cudaPeekAtLastError();       \
    if (e__ != cudaSuccess) return (int)e__;       \
  } while (0)

// ---- programmatic dependent launch (PDL) ------------------------------------------------
// Kernels of the training iteration are launched with programmatic stream serialisation: the
// next kernel's CTAs may become resident while the current one drains, and block in
// `griddepcontrol.wait` until the predecessor's memory is visible.  Hides launch latency and
// prologues of the ~50 small dependent kernels of an iteration (also inside CUDA graphs).
extern int g_ftg_pdl;            // defined in elementwise.cu; FEDTGAN_PDL=0 disables
#define FTG_PDL_ENTER()                                        \
  do {                                                         \
    asm volatile("griddepcontrol.launch_dependents;");         \
    asm volatile("griddepcontrol.wait;" ::: "memory");         \
  } while (0)

template <typename... KArgs, typename... Args>
static inline cudaError_t ftg_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_ftg_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

static inline int ftg_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---------------------------------------------------------------- Philox4x32-10
// Counter-based RNG: (seed, step) come from a device-resident int64[2] so that CUDA-graph
// replays draw fresh numbers; `stream` separates consumers inside one step and `idx`
// indexes the 4-vectors of a consumer.
struct Philox4 {
  uint32_t x, y, z, w;
};

__device__ __forceinline__ Philox4 philox4x32_10(uint64_t seed, uint64_t step, uint32_t stream, uint32_t idx) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32) ^ (stream * 0x85EBCA6Bu + 0x27D4EB2Fu);
  uint32_t c0 = idx, c1 = (uint32_t)step, c2 = (uint32_t)(step >> 32), c3 = stream;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  return Philox4{c0, c1, c2, c3};
}

__device__ __forceinline__ float u01(uint32_t r) { return (float)(r >> 8) * (1.0f / 16777216.0f); }

// one uniform in [0,1) for a scalar element index
__device__ __forceinline__ float philox_uniform(const int64_t* state, int stream, uint32_t elem) {
  Philox4 p = philox4x32_10((uint64_t)state[0], (uint64_t)state[1], (uint32_t)stream, elem >> 2);
  uint32_t r = (elem & 3u) == 0 ? p.x : (elem & 3u) == 1 ? p.y : (elem & 3u) == 2 ? p.z : p.w;
  return u01(r);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// block-wide sum for blockDim.x <= 1024 (result valid in all threads)
__device__ __forceinline__ float block_sum(float v, float* smem32) {
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) smem32[wid] = v;
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  float t = (threadIdx.x < nw) ? smem32[threadIdx.x] : 0.f;
  if (wid == 0) t = warp_sum(t);
  if (threadIdx.x == 0) smem32[0] = t;
  __syncthreads();
  return smem32[0];
}

// ------------------------------------------------------------- GEMM epilogue
#include "../include/fedtgan_b200.h"

// Dropout uniform of element (m, n): Philox block (m, n/4), component n%4 -- so that a thread
// owning 4 consecutive columns needs one Philox evaluation.
__device__ __forceinline__ Philox4 ftg_dropout_block(const ftg_gemm_args& p, int m, int n4) {
  return philox4x32_10((uint64_t)p.rng_state[0], (uint64_t)p.rng_state[1], (uint32_t)p.rng_stream,
                       (uint32_t)(m * ((p.N + 3) >> 2) + n4));
}

// Applies the fused epilogue to accumulator value `acc` of element (m, n) and stores it.
__device__ __forceinline__ void ftg_epilogue_store(const ftg_gemm_args& p, int m, int n, float acc) {
  float v = p.alpha * acc;
  if (p.bias) v += __ldg(p.bias + n);
  if (p.act) {
    float mult = v > 0.f ? 1.f : p.slope;
    if (p.act == 2) {
      float u;
      if (p.drop_u) {
        u = p.drop_u[(size_t)m * p.ld_drop_u + n];
      } else {
        Philox4 r = ftg_dropout_block(p, m, n >> 2);
        uint32_t w = (n & 3) == 0 ? r.x : (n & 3) == 1 ? r.y : (n & 3) == 2 ? r.z : r.w;
        u = u01(w);
      }
      mult = (u >= p.drop_p) ? mult / (1.f - p.drop_p) : 0.f;
    }
    if (p.mask_out) p.mask_out[(size_t)m * p.ld_mask_out + n] = mult;
    v *= mult;
  }
  if (p.mask_in) v *= p.mask_in[(size_t)m * p.ld_mask_in + n];
  float* dst = p.d + (size_t)m * p.ldd + n;
  if (p.accumulate) v += *dst;
  *dst = v;
}
