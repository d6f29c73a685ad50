// Fused federated weighted average over peer memory.
//
//   theta <- sum_i w_i * theta_i   (fixed client order, fp32)   written into EVERY replica
//
// Replaces, per round, the reference's D2H of both models, 2N pickled RPC transfers through
// rank 0, the Python-loop CPU average and the H2D on the way back
// (Server/dtds/distributed.py:86-106, 799-823, 414-417).  No NCCL on this path: the kernel
// itself loads peer replicas over NVLink (peer-mapped pointers obtained through CUDA IPC),
// applies the similarity weights, accumulates in fp32, truncates the integer slots
// (num_batches_tracked) and stores the result into every replica, with the cross-GPU
// barriers done by system-scope flags inside the same kernel.
//
//   two-shot  rank r reduces shard r of all replicas and broadcasts it (in place; traffic
//             per rank (N-1)/N*P in + (N-1)/N*P out over NVLink)         <- default
//   one-shot  every rank reduces the whole arena into a local staging buffer, barrier,
//             then copies back (latency-optimal for tiny models, N*P in)
//   local     all replicas on one GPU: one launch over all ranks' data (LocalGroup and the
//             single-GPU emulation used by the tests)
#include "common.cuh"

namespace {

constexpr int MAX_WORLD = 16;
constexpr int MAX_INT = 16;

struct IntSlots {
  int n;
  long long off[MAX_INT];
};

__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// spin until *p >= want; returns false on timeout
__device__ __forceinline__ bool wait_flag(const int* p, int want, unsigned long long deadline) {
  while (ld_acquire_sys(p) < want) {
    if (globaltimer_ns() > deadline) return false;
    __nanosleep(64);
  }
  return true;
}

__device__ __forceinline__ float4 trunc_slots(float4 v, long long e, const IntSlots& s) {
  for (int k = 0; k < s.n; ++k) {
    long long d = s.off[k] - e;
    if (d >= 0 && d < 4) {
      float* f = &v.x;
      f[d] = truncf(f[d]);
    }
  }
  return v;
}

struct Peers {
  float* p[MAX_WORLD];
  int* f[MAX_WORLD];
  float w[MAX_WORLD];
  const float* wdev;        // non-null: the weights live in device memory (output of ftg_fed_stats_to_weights)
};
__device__ __forceinline__ float peer_weight(const Peers& P, int i) { return P.wdev ? __ldg(P.wdev + i) : P.w[i]; }

// ------------------------------------------------------------------- local
__global__ void k_fedavg_local(Peers P, int world, long long n4, IntSlots slots) {
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int i = 0; i < world; ++i) {
      float4 v = reinterpret_cast<const float4*>(P.p[i])[q];
      float w = peer_weight(P, i);
      acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y); acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
    }
    if (slots.n) acc = trunc_slots(acc, q * 4, slots);
    for (int i = 0; i < world; ++i) reinterpret_cast<float4*>(P.p[i])[q] = acc;
  }
}

// cross-GPU barrier: every rank stores `round_id` into slot [phase*world + rank] of every
// peer's flag page and waits until its own page shows all peers.
__device__ __forceinline__ void signal_peers(const Peers& P, int rank, int world, int phase, int round_id) {
  if (threadIdx.x < world) st_release_sys(P.f[threadIdx.x] + phase * world + rank, round_id);
}
__device__ __forceinline__ bool wait_peers(const Peers& P, int rank, int world, int phase, int round_id,
                                           unsigned long long deadline) {
  bool ok = true;
  if (threadIdx.x < world) ok = wait_flag(P.f[rank] + phase * world + threadIdx.x, round_id, deadline);
  return __syncthreads_and(ok);
}

// ----------------------------------------------------------------- two-shot
// W = compile-time upper bound of the world size, U = float4 chunks per thread per trip:
// W * U = 16 independent 16 B peer loads in flight per thread (NVLink latency ~2 us needs
// >1 MB in flight per GPU to approach link bandwidth).
template <int W, int U>
__global__ void __launch_bounds__(256)
k_fedavg_twoshot(Peers P, int rank, int world, long long n4, IntSlots slots, int round_id, int* status,
                 unsigned long long timeout_ns) {
  const unsigned long long deadline = globaltimer_ns() + timeout_ns;
  int* myflags = P.f[rank];
  // barrier-in: my arena is final (previous kernels on this stream completed)
  if (blockIdx.x == 0) {
    __threadfence_system();
    signal_peers(P, rank, world, 0, round_id);
  }
  if (!wait_peers(P, rank, world, 0, round_id, deadline)) {
    if (threadIdx.x == 0) atomicExch(status, 1);
    return;
  }
  // reduce-scatter + all-gather of shard `rank`, element by element, in place
  const long long chunk = (n4 + world - 1) / world;
  const long long lo = (long long)rank * chunk, hi = min(n4, lo + chunk);
  const long long T = (long long)gridDim.x * blockDim.x;
  for (long long q0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; q0 < hi; q0 += T * U) {
    float4 v[U][W];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long q = q0 + u * T;
#pragma unroll
      for (int i = 0; i < W; ++i)
        if (i < world && q < hi) v[u][i] = reinterpret_cast<const float4*>(P.p[i])[q];   // peer loads over NVLink
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long q = q0 + u * T;
      if (q >= hi) continue;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int i = 0; i < W; ++i)
        if (i < world) {
          const float w = peer_weight(P, i);
          acc.x = fmaf(w, v[u][i].x, acc.x); acc.y = fmaf(w, v[u][i].y, acc.y);
          acc.z = fmaf(w, v[u][i].z, acc.z); acc.w = fmaf(w, v[u][i].w, acc.w);
        }
      if (slots.n) acc = trunc_slots(acc, q * 4, slots);
#pragma unroll
      for (int i = 0; i < W; ++i)
        if (i < world) reinterpret_cast<float4*>(P.p[i])[q] = acc;                        // peer stores over NVLink
    }
  }
  // barrier-out: all my stores are visible system-wide before peers may continue
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    int done = atomicAdd(myflags + 2 * world, 1);
    s_last = (done == (int)gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  if (threadIdx.x == 0) myflags[2 * world] = 0;
  __threadfence_system();
  signal_peers(P, rank, world, 1, round_id);
  if (!wait_peers(P, rank, world, 1, round_id, deadline)) {
    if (threadIdx.x == 0) atomicExch(status, 2);
  }
}

// ----------------------------------------------------------------- one-shot
__global__ void __launch_bounds__(256)
k_fedavg_oneshot(Peers P, float* staging, int rank, int world, long long n4, IntSlots slots, int round_id, int* status,
                 unsigned long long timeout_ns) {
  const unsigned long long deadline = globaltimer_ns() + timeout_ns;
  int* myflags = P.f[rank];
  if (blockIdx.x == 0) {
    __threadfence_system();
    signal_peers(P, rank, world, 0, round_id);
  }
  if (!wait_peers(P, rank, world, 0, round_id, deadline)) {
    if (threadIdx.x == 0) atomicExch(status, 1);
    return;
  }
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int i = 0; i < world; ++i) {
      float4 v = reinterpret_cast<const float4*>(P.p[i])[q];
      float w = peer_weight(P, i);
      acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y); acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
    }
    if (slots.n) acc = trunc_slots(acc, q * 4, slots);
    reinterpret_cast<float4*>(staging)[q] = acc;
  }
  // grid-wide + cross-GPU barrier: nobody overwrites an arena that a peer may still be reading
  __threadfence_system();
  __syncthreads();
  __shared__ int s_ok;
  if (threadIdx.x == 0) {
    int done = atomicAdd(myflags + 2 * world, 1);
    s_ok = 1;
    if (done == (int)gridDim.x - 1) {
      myflags[2 * world] = 0;
      for (int q = 0; q < world; ++q) st_release_sys(P.f[q] + world + rank, round_id);
      for (int q = 0; q < world; ++q)
        if (!wait_flag(myflags + world + q, round_id, deadline)) { atomicExch(status, 2); }
      st_release_sys(myflags + 2 * world + 1, round_id);                 // local "go"
    } else {
      if (!wait_flag(myflags + 2 * world + 1, round_id, deadline)) s_ok = 0;
    }
  }
  __syncthreads();
  if (!s_ok) return;
  float* mine = P.p[rank];
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x)
    reinterpret_cast<float4*>(mine)[q] = reinterpret_cast<const float4*>(staging)[q];
}

// ------------------------------------------------------------------- NVLS
// Same aggregation through the NVSwitch multicast engine: every rank first scales its own
// replica by w_rank in place, then rank r pulls shard r already summed by the switch
// (multimem.ld_reduce) and pushes the result to all replicas with one multicast store
// (multimem.st).  Per GPU that halves the NVLink traffic of the P2P two-shot kernel
// ((N-1)/N*P out, P/N + (N-1)/N*P in).  The reduction order inside the switch is not
// specified, so this variant is not bitwise reproducible across runs (replicas are still
// identical to each other: everybody receives the same multicast value).
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__global__ void __launch_bounds__(512)
k_fedavg_nvls(Peers P, float* mc_base, int rank, int world, long long n4, IntSlots slots, int round_id, int* status,
              unsigned long long timeout_ns, int uniform) {
  const unsigned long long deadline = globaltimer_ns() + timeout_ns;
  int* myflags = P.f[rank];
  __shared__ int s_last;
  // phase 0: theta_rank *= w_rank (whole local replica), then tell the peers.  With equal weights (IID
  // shards: w_i = 1/N exactly) the switch sums the raw replicas and the common factor is applied to the
  // reduced value instead: no extra pass over the arena.
  if (uniform) {
    if (blockIdx.x == 0) {
      __threadfence_system();
      signal_peers(P, rank, world, 0, round_id);
    }
  } else {
    float4* mine = reinterpret_cast<float4*>(P.p[rank]);
    const float w = peer_weight(P, rank);
    const long long stride0 = (long long)gridDim.x * blockDim.x;
    for (long long q0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; q0 < n4; q0 += stride0 * 4) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u * stride0 < n4) v[u] = mine[q0 + u * stride0];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (q0 + u * stride0 < n4) {
          v[u].x *= w; v[u].y *= w; v[u].z *= w; v[u].w *= w;
          mine[q0 + u * stride0] = v[u];
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      int done = atomicAdd(myflags + 2 * world, 1);
      s_last = (done == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
      if (threadIdx.x == 0) myflags[2 * world] = 0;
      __threadfence_system();
      signal_peers(P, rank, world, 0, round_id);
    }
  }
  if (!wait_peers(P, rank, world, 0, round_id, deadline)) {
    if (threadIdx.x == 0) atomicExch(status, 1);
    return;
  }
  // phase 1: in-switch reduce of my shard, multicast the result into every replica
  const long long chunk = (n4 + world - 1) / world;
  const long long lo = (long long)rank * chunk, hi = min(n4, lo + chunk);
  {
    constexpr int U = 8;                      // independent switch reductions in flight per thread
    const float post = uniform ? peer_weight(P, rank) : 1.f;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; q0 < hi; q0 += stride * U) {
      float4 acc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long q = q0 + u * stride;
        if (q < hi) acc[u] = multimem_ld_reduce_add(mc_base + q * 4);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long q = q0 + u * stride;
        if (q < hi) {
          acc[u].x *= post; acc[u].y *= post; acc[u].z *= post; acc[u].w *= post;
          if (slots.n) acc[u] = trunc_slots(acc[u], q * 4, slots);
          multimem_st(mc_base + q * 4, acc[u]);
        }
      }
    }
  }
  // phase 2: barrier-out
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    int done = atomicAdd(myflags + 2 * world + 2, 1);
    s_last = (done == (int)gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  if (threadIdx.x == 0) myflags[2 * world + 2] = 0;
  __threadfence_system();
  signal_peers(P, rank, world, 1, round_id);
  if (!wait_peers(P, rank, world, 1, round_id, deadline)) {
    if (threadIdx.x == 0) atomicExch(status, 2);
  }
}

int fill(Peers& P, IntSlots& S, float* const* arenas, int* const* flags, int world, const float* weights,
         const int* int_offsets, int n_int) {
  if (world < 1 || world > MAX_WORLD || n_int > MAX_INT) return FTG_ERR_UNSUPPORTED;
  for (int i = 0; i < MAX_WORLD; ++i) { P.p[i] = nullptr; P.f[i] = nullptr; P.w[i] = 0.f; }
  // `weights` may be a device pointer: then the kernel reads it at run time
  cudaPointerAttributes at;
  const bool on_device = cudaPointerGetAttributes(&at, weights) == cudaSuccess &&
                         (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged);
  cudaGetLastError();       // a plain host pointer makes older runtimes report an (ignorable) error
  P.wdev = on_device ? weights : nullptr;
  for (int i = 0; i < world; ++i) {
    P.p[i] = arenas[i];
    P.f[i] = flags ? flags[i] : nullptr;
    if (!on_device) P.w[i] = weights[i];
  }
  S.n = n_int;
  for (int k = 0; k < n_int; ++k) S.off[k] = int_offsets[k];
  return 0;
}

}  // namespace

// NOTE: `arenas`, `flags`, `weights`, `int_offsets` are HOST arrays here (tiny; they are
// passed to the kernel by value), the pointers inside are device / peer-mapped pointers.
extern "C" int ftg_fedavg_local(float* const* arenas, int world, const float* weights, long long n,
                                const int* int_offsets, int n_int, ftg_stream_t stream) {
  Peers P; IntSlots S;
  int rc = fill(P, S, arenas, nullptr, world, weights, int_offsets, n_int);
  if (rc) return rc;
  if (n & 3) return FTG_ERR_UNSUPPORTED;
  long long n4 = n / 4;
  int blocks = ftg_cdiv(n4, 256);
  if (blocks > 148 * 4) blocks = 148 * 4;
  k_fedavg_local<<<blocks, 256, 0, (cudaStream_t)stream>>>(P, world, n4, S);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_fedavg_twoshot(float* const* peers, int* const* flags, int rank, int world, const float* weights,
                                  long long n, const int* int_offsets, int n_int, int round_id, int* status,
                                  int timeout_ms, ftg_stream_t stream) {
  Peers P; IntSlots S;
  int rc = fill(P, S, peers, flags, world, weights, int_offsets, n_int);
  if (rc) return rc;
  if (n & 3) return FTG_ERR_UNSUPPORTED;
  long long n4 = n / 4;
  long long per = (n4 + world - 1) / world;
  const int U = world <= 2 ? 8 : world <= 4 ? 4 : world <= 8 ? 2 : 1;
  int blocks = ftg_cdiv(per, 256 * U);
  if (blocks > 144) blocks = 144;          // all CTAs co-resident (they meet at the exit barrier)
  if (blocks < 1) blocks = 1;
  const unsigned long long tns = (unsigned long long)timeout_ms * 1000000ull;
  cudaStream_t st = (cudaStream_t)stream;
  if (world <= 2) k_fedavg_twoshot<2, 8><<<blocks, 256, 0, st>>>(P, rank, world, n4, S, round_id, status, tns);
  else if (world <= 4) k_fedavg_twoshot<4, 4><<<blocks, 256, 0, st>>>(P, rank, world, n4, S, round_id, status, tns);
  else if (world <= 8) k_fedavg_twoshot<8, 2><<<blocks, 256, 0, st>>>(P, rank, world, n4, S, round_id, status, tns);
  else k_fedavg_twoshot<16, 1><<<blocks, 256, 0, st>>>(P, rank, world, n4, S, round_id, status, tns);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_fedavg_oneshot(float* const* peers, int* const* flags, float* staging, int rank, int world,
                                  const float* weights, long long n, const int* int_offsets, int n_int, int round_id,
                                  int* status, int timeout_ms, ftg_stream_t stream) {
  Peers P; IntSlots S;
  int rc = fill(P, S, peers, flags, world, weights, int_offsets, n_int);
  if (rc) return rc;
  if (n & 3) return FTG_ERR_UNSUPPORTED;
  long long n4 = n / 4;
  int blocks = ftg_cdiv(n4, 256);
  if (blocks > 128) blocks = 128;          // CTAs spin on a local flag: must all be resident
  k_fedavg_oneshot<<<blocks, 256, 0, (cudaStream_t)stream>>>(P, staging, rank, world, n4, S, round_id, status,
                                                             (unsigned long long)timeout_ms * 1000000ull);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_fedavg_nvls(float* const* peers, int* const* flags, float* multicast_base, int rank, int world,
                               const float* weights, long long n, const int* int_offsets, int n_int, int round_id,
                               int* status, int timeout_ms, ftg_stream_t stream) {
  Peers P; IntSlots S;
  int rc = fill(P, S, peers, flags, world, weights, int_offsets, n_int);
  if (rc) return rc;
  if ((n & 3) || !multicast_base) return FTG_ERR_UNSUPPORTED;
  long long n4 = n / 4;
  int blocks = ftg_cdiv(n4, 512 * 4);
  if (blocks > 144) blocks = 144;          // CTAs meet at grid-wide counters: keep them co-resident
  if (blocks < 1) blocks = 1;
  // equal host-side weights (bitwise): skip the pre-scale pass, scale the reduced sum instead
  int uniform = P.wdev == nullptr;
  for (int i = 1; i < world && uniform; ++i) uniform = (P.w[i] == P.w[0]);
  k_fedavg_nvls<<<blocks, 512, 0, (cudaStream_t)stream>>>(P, multicast_base, rank, world, n4, S, round_id, status,
                                                          (unsigned long long)timeout_ms * 1000000ull, uniform);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------- peer-memory plumbing
extern "C" int ftg_symm_alloc(long long bytes, void** ptr) {
  cudaError_t e = cudaMalloc(ptr, (size_t)bytes);
  if (e != cudaSuccess) return (int)e;
  return (int)cudaMemset(*ptr, 0, (size_t)bytes);
}
extern "C" int ftg_symm_free(void* ptr) { return (int)cudaFree(ptr); }
extern "C" int ftg_symm_export(void* ptr, unsigned char handle_out[64]) {
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, ptr);
  if (e != cudaSuccess) return (int)e;
  static_assert(sizeof(h) == 64, "IPC handle size");
  memcpy(handle_out, &h, 64);
  return 0;
}
extern "C" int ftg_symm_open(const unsigned char handle[64], void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  return (int)cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}
extern "C" int ftg_symm_close(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }
extern "C" int ftg_enable_peer_access(int device, int peer) {
  int can = 0;
  cudaError_t e = cudaDeviceCanAccessPeer(&can, device, peer);
  if (e != cudaSuccess) return (int)e;
  if (!can) return FTG_ERR_UNSUPPORTED;
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(device);
  e = cudaDeviceEnablePeerAccess(peer, 0);
  cudaSetDevice(cur);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
  return (int)e;
}
