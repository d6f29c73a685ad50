// Fused middle of the PacGAN discriminator for the packed batch (M = 50 ... 150 rows), SURVEY K2/K3/K7.
//
// Reference: Discriminator.forward (Server/dtds/synthesizers/ctgan.py:28-30) and its autograd backward.  With
// the standard two hidden layers the chain after the first (wide, tensor-core) layer is tiny and ROW LOCAL --
// there is no BatchNorm in D:
//     a1 = dropout(leaky(a0 . W1^T + b1))          [M, H1]   (+ the applied multiplier mask1)
//     y  = a1 . w + b                               [M]       (+ loss = sum_m rs_loss[m] * y[m])
//     Z1 = rs_sig (x) w (.) mask1                   [M, H1]   d loss / d pre-activation of layer 1
//     Z0 = (Z1 . W1) (.) mask0                      [M, H0]   d loss / d pre-activation of layer 0
// As three launches (tcgen05 GEMM, k_d_head, tcgen05 GEMM) these cost 22-24 us of dependent-launch latency per
// step for 6.6 MFLOP.  Here one CTA owns DM_RPB rows and runs the whole chain out of shared memory in exact fp32
// FMA: W1 (256 KB for 256 x 256) is streamed from L2 ONCE per CTA, used both ways (a warp per output for
// the forward dot products, a thread per input column for the transposed product), through a cp.async ring.  The scalar loss is summed in
// fixed row order by the last CTA to finish (deterministic).  Used when both hidden widths are <= 512.
#include "common.cuh"

namespace {

constexpr int DM_THREADS = 512;     // 16 warps: the chain is instruction-latency bound, 4 warps per scheduler hide it
constexpr int DM_RPB = 2;          // rows per CTA (75 / 25 CTAs for the 150- / 50-row packed batches)
static_assert(DM_RPB == 2, "the transposed Z1 staging is read as one float2 per n");
constexpr int DM_MAXH = 512;
constexpr int DM_CH = 32;           // W1 rows per shared-memory chunk

struct DMidArgs {
  const float* a0; int lda0;               // [M, H0] post-activation output of layer 0
  const float* w1; int ldw1;               // [H1, H0]
  const float* b1;                         // [H1]
  float* a1; int lda1;                     // [M, H1] out
  float* mask1; int ldm1;                  // [M, H1] out
  const float* mask0; int ldm0;            // [M, H0] multiplier of layer 0 (from its epilogue)
  const float* w_head; const float* b_head;
  float* y;                                // [M]
  const float* rs_loss; float* loss_out;   // loss = sum rs_loss[m] * y[m]
  const float* rs_sig;                     // [M]
  float* z1; int ldz1;                     // [M, H1] out
  float* z0; int ldz0;                     // [M, H0] out
  const float* drop_u; int ldu;            // optional uniforms [M, H1] (tests); else Philox
  const int64_t* rng_state; int rng_stream;
  float slope, drop_p;
  int M, H0, H1;
  int* counter;                            // zero-initialised int, re-armed by the kernel
};

// optional clock64 timeline of CTA 0 (latency analysis, like ftg_gemm_tc_timeline)
__device__ long long g_dmid_timeline[8];
__device__ int g_dmid_timeline_on = 0;
#define DTL(i) do { if (g_dmid_timeline_on && blockIdx.x == 0 && threadIdx.x == 0) g_dmid_timeline[i] = clock64(); } while (0)

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// W1 is streamed ONCE per CTA in chunks of DM_CH rows through a double-buffered cp.async ring (all 16 B requests of a
// chunk in flight together: the first version issued dependent L2 loads from the loops and took 76 us).  Each chunk
// serves both directions: the forward dot products of its DM_CH outputs n (a warp per n) and, as soon as their
// Z1[., n] are known, the transposed accumulation Z0[., k] += Z1[., n] * W1[n, k] (a thread per k).
__global__ void __launch_bounds__(DM_THREADS) k_d_mid(const DMidArgs P) {
  DTL(0);
  FTG_PDL_ENTER();
  DTL(1);
  extern __shared__ __align__(16) float wbuf[];                 // [nbuf][DM_CH][H0]
  __shared__ __align__(16) float a0s[DM_RPB][DM_MAXH];
  __shared__ __align__(16) float a1s[DM_RPB][DM_MAXH];
  __shared__ __align__(16) float z1t[DM_MAXH][DM_RPB];           // Z1 of this CTA's rows, [n][row]: one 16 B broadcast per n
  __shared__ float b1s[DM_MAXH], whs[DM_MAXH], us[DM_RPB][DM_MAXH], rss[DM_RPB];
  __shared__ float red[32];
  __shared__ int s_last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_base = blockIdx.x * DM_RPB;
  const int H0 = P.H0, H1 = P.H1;
  const int n_chunks = (H1 + DM_CH - 1) / DM_CH;
  const int row4 = H0 >> 2;                                     // float4 per W1 row
  // ring depth: a chunk needs ~2 us from L2 (measured: with 2 buffers every chunk waited 4-5 k cycles for its data)
  const int nbuf = (H0 <= 256) ? 4 : 2;
  // per-thread copy pattern of a chunk (identical for every chunk: only the base row moves)
  constexpr int CPT = (DM_CH * (DM_MAXH / 4) + DM_THREADS - 1) / DM_THREADS;     // float4 copies per thread, worst case
  int cp_n[CPT], cp_k[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    const int i = threadIdx.x + j * DM_THREADS;
    cp_n[j] = i < DM_CH * row4 ? i / row4 : -1;
    cp_k[j] = (i - (i / row4) * row4) * 4;
  }
  auto load_chunk = [&](int c) {
    float* dst = wbuf + (size_t)(c % nbuf) * DM_CH * H0;
    const int n0 = c * DM_CH;
    if (c < n_chunks) {
#pragma unroll
      for (int j = 0; j < CPT; ++j)
        if (cp_n[j] >= 0 && n0 + cp_n[j] < H1)
          cp_async16(dst + cp_n[j] * H0 + cp_k[j], P.w1 + (size_t)(n0 + cp_n[j]) * P.ldw1 + cp_k[j]);
    }
    cp_async_commit();                                          // (an empty group past the end keeps the count uniform)
  };
  for (int c = 0; c < nbuf - 1; ++c) load_chunk(c);
  for (int e = threadIdx.x; e < DM_RPB * H0; e += DM_THREADS) {
    const int r = e / H0, k = e - r * H0, m = m_base + r;
    a0s[r][k] = m < P.M ? P.a0[(size_t)m * P.lda0 + k] : 0.f;
  }
  for (int n = threadIdx.x; n < H1; n += DM_THREADS) {
    b1s[n] = P.b1[n];
    whs[n] = P.w_head[n];
  }
  if (P.drop_u)
    for (int e = threadIdx.x; e < DM_RPB * H1; e += DM_THREADS) {
      const int r = e / H1, n = e - r * H1, m = m_base + r;
      us[r][n] = m < P.M ? P.drop_u[(size_t)m * P.ldu + n] : 1.f;
    }
  if (threadIdx.x < DM_RPB) rss[threadIdx.x] = (m_base + threadIdx.x < P.M) ? P.rs_sig[m_base + threadIdx.x] : 0.f;
  const uint64_t seed = P.rng_state ? (uint64_t)P.rng_state[0] : 0ull, step = P.rng_state ? (uint64_t)P.rng_state[1] : 0ull;
  const float keep_scale = 1.f / (1.f - P.drop_p);
  float zacc[2][DM_RPB];                                        // partial Z0: k = (tid & 255) and + 256, this half's rows
#pragma unroll
  for (int j = 0; j < 2; ++j)
#pragma unroll
    for (int r = 0; r < DM_RPB; ++r) zacc[j][r] = 0.f;
  for (int c = 0; c < n_chunks; ++c) {
    load_chunk(c + nbuf - 1);                                   // refills the buffer chunk c - 1 has just released
    if (nbuf == 4) cp_async_wait<3>(); else cp_async_wait<1>();
    __syncthreads();                                            // chunk c (and, first time, a0s) visible to all
    if (c == 0) DTL(2);
    if (c == 1) DTL(3);
    const float* wc = wbuf + (size_t)(c % nbuf) * DM_CH * H0;
    const int n0 = c * DM_CH;
    // ---- forward: a warp owns DM_CH / 8 = 4 outputs n of the chunk; all operands come from shared memory
    // (b1, w_head, the dropout uniforms and the row scales were staged up front: a global load per output in
    // this loop costs an L2 round trip each and made the first version latency bound)
    {
      constexpr int NPW = DM_CH / (DM_THREADS / 32);            // outputs per warp per chunk
      float acc[NPW][DM_RPB];
#pragma unroll
      for (int q = 0; q < NPW; ++q)
#pragma unroll
        for (int r = 0; r < DM_RPB; ++r) acc[q][r] = 0.f;
      for (int k = lane * 4; k < H0; k += 128) {
        float4 a[DM_RPB];
#pragma unroll
        for (int r = 0; r < DM_RPB; ++r) a[r] = *reinterpret_cast<const float4*>(&a0s[r][k]);
#pragma unroll
        for (int q = 0; q < NPW; ++q) {
          const float4 w = *reinterpret_cast<const float4*>(wc + (warp * NPW + q) * H0 + k);
#pragma unroll
          for (int r = 0; r < DM_RPB; ++r)
            acc[q][r] = fmaf(a[r].x, w.x, fmaf(a[r].y, w.y, fmaf(a[r].z, w.z, fmaf(a[r].w, w.w, acc[q][r]))));
        }
      }
#pragma unroll
      for (int q = 0; q < NPW; ++q)
#pragma unroll
        for (int r = 0; r < DM_RPB; ++r) acc[q][r] = warp_sum(acc[q][r]);
      // lane (q, r) = (lane / RPB, lane % RPB) finishes output n = n0 + warp * NPW + q of row r
      if (lane < NPW * DM_RPB) {
        const int q = lane / DM_RPB, r = lane % DM_RPB;
        const int n = n0 + warp * NPW + q, m = m_base + r;
        float pre = 0.f;
#pragma unroll
        for (int qq = 0; qq < NPW; ++qq)
#pragma unroll
          for (int rr = 0; rr < DM_RPB; ++rr)
            if (qq == q && rr == r) pre = acc[qq][rr];
        if (n < H1) {
          pre += b1s[n];
          float mult = pre > 0.f ? 1.f : P.slope;
          float v = 0.f, z = 0.f;
          if (m < P.M) {
            float u;
            if (P.drop_u) {
              u = us[r][n];
            } else {
              const Philox4 ph = philox4x32_10(seed, step, (uint32_t)P.rng_stream, (uint32_t)(m * ((H1 + 3) >> 2) + (n >> 2)));
              const uint32_t wv = (n & 3) == 0 ? ph.x : (n & 3) == 1 ? ph.y : (n & 3) == 2 ? ph.z : ph.w;
              u = u01(wv);
            }
            mult = (u >= P.drop_p) ? mult * keep_scale : 0.f;
            v = pre * mult;
            z = rss[r] * whs[n] * mult;
            P.a1[(size_t)m * P.lda1 + n] = v;
            P.mask1[(size_t)m * P.ldm1 + n] = mult;
            P.z1[(size_t)m * P.ldz1 + n] = z;
          }
          a1s[r][n] = v;
          z1t[n][r] = z;
        }
      }
    }
    if (c == 0) DTL(6);
    __syncthreads();                                            // Z1[., n0 .. n0+CH) of this CTA's rows are in z1s
    if (c == 0) DTL(7);
    // ---- transposed accumulation on the same chunk: thread (k, half) covers half of the chunk's rows for input
    // column k (and k + 256 when H0 > 256); the two halves are combined once, after the last chunk
    {
      const int n_hi = min(DM_CH, H1 - n0);
      const int half = threadIdx.x >> 8, kk = threadIdx.x & 255;
      const int nl0 = half * (DM_CH / 2), nl1 = min(n_hi, nl0 + DM_CH / 2);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int k = kk + j * 256;
        if (k < H0) {
#pragma unroll 8
          for (int nl = nl0; nl < nl1; ++nl) {
            const float w = wc[nl * H0 + k];
            const float2 z = *reinterpret_cast<const float2*>(&z1t[n0 + nl][0]);
            zacc[j][0] = fmaf(z.x, w, zacc[j][0]); zacc[j][1] = fmaf(z.y, w, zacc[j][1]);
          }
        }
      }
    }
    __syncthreads();                                            // this chunk's buffer may be refilled now
  }
  DTL(4);
  // ---- head: y[m] = a1[m,:] . w + b   (warp r owns row r)
  if (warp < DM_RPB) {
    const int m = m_base + warp;
    float s = 0.f;
    for (int n = lane; n < H1; n += 32) s = fmaf(a1s[warp][n], whs[n], s);
    s = warp_sum(s);
    if (lane == 0 && m < P.M) P.y[m] = s + (P.b_head ? P.b_head[0] : 0.f);
  }
  // combine the two halves' partial Z0 through shared memory (a0s is free now), apply mask0, store
  {
    const int half = threadIdx.x >> 8, kk = threadIdx.x & 255;
    if (half == 1) {
#pragma unroll
      for (int j = 0; j < 2; ++j)
        if (kk + j * 256 < H0) { a0s[0][kk + j * 256] = zacc[j][0]; a0s[1][kk + j * 256] = zacc[j][1]; }
    }
    __syncthreads();
    if (half == 0) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int k = kk + j * 256;
        if (k < H0) {
#pragma unroll
          for (int r = 0; r < DM_RPB; ++r) {
            const int m = m_base + r;
            if (m < P.M) P.z0[(size_t)m * P.ldz0 + k] = (zacc[j][r] + a0s[r][k]) * P.mask0[(size_t)m * P.ldm0 + k];
          }
        }
      }
    }
  }
  DTL(5);
  // ---- loss: the last CTA to finish sums rs_loss[m] * y[m] over all rows in fixed order
  if (P.loss_out == nullptr) return;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int done = atomicAdd(P.counter, 1);
    s_last = (done == (int)gridDim.x - 1);
    if (s_last) *P.counter = 0;              // re-arm for the next launch
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  float cs = 0.f;
  for (int m = threadIdx.x; m < P.M; m += DM_THREADS) cs += P.rs_loss[m] * __ldcg(P.y + m);
  const float tot = block_sum(cs, red);
  if (threadIdx.x == 0) P.loss_out[0] = tot;
}

}  // namespace

extern "C" int ftg_d_mid_timeline(int enable, long long* out8) {
  cudaError_t e = cudaMemcpyToSymbol(g_dmid_timeline_on, &enable, sizeof(int));
  if (e != cudaSuccess) return (int)e;
  if (out8) e = cudaMemcpyFromSymbol(out8, g_dmid_timeline, 8 * sizeof(long long));
  return (int)e;
}

extern "C" int ftg_d_mid_supported(int H0, int H1) {
  return (H0 > 0 && H1 > 0 && H0 <= DM_MAXH && H1 <= DM_MAXH && (H0 & 3) == 0) ? 1 : 0;
}

extern "C" int ftg_d_mid(const float* a0, int lda0, int M, int H0, int H1, const float* w1, int ldw1, const float* b1,
                         float* a1, int lda1, float* mask1, int ldm1, const float* mask0, int ldm0, const float* w_head,
                         const float* b_head, float* y, const float* rs_loss, float* loss_out, const float* rs_sig,
                         float* z1, int ldz1, float* z0, int ldz0, const float* drop_u, int ldu,
                         const int64_t* rng_state, int rng_stream, float slope, float drop_p, int* counter,
                         ftg_stream_t stream) {
  if (!ftg_d_mid_supported(H0, H1) || M <= 0) return FTG_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(w1) & 15) || (ldw1 & 3)) return FTG_ERR_UNSUPPORTED;
  if (loss_out && !counter) return FTG_ERR_UNSUPPORTED;
  if (!drop_u && !rng_state) return FTG_ERR_UNSUPPORTED;
  DMidArgs P;
  P.a0 = a0; P.lda0 = lda0; P.w1 = w1; P.ldw1 = ldw1; P.b1 = b1; P.a1 = a1; P.lda1 = lda1; P.mask1 = mask1; P.ldm1 = ldm1;
  P.mask0 = mask0; P.ldm0 = ldm0; P.w_head = w_head; P.b_head = b_head; P.y = y; P.rs_loss = rs_loss; P.loss_out = loss_out;
  P.rs_sig = rs_sig; P.z1 = z1; P.ldz1 = ldz1; P.z0 = z0; P.ldz0 = ldz0; P.drop_u = drop_u; P.ldu = ldu;
  P.rng_state = rng_state; P.rng_stream = rng_stream; P.slope = slope; P.drop_p = drop_p; P.M = M; P.H0 = H0; P.H1 = H1;
  P.counter = counter;
  const size_t dyn = (size_t)(H0 <= 256 ? 4 : 2) * DM_CH * H0 * sizeof(float);
  static std::atomic<bool> attr_done[64];
  int dev_id = 0;
  if (cudaGetDevice(&dev_id) != cudaSuccess || dev_id < 0 || dev_id >= 64) return FTG_ERR_UNSUPPORTED;
  if (!attr_done[dev_id].load(std::memory_order_acquire)) {
    cudaError_t e = cudaFuncSetAttribute(k_d_mid, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)((size_t)2 * DM_CH * DM_MAXH * sizeof(float)));
    if (e != cudaSuccess) return (int)e;
    attr_done[dev_id].store(true, std::memory_order_release);
  }
  ftg_launch(k_d_mid, dim3(ftg_cdiv(M, DM_RPB)), dim3(DM_THREADS), dyn, (cudaStream_t)stream, P);
  FTG_LAUNCH_CHECK();
  return 0;
}
