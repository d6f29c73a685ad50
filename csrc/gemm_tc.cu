// tcgen05 / TMEM / TMA GEMM for sm_100a -- kind::tf32, fp32 accumulation in tensor memory.
//
//   D[M,N] (+)= epilogue( A[M,K] . B[N,K]^T )      (epilogue: common.cuh::ftg_epilogue_store)
//
// * operands are fp32 in global memory; TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) stages
//   them into shared memory (K-major: SWIZZLE_128B; MN-major: SWIZZLE_128B_ATOM_32B, the
//   only layout tf32 accepts for transposed operands), the tensor core reads them as tf32, the accumulator tile
//   (128 lanes x 64 columns fp32) lives in TMEM and is read back with tcgen05.ld;
// * both operand majors are supported natively (K-major: [rows, K]; MN-major: [K, rows]),
//   so forward (X.W^T), dgrad (dY.W) and wgrad (dY^T.X) need no transposed copies;
// * warp roles: warp 0 lane 0 = TMA producer, warp 1 = TMEM allocator + lane 0 MMA issuer,
//   warps 2..5 = epilogue (each owns one 32-lane TMEM quadrant); 6-stage mbarrier ring;
// * fp32x3 mode (template X3, ftg_gemm_args::x3): fp32-equivalent products on the tensor cores.  After a
//   stage has landed, the four epilogue warps split it in shared memory into hi = rna_tf32(x) (in
//   place) and lo = rna_tf32(x - hi) (second buffer of the stage) and the issuer accumulates
//   A_lo.B_hi + A_hi.B_lo + A_hi.B_hi into the same TMEM tile (3xTF32; the dropped lo.lo term is
//   ~2^-22 relative).  Matches the reference's fp32 matmuls (torch allow_tf32=False) to fp32
//   round-off while staying on tcgen05.
// * split-K: grid.z CTAs write partial tiles to a caller-provided workspace; the CTA that
//   arrives last at the tile's counter reduces them in fixed z order (deterministic) and
//   applies the fused epilogue.  The CTGAN shapes are skinny (M = 50..500, K up to ~8000),
//   so split-K is what fills the 148 SMs.
#include "common.cuh"
#include <cuda.h>
#include <mutex>
#include <cstdio>
#include <cstdlib>

namespace {

constexpr int BM = 128, BN = 64, BK = 32;       // BK floats = 128 B = one swizzle row
constexpr int MAX_STAGES = 6;                   // ring depth is a launch parameter (2..6), see choose_stages()
constexpr int A_STAGE_BYTES = BM * BK * 4;      // 16 KB
constexpr int B_STAGE_BYTES = BN * BK * 4;      //  8 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int TMEM_COLS = 64;
constexpr int NUM_THREADS = 192;
constexpr int MAX_STAGES_X3 = 4;                // fp32x3: a stage carries hi and lo copies (48 KB)
constexpr int smem_bytes(int stages, bool x3 = false, bool push = false) {
  return stages * STAGE_BYTES * (x3 ? 2 : 1) + 1024 /*align slack*/ + 256 /*barriers*/ +
         (push ? BM * BN * 4 + 8 * 3 * BN * 4 : 0) /*receive buffers of the push-based cluster reduction*/;
}
constexpr int SMEM_BYTES = smem_bytes(MAX_STAGES);
static_assert(smem_bytes(MAX_STAGES_X3, true) <= 227 * 1024, "fp32x3 ring does not fit shared memory");
static_assert(smem_bytes(MAX_STAGES_X3 - 1, true, true) <= 227 * 1024, "fp32x3 cluster ring does not fit shared memory");
static_assert(smem_bytes(MAX_STAGES, false, true) <= 227 * 1024, "tf32 cluster ring does not fit shared memory");

// hi/lo split of one fp32 for 3xTF32: hi = round-to-nearest tf32, lo = round-to-nearest tf32 of the
// (exactly representable) remainder.
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  uint32_t h, l;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  hi = __uint_as_float(h);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(x - hi));
  lo = __uint_as_float(l);
}
constexpr uint32_t SPIN_LIMIT = 1u << 24;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0, spins = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!done && ++spins > SPIN_LIMIT) __trap();   // never hang the GPU on a protocol bug
  }
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

// Asynchronous 16 B store into ANOTHER CTA's shared memory (same cluster) that signals `bytes` on an
// mbarrier of that CTA when it lands: the receiving side waits on its own barrier, nobody synchronises
// the whole cluster.  Addresses are shared::cluster addresses (mapa).
__device__ __forceinline__ void st_async_v4(uint32_t remote_addr, float4 v, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];"
               ::"r"(remote_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(remote_bar) : "memory");
}
__device__ __forceinline__ uint32_t map_cluster(uint32_t local_addr, int cta_rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(cta_rank));
  return remote;
}
constexpr int RECV_BYTES = BM * BN * 4;               // push split-K: one slot per source rank, R rows x 64 floats each
constexpr int STAT_RECV_BYTES = 8 * 3 * BN * 4;       // BatchNorm statistics of up to 8 cluster CTAs
constexpr int CLUSTER_EXTRA_BYTES = RECV_BYTES + STAT_RECV_BYTES;

// lo part relative to the hardware's own truncation of x to tf32 (x3 mode 2)
__device__ __forceinline__ float lo_trunc_tf32(float x) {
  const float hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  uint32_t l;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(x - hi));
  return __uint_as_float(l);
}

// Shared-memory matrix descriptor (sm_100 descriptor version 1).
//   bits [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout
//   layout 2 = SWIZZLE_128B (16 B atomicity)   : K-major operands
//   layout 1 = SWIZZLE_128B with 32 B atomicity : the only legal layout for MN-major tf32
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                              uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

constexpr int SLD = BN + 4;                      // staged tile row stride (floats): 16 B aligned, conflict-free
constexpr int CHUNKS = (BM * BN / 4) / 128;      // float4 chunks per epilogue thread (16)
constexpr int EPI_BATCH = CHUNKS / 2;

// Epilogue specialisations (compile-time, so the unrolled code stays tiny and branch-free).
enum { EPI_PLAIN = 0, EPI_ACC = 1, EPI_BIAS = 2, EPI_BIAS_ACT = 3, EPI_MASK = 4, EPI_GENERIC = 5, EPI_BN = 6 };

// Ragged / unaligned chunk (and the generic flag combination): element-wise path, out of line
// because it is cold; `p` by value so that the kernel itself never spills its parameters.
__device__ __noinline__ void epilogue_scalar4(const ftg_gemm_args p, int m, int n, float4 a) {
#pragma unroll 1
  for (int e = 0; e < 4; ++e) {
    const float v = e == 0 ? a.x : e == 1 ? a.y : e == 2 ? a.z : a.w;
    if (n + e < p.N) ftg_epilogue_store(p, m, n + e, v);
  }
}

// Fused epilogue: thread `et` owns float4 chunks f = it*128 + et of the staged 128 x 64 tile.
// Two batches of 8 chunks: all global loads of a batch are issued before its math and stores
// (the compiler cannot hoist loads over the stores itself: no-alias is unknown).
template <int EPI>
__device__ __forceinline__ void epilogue_tile(const ftg_gemm_args& p, bool vec_ok, const float* stage, int et, int m0,
                                              int n0, int f_base, int n_iter) {
  // chunks f = f_base + it*128 + et, it < n_iter (whole tile: f_base 0, n_iter 16)
#pragma unroll 1
  for (int batch = 0; batch * EPI_BATCH < n_iter; ++batch) {
    float4 acc[EPI_BATCH], aux[EPI_BATCH];
#pragma unroll
    for (int b = 0; b < EPI_BATCH; ++b) {
      const int it = batch * EPI_BATCH + b;
      const int f = f_base + it * 128 + et, row = f >> 4, c4 = f & 15;
      const int m = (it < n_iter) ? m0 + row : p.M, n = n0 + c4 * 4;
      const bool full = vec_ok && m < p.M && (n + 3 < p.N);
      acc[b] = (it < n_iter) ? *reinterpret_cast<const float4*>(stage + row * SLD + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      aux[b] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (full) {
        if (EPI == EPI_ACC) aux[b] = *reinterpret_cast<const float4*>(p.d + (size_t)m * p.ldd + n);
        if (EPI == EPI_MASK) aux[b] = *reinterpret_cast<const float4*>(p.mask_in + (size_t)m * p.ld_mask_in + n);
        if (EPI == EPI_BIAS_ACT && p.drop_u) aux[b] = *reinterpret_cast<const float4*>(p.drop_u + (size_t)m * p.ld_drop_u + n);
      }
    }
#pragma unroll
    for (int b = 0; b < EPI_BATCH; ++b) {
      const int it = batch * EPI_BATCH + b;
      const int f = f_base + it * 128 + et, row = f >> 4, c4 = f & 15;
      const int m = (it < n_iter) ? m0 + row : p.M, n = n0 + c4 * 4;
      if (m >= p.M || n >= p.N) continue;
      if (EPI == EPI_GENERIC || !(vec_ok && (n + 3 < p.N))) {
        epilogue_scalar4(p, m, n, acc[b]);
        continue;
      }
      float v0 = p.alpha * acc[b].x, v1 = p.alpha * acc[b].y, v2 = p.alpha * acc[b].z, v3 = p.alpha * acc[b].w;
      if (EPI == EPI_BIAS || EPI == EPI_BIAS_ACT) {
        const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias + n));
        v0 += bb.x; v1 += bb.y; v2 += bb.z; v3 += bb.w;
      }
      if (EPI == EPI_BIAS_ACT) {
        float4 u = aux[b];
        if (p.act == 2 && !p.drop_u) {
          const Philox4 r = ftg_dropout_block(p, m, n >> 2);
          u = make_float4(u01(r.x), u01(r.y), u01(r.z), u01(r.w));
        }
        float k0 = v0 > 0.f ? 1.f : p.slope, k1 = v1 > 0.f ? 1.f : p.slope, k2 = v2 > 0.f ? 1.f : p.slope,
              k3 = v3 > 0.f ? 1.f : p.slope;
        if (p.act == 2) {
          const float keep = 1.f / (1.f - p.drop_p);
          k0 = (u.x >= p.drop_p) ? k0 * keep : 0.f;
          k1 = (u.y >= p.drop_p) ? k1 * keep : 0.f;
          k2 = (u.z >= p.drop_p) ? k2 * keep : 0.f;
          k3 = (u.w >= p.drop_p) ? k3 * keep : 0.f;
        }
        if (p.mask_out) *reinterpret_cast<float4*>(p.mask_out + (size_t)m * p.ld_mask_out + n) = make_float4(k0, k1, k2, k3);
        v0 *= k0; v1 *= k1; v2 *= k2; v3 *= k3;
      }
      if (EPI == EPI_MASK) { v0 *= aux[b].x; v1 *= aux[b].y; v2 *= aux[b].z; v3 *= aux[b].w; }
      if (EPI == EPI_ACC) { v0 += aux[b].x; v1 += aux[b].y; v2 += aux[b].z; v3 += aux[b].w; }
      *reinterpret_cast<float4*>(p.d + (size_t)m * p.ldd + n) = make_float4(v0, v1, v2, v3);
    }
  }
}


// ---------------------------------------------------------------------------------------------
// EPI_BN: Linear -> BatchNorm1d(train) -> ReLU -> concat-store as the GEMM's epilogue (SURVEY K2;
// reference Residual, S/dtds/synthesizers/ctgan.py:33-44).
//
// The CTAs of one 64-column panel -- all M-tiles (cluster dim y = gridDim.y <= 4) times the split-K
// factor (cluster dim z) -- form one thread-block cluster.  After the split-K partials have been
// summed through distributed shared memory, CTA (y, z) owns rows [z*128/S, (z+1)*128/S) of M-tile y.
// It computes (count, mean, M2) of its rows per column (two local passes over its shared-memory
// slice), publishes them in shared memory, and after one cluster barrier every CTA merges the
// triplets of all cluster CTAs with Chan's parallel-variance formula (fixed order: deterministic,
// no E[x^2]-mean^2 cancellation).  Then each CTA normalises its own slice and stores BOTH the
// pre-activation (kept for the backward pass) and relu(bn(.)) into its column window of the wide
// activation buffer.  CTA (0, 0) of the cluster updates running_mean / running_var (momentum,
// unbiased variance) and the saved batch statistics.
constexpr int BN_STAT_OFF = BM * SLD * 4;            // bytes: red[2][4][64] floats, then stat[3][64] floats
constexpr int BN_RED_FLOATS = 2 * 4 * BN;
static_assert(BN_STAT_OFF + (BN_RED_FLOATS + 3 * BN) * 4 <= 2 * STAGE_BYTES, "BN scratch must fit the idle operand ring");

__device__ __forceinline__ float4 bn_colreduce(float4 v, float* red, int et) {
  // threads with equal (et & 15) own the same 4 columns; 8 row groups = 2 per warp x 4 warps
  v.x += __shfl_xor_sync(0xffffffffu, v.x, 16); v.y += __shfl_xor_sync(0xffffffffu, v.y, 16);
  v.z += __shfl_xor_sync(0xffffffffu, v.z, 16); v.w += __shfl_xor_sync(0xffffffffu, v.w, 16);
  const int w = et >> 5, c4 = et & 15;
  asm volatile("bar.sync 1, 128;" ::: "memory");               // previous use of `red` is over
  if ((et & 31) < 16) *reinterpret_cast<float4*>(red + w * BN + c4 * 4) = v;
  asm volatile("bar.sync 1, 128;" ::: "memory");
  float4 t = *reinterpret_cast<const float4*>(red + c4 * 4);
#pragma unroll
  for (int k = 1; k < 4; ++k) {
    const float4 u = *reinterpret_cast<const float4*>(red + k * BN + c4 * 4);
    t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
  }
  return t;
}

__device__ __forceinline__ void bn_epilogue(const ftg_gemm_args& p, const ftg_bn_args& bn, bool vec_ok,
                                            unsigned char* gen_base, uint32_t base_u32, int et, int m0, int n0,
                                            int r, int S, int CY, int y, float* stat_recv, uint32_t stat_recv_u32,
                                            uint32_t stat_bar) {
  float* stage = reinterpret_cast<float*>(gen_base);
  float* red = reinterpret_cast<float*>(gen_base + BN_STAT_OFF);
  float* stat = red + BN_RED_FLOATS;
  const int R = BM / S;                                // rows of the tile owned by this CTA
  const int c4 = et & 15, g = et >> 4;
  const int n = n0 + c4 * 4;
  const int row0 = r * R;
  const int nj = R >> 3;                               // rows per thread (stride 8)
  int cnt_i = p.M - (m0 + row0);
  cnt_i = cnt_i < 0 ? 0 : (cnt_i > R ? R : cnt_i);
  const float cnt = (float)cnt_i;
  // bias (column-wise shift: does not change the variance, but the stored pre-activation carries it)
  float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.bias) {
    bb.x = n + 0 < p.N ? __ldg(p.bias + n + 0) : 0.f; bb.y = n + 1 < p.N ? __ldg(p.bias + n + 1) : 0.f;
    bb.z = n + 2 < p.N ? __ldg(p.bias + n + 2) : 0.f; bb.w = n + 3 < p.N ? __ldg(p.bias + n + 3) : 0.f;
  }
  // ---- pass A: local column sums (and fold alpha / bias into the staged slice)
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int j = 0; j < nj; ++j) {
    const int row = row0 + g + 8 * j;
    float4* q = reinterpret_cast<float4*>(stage + row * SLD + c4 * 4);
    float4 v = *q;
    v.x = p.alpha * v.x + bb.x; v.y = p.alpha * v.y + bb.y; v.z = p.alpha * v.z + bb.z; v.w = p.alpha * v.w + bb.w;
    *q = v;
    if (m0 + row < p.M) { s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
  }
  s = bn_colreduce(s, red, et);
  const float icnt = cnt_i > 0 ? 1.f / cnt : 0.f;
  const float4 mean_loc = make_float4(s.x * icnt, s.y * icnt, s.z * icnt, s.w * icnt);
  // ---- pass B: local M2 = sum (v - mean_loc)^2
  float4 q2 = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int j = 0; j < nj; ++j) {
    const int row = row0 + g + 8 * j;
    if (m0 + row < p.M) {
      const float4 v = *reinterpret_cast<const float4*>(stage + row * SLD + c4 * 4);
      const float dx = v.x - mean_loc.x, dy = v.y - mean_loc.y, dz = v.z - mean_loc.z, dw = v.w - mean_loc.w;
      q2.x += dx * dx; q2.y += dy * dy; q2.z += dz * dz; q2.w += dw * dw;
    }
  }
  q2 = bn_colreduce(q2, red + 4 * BN, et);
  const int nct = CY * S;
  float4 mu = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 pm[8], pq[8];
  float pc[8];
  if (stat_recv != nullptr) {
    // ---- push: store my (count, mean, M2) into slot `me` of every cluster CTA's statistics buffer
    // (st.async signals that CTA's stat_bar), then wait for all nct slots of my own buffer
    const int me = y + CY * r;
    if (g == 0) {
      const float4 cc = make_float4(cnt, cnt, cnt, cnt);
      const uint32_t slot = stat_recv_u32 + (uint32_t)((me * 3 * BN + c4 * 4) * 4);
      for (int c = 0; c < nct; ++c) {
        const uint32_t rb = map_cluster(stat_bar, c);
        st_async_v4(map_cluster(slot, c), cc, rb);
        st_async_v4(map_cluster(slot + BN * 4, c), mean_loc, rb);
        st_async_v4(map_cluster(slot + 2 * BN * 4, c), q2, rb);
      }
    }
    mbar_wait(stat_bar, 0);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      pc[c] = 0.f; pm[c] = make_float4(0.f, 0.f, 0.f, 0.f); pq[c] = pm[c];
      if (c < nct) {
        const float* sl = stat_recv + c * 3 * BN + c4 * 4;
        pc[c] = sl[0];
        pm[c] = *reinterpret_cast<const float4*>(sl + BN);
        pq[c] = *reinterpret_cast<const float4*>(sl + 2 * BN);
        mu.x += pc[c] * pm[c].x; mu.y += pc[c] * pm[c].y; mu.z += pc[c] * pm[c].z; mu.w += pc[c] * pm[c].w;
      }
    }
  } else {
  if (g == 0) {
    *reinterpret_cast<float4*>(stat + 0 * BN + c4 * 4) = make_float4(cnt, cnt, cnt, cnt);
    *reinterpret_cast<float4*>(stat + 1 * BN + c4 * 4) = mean_loc;
    *reinterpret_cast<float4*>(stat + 2 * BN + c4 * 4) = q2;
  }
  // ---- publish, then merge the statistics of every CTA of the cluster (same order everywhere)
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  const uint32_t stat_u32 = base_u32 + BN_STAT_OFF + BN_RED_FLOATS * 4 + c4 * 16;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    pc[c] = 0.f; pm[c] = make_float4(0.f, 0.f, 0.f, 0.f); pq[c] = pm[c];
    if (c < nct) {
      uint32_t remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(stat_u32), "r"(c));
      float4 cc;
      asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(cc.x), "=f"(cc.y), "=f"(cc.z), "=f"(cc.w) : "r"(remote) : "memory");
      asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(pm[c].x), "=f"(pm[c].y), "=f"(pm[c].z), "=f"(pm[c].w) : "r"(remote + BN * 4) : "memory");
      asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(pq[c].x), "=f"(pq[c].y), "=f"(pq[c].z), "=f"(pq[c].w) : "r"(remote + 2 * BN * 4) : "memory");
      pc[c] = cc.x;
      mu.x += cc.x * pm[c].x; mu.y += cc.x * pm[c].y; mu.z += cc.x * pm[c].z; mu.w += cc.x * pm[c].w;
    }
  }
  }
  const float invM = 1.f / (float)p.M;
  mu.x *= invM; mu.y *= invM; mu.z *= invM; mu.w *= invM;
  float4 var = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    if (c < nct) {
      const float dx = pm[c].x - mu.x, dy = pm[c].y - mu.y, dz = pm[c].z - mu.z, dw = pm[c].w - mu.w;
      var.x += pq[c].x + pc[c] * dx * dx; var.y += pq[c].y + pc[c] * dy * dy;
      var.z += pq[c].z + pc[c] * dz * dz; var.w += pq[c].w + pc[c] * dw * dw;
    }
  }
  var.x *= invM; var.y *= invM; var.z *= invM; var.w *= invM;                 // biased (normalisation)
  const float4 inv = make_float4(rsqrtf(var.x + bn.eps), rsqrtf(var.y + bn.eps), rsqrtf(var.z + bn.eps), rsqrtf(var.w + bn.eps));
  float ga[4], be[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    ga[e] = n + e < p.N ? __ldg(bn.gamma + n + e) : 0.f;
    be[e] = n + e < p.N ? __ldg(bn.beta + n + e) : 0.f;
  }
  if (y == 0 && r == 0 && g == 0) {
    const float unb = (float)p.M / (float)(p.M > 1 ? p.M - 1 : 1);
    const float mus[4] = {mu.x, mu.y, mu.z, mu.w}, vars[4] = {var.x, var.y, var.z, var.w},
                invs[4] = {inv.x, inv.y, inv.z, inv.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (n + e < p.N) {
        bn.running_mean[n + e] = (1.f - bn.momentum) * bn.running_mean[n + e] + bn.momentum * mus[e];
        bn.running_var[n + e] = (1.f - bn.momentum) * bn.running_var[n + e] + bn.momentum * vars[e] * unb;
        if (bn.save_mean) { bn.save_mean[n + e] = mus[e]; bn.save_invstd[n + e] = invs[e]; }
      }
    }
    if (blockIdx.x == 0 && et == 0 && bn.nbt) bn.nbt[0] += 1.f;
  }
  // ---- normalise own slice: pre-activation -> p.d, relu(bn) -> bn.out (column window of the wide buffer)
  const bool vec = vec_ok && (n + 3 < p.N);
  for (int j = 0; j < nj; ++j) {
    const int row = row0 + g + 8 * j;
    const int m = m0 + row;
    if (m >= p.M || n >= p.N) continue;
    const float4 v = *reinterpret_cast<const float4*>(stage + row * SLD + c4 * 4);
    float4 o;
    o.x = fmaxf((v.x - mu.x) * inv.x * ga[0] + be[0], 0.f); o.y = fmaxf((v.y - mu.y) * inv.y * ga[1] + be[1], 0.f);
    o.z = fmaxf((v.z - mu.z) * inv.z * ga[2] + be[2], 0.f); o.w = fmaxf((v.w - mu.w) * inv.w * ga[3] + be[3], 0.f);
    if (vec) {
      *reinterpret_cast<float4*>(p.d + (size_t)m * p.ldd + n) = v;
      *reinterpret_cast<float4*>(bn.out + (size_t)m * bn.ldo + n) = o;
    } else {
      const float vs[4] = {v.x, v.y, v.z, v.w}, os[4] = {o.x, o.y, o.z, o.w};
      for (int e = 0; e < 4 && n + e < p.N; ++e) {
        p.d[(size_t)m * p.ldd + n + e] = vs[e];
        bn.out[(size_t)m * bn.ldo + n + e] = os[e];
      }
    }
  }
}

// Optional in-kernel timeline of CTA (0,0,0) (clock64 at phase boundaries) for latency analysis.
__device__ long long g_gemm_timeline[16];
__device__ int g_gemm_timeline_on = 0;
#define TL(i) do { if (g_gemm_timeline_on && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) g_gemm_timeline[i] = clock64(); } while (0)

// 112 registers: two (or three) CTAs fit one SM when the ring is shallow, so that on multi-wave
// grids one CTA's prologue / epilogue latency hides under its neighbours' main loops.
template <bool A_MN, bool B_MN, int EPI, bool X3>
__global__ void __maxnreg__(112)
k_gemm_tc(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
          const ftg_gemm_args p, int split_k, float* __restrict__ workspace, int* __restrict__ tile_counters,
          int vec_ok_i, int cluster_mode, int STAGES, const ftg_bn_args bn) {
  const bool vec_ok = vec_ok_i != 0;
  constexpr int STG = X3 ? 2 * STAGE_BYTES : STAGE_BYTES;   // fp32x3: [A_hi | B_hi | A_lo | B_lo] per stage
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;            // SWIZZLE_128B tiles need 1024 B alignment
  unsigned char* gen_base = smem_raw + (base - raw);
  const uint32_t bars = base + STAGES * STG;               // full[S], empty[S], tmem_full, tmem_slot, flag, split[S]
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gen_base + STAGES * STG + 8 * (2 * STAGES + 1));
  volatile int* last_flag = reinterpret_cast<volatile int*>(gen_base + STAGES * STG + 8 * (2 * STAGES + 1) + 8);
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  const uint32_t tmem_full_bar = bars + 8u * (2 * STAGES);
  auto split_bar = [&](int s) { return bars + 8u * (2 * STAGES + 3 + s); };   // hi/lo copies of stage s are ready
  // push-based cluster reduction (cluster_mode == 2): peers store their partial of MY slice into `recv`
  // (st.async) and signal recv_bar; BatchNorm statistics of all cluster CTAs land in `stat_recv` / stat_bar
  const uint32_t recv_bar = bars + 8u * (3 * STAGES + 3), stat_bar = bars + 8u * (3 * STAGES + 4);
  const uint32_t recv_u32 = bars + 256u, stat_recv_u32 = recv_u32 + RECV_BYTES;
  float* recv = reinterpret_cast<float*>(gen_base + STAGES * STG + 256);
  float* stat_recv = reinterpret_cast<float*>(gen_base + STAGES * STG + 256 + RECV_BYTES);
  const bool push = cluster_mode == 2;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM, z = blockIdx.z;
  const int nkb = (p.K + BK - 1) / BK;
  const int kb0 = (int)((long long)z * nkb / split_k), kb1 = (int)((long long)(z + 1) * nkb / split_k);

  // PDL: let the next kernel become resident now; our own prologue (barrier init, TMEM
  // allocation, descriptor prefetch) overlaps the predecessor's tail, then we wait for its data.
  asm volatile("griddepcontrol.launch_dependents;");
  if (threadIdx.x == 0) {
    TL(0);
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    if (X3) for (int s = 0; s < STAGES; ++s) mbar_init(split_bar(s), 4);      // one arrival per splitter warp
    mbar_init(tmem_full_bar, 1);
    if (push) {
      // one arrival (this thread) + the bytes the peers will push: (S-1) slices of BM/S rows x 64 floats,
      // and for the BatchNorm epilogue 768 B of statistics from each of the CY*S cluster CTAs (self included)
      mbar_init(recv_bar, 1);
      mbar_init(stat_bar, 1);
      mbar_expect_tx(recv_bar, (uint32_t)((split_k - 1) * (BM / split_k) * BN * 4));
      if (EPI == EPI_BN) mbar_expect_tx(stat_bar, (uint32_t)(split_k * (int)gridDim.y * 3 * BN * 4));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  // push protocol: tell the cluster that this CTA's receive barriers exist (non-blocking; the matching
  // wait sits right before the first remote store, a whole main loop later)
  // (.relaxed: fence.mbarrier_init.release.cluster above already publishes the initialised barriers)
  if (push) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");        // predecessor's outputs are visible from here on

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer
    int s = 0;
    uint32_t phase = 0;
    TL(1);
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(empty_bar(s), phase ^ 1u);
      const uint32_t sa = base + s * STG, sb = sa + A_STAGE_BYTES;
      mbar_expect_tx(full_bar(s), STAGE_BYTES);
      const int k = kb * BK;
      if (A_MN) {
#pragma unroll
        for (int c = 0; c < BM / 32; ++c) tma_load_2d(sa + c * (BK * 128), &map_a, full_bar(s), m0 + c * 32, k);
      } else {
        tma_load_2d(sa, &map_a, full_bar(s), k, m0);
      }
      if (B_MN) {
#pragma unroll
        for (int c = 0; c < BN / 32; ++c) tma_load_2d(sb + c * (BK * 128), &map_b, full_bar(s), n0 + c * 32, k);
      } else {
        tma_load_2d(sb, &map_b, full_bar(s), k, n0);
      }
      if (++s == STAGES) { s = 0; phase ^= 1u; }
    }
    TL(2);
  } else if (warp == 1 && lane == 0) {
    // -------------------------------------------------------------- MMA issuer
    // instruction descriptor: D=f32, A=B=tf32, majors, N=64, M=128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    int s = 0;
    uint32_t phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(X3 ? split_bar(s) : full_bar(s), phase);
      if (kb == kb0) TL(3);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t sa = base + s * STG, sb = sa + A_STAGE_BYTES;
#pragma unroll
      for (int ks = 0; ks < BK / 8; ++ks) {
        // K-major : 8 tf32 = 32 B step inside the 128 B swizzle row; SBO = 8 rows * 128 B
        // MN-major: one MMA (K = 8) spans two 4-row swizzle atoms (SBO = 512 B apart); the next
        //           32-element MN chunk is the next TMA box (LBO = BK * 128 B); 8 k-rows = 1024 B step
        uint64_t ad = A_MN ? make_desc(sa + ks * 1024, BK * 128, 512, 1) : make_desc(sa + ks * 32, 16, 1024, 2);
        uint64_t bd = B_MN ? make_desc(sb + ks * 1024, BK * 128, 512, 1) : make_desc(sb + ks * 32, 16, 1024, 2);
        const uint32_t acc0 = (kb > kb0 || ks > 0) ? 1u : 0u;
        if (X3) {
          // lo copies sit STAGE_BYTES above their hi twins; small cross terms first
          const uint32_t la = sa + STAGE_BYTES, lb = sb + STAGE_BYTES;
          uint64_t adl = A_MN ? make_desc(la + ks * 1024, BK * 128, 512, 1) : make_desc(la + ks * 32, 16, 1024, 2);
          uint64_t bdl = B_MN ? make_desc(lb + ks * 1024, BK * 128, 512, 1) : make_desc(lb + ks * 32, 16, 1024, 2);
          umma_tf32(tmem_base, adl, bd, idesc, acc0);
          umma_tf32(tmem_base, ad, bdl, idesc, 1u);
          umma_tf32(tmem_base, ad, bd, idesc, 1u);
        } else {
          umma_tf32(tmem_base, ad, bd, idesc, acc0);
        }
      }
      umma_commit(empty_bar(s));                // frees the smem slot when these MMAs retire
      if (++s == STAGES) { s = 0; phase ^= 1u; }
    }
    umma_commit(tmem_full_bar);                 // accumulator complete
    TL(4);
  } else if (warp >= 2) {
    // ---------------------------------------------------------------- epilogue
    const int q = warp & 3;                      // TMEM lane quadrant owned by this warp
    const int row_in_tile = q * 32 + lane;
    const int m = m0 + row_in_tile;
    if (push) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");   // peers' receive barriers exist
    if (X3) {
      // ---- fp32x3 splitter: as each stage lands, rewrite it as hi (in place) + lo (upper half of the stage)
      const int st = threadIdx.x - 64;
      int s = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar(s), phase);
        float4* hi4 = reinterpret_cast<float4*>(gen_base + s * STG);
        float4* lo4 = reinterpret_cast<float4*>(gen_base + s * STG + STAGE_BYTES);
        constexpr int PER = STAGE_BYTES / 16 / 128;        // 12 float4 per thread
        float4 v[PER];
#pragma unroll
        for (int i = 0; i < PER; ++i) v[i] = hi4[i * 128 + st];
        if (p.x3 == 2) {
          // The tensor core reads an fp32 word as tf32 by dropping its 13 low mantissa bits, so the
          // stage as TMA wrote it already IS the hi operand; only lo = rna_tf32(x - trunc_tf32(x))
          // has to be written (one third less shared-memory traffic per stage).
#pragma unroll
          for (int i = 0; i < PER; ++i) {
            float4 l;
            l.x = lo_trunc_tf32(v[i].x); l.y = lo_trunc_tf32(v[i].y);
            l.z = lo_trunc_tf32(v[i].z); l.w = lo_trunc_tf32(v[i].w);
            lo4[i * 128 + st] = l;
          }
        } else {
#pragma unroll
          for (int i = 0; i < PER; ++i) {
            float4 h, l;
            split_tf32(v[i].x, h.x, l.x); split_tf32(v[i].y, h.y, l.y);
            split_tf32(v[i].z, h.z, l.z); split_tf32(v[i].w, h.w, l.w);
            hi4[i * 128 + st] = h;
            lo4[i * 128 + st] = l;
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
        __syncwarp();
        if (lane == 0) mbar_arrive(split_bar(s));
        if (++s == STAGES) { s = 0; phase ^= 1u; }
      }
    }
    mbar_wait(tmem_full_bar, 0);
    if (threadIdx.x == 64) TL(5);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t r[64];
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    tmem_ld32(taddr, r);
    tmem_ld32(taddr + 32, r + 32);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    // Stage the accumulator tile through shared memory (the operand ring is idle once the
    // last MMA has retired): afterwards thread `et` owns float4 chunks f = it*128 + et of the
    // row-major tile, so all global traffic is 16 B per lane and fully coalesced, and the
    // epilogue's loads (masks, uniforms, accumulate) are batched 8 deep for latency hiding.
    float* stage = reinterpret_cast<float*>(gen_base);
    {
      float4* myrow = reinterpret_cast<float4*>(stage + row_in_tile * SLD);
#pragma unroll
      for (int j = 0; j < BN / 4; ++j)
        myrow[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                               __uint_as_float(r[4 * j + 3]));
    }
    const int et = threadIdx.x - 64;                           // 0..127 among the epilogue threads
    asm volatile("bar.sync 1, 128;" ::: "memory");             // whole tile staged
    if (threadIdx.x == 64) TL(6);
    bool do_epilogue = !cluster_mode;
    if (split_k > 1 && !cluster_mode) {
      // ---- split-K: publish the partial tile, last-arriving CTA reduces all partials
      const int tile = blockIdx.y * gridDim.x + blockIdx.x;
      const int n_tiles = gridDim.x * gridDim.y;
      float4* ws4 = reinterpret_cast<float4*>(workspace + ((size_t)z * n_tiles + tile) * (BM * BN));
#pragma unroll
      for (int it = 0; it < CHUNKS; ++it) {
        const int f = it * 128 + et;
        ws4[f] = *reinterpret_cast<const float4*>(stage + (f >> 4) * SLD + (f & 15) * 4);
      }
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (et == 0) {
        int old = atomicAdd(&tile_counters[tile], 1);
        *last_flag = (old == split_k - 1) ? 1 : 0;
        if (old == split_k - 1) tile_counters[tile] = 0;       // re-arm for the next launch
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      do_epilogue = (*last_flag != 0);
      if (do_epilogue) {
        __threadfence();
        const size_t zstride4 = (size_t)n_tiles * (BM * BN / 4);
        const float4* wst4 = reinterpret_cast<const float4*>(workspace + (size_t)tile * (BM * BN));
        float4 acc[CHUNKS];
#pragma unroll
        for (int it = 0; it < CHUNKS; ++it) acc[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int zz = 0; zz < split_k; ++zz) {                 // fixed order => deterministic
#pragma unroll
          for (int it = 0; it < CHUNKS; ++it) {                // 16 independent 16 B loads in flight
            const float4 v = __ldcg(wst4 + zz * zstride4 + it * 128 + et);
            acc[it].x += v.x; acc[it].y += v.y; acc[it].z += v.z; acc[it].w += v.w;
          }
        }
#pragma unroll
        for (int it = 0; it < CHUNKS; ++it) {
          const int f = it * 128 + et;
          *reinterpret_cast<float4*>(stage + (f >> 4) * SLD + (f & 15) * 4) = acc[it];   // own chunks only
        }
      }
    }
    if (threadIdx.x == 64) TL(7);
    if (do_epilogue) {
      epilogue_tile<EPI>(p, vec_ok, stage, et, m0, n0, 0, CHUNKS);
    }
    if (threadIdx.x == 64) TL(8);
    (void)m;
  }
  if (push) {
    // ---- split-K inside a thread-block cluster, PUSH protocol (FTG_CLUSTER_PUSH=1; alternative to pull).  Every CTA has staged its
    // partial tile in its own shared memory.  CTA r owns rows [r*BM/S, (r+1)*BM/S) of the tile: each
    // CTA stores the other ranks' slices of its partial straight into the owners' receive buffers with
    // st.async (the store signals the owner's mbarrier when it lands), then waits on its OWN barrier
    // for the S-1 inbound slices, sums them in rank order (deterministic) and runs the fused epilogue
    // on its slice.  The data hand-off needs no cluster-wide barrier and no remote loads: one one-way
    // DSMEM trip instead of barrier + round trip + barrier (see profiles/gemm_timeline.json).
    const int S = split_k, r = z;
    const int CY = (EPI == EPI_BN) ? (int)gridDim.y : 1;    // cluster = (1, CY, S); ctarank(y', z') = y' + CY * z'
    const int ybase = (EPI == EPI_BN) ? (int)blockIdx.y : 0;
    // (phase-1 wait -- "every peer's receive barriers exist" -- of the epilogue warps sits at the start of their
    // role, hidden under the main loop; the producer / issuer warps take theirs here, after their work)
    if (warp < 2) {
      __syncwarp();
      asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    if (threadIdx.x == 64) TL(10);
    if (warp >= 2) {
      const int et = threadIdx.x - 64;
      float* stage = reinterpret_cast<float*>(gen_base);
      const int log2S = 31 - __clz(S);
      const int cpr = (BM * BN / 4) / S;                   // float4 chunks per rank slice
      const int R = BM / S;
      // push: chunk t of the other ranks' slices (t enumerates (owner o != r, j))
      if (S > 1) {
#pragma unroll
        for (int t = 0; t < CHUNKS; ++t) {
          const int j = t >> log2S, o = t & (S - 1);
          if (o == r) continue;
          const int f = o * cpr + j * 128 + et;
          const int row = f >> 4, c4 = f & 15;
          const float4 v = *reinterpret_cast<const float4*>(stage + row * SLD + c4 * 4);
          const uint32_t dst_local = recv_u32 + (uint32_t)(((r * R + (row - o * R)) * BN + c4 * 4) * 4);
          st_async_v4(map_cluster(dst_local, ybase + CY * o), v, map_cluster(recv_bar, ybase + CY * o));
        }
      }
      if (S > 1) mbar_wait(recv_bar, 0);                                       // my inbound slices have landed
      if (threadIdx.x == 64) TL(11);
      // own slice: sum the S partials in rank order
#pragma unroll
      for (int j = 0; j < CHUNKS; ++j) {
        if (j < (CHUNKS >> log2S)) {
          const int f = r * cpr + j * 128 + et;
          const int row = f >> 4, c4 = f & 15;
          float4* own = reinterpret_cast<float4*>(stage + row * SLD + c4 * 4);
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int q = 0; q < S; ++q) {
            const float4 v = (q == r) ? *own
                                      : *reinterpret_cast<const float4*>(recv + (q * R + (row - r * R)) * BN + c4 * 4);
            if (q == 0) acc = v;
            else { acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
          }
          *own = acc;
        }
      }
      if (EPI == EPI_BN) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
        bn_epilogue(p, bn, vec_ok, gen_base, base, et, m0, n0, r, S, CY, (int)blockIdx.y, stat_recv, stat_recv_u32, stat_bar);
      } else {
        epilogue_tile<EPI>(p, vec_ok, stage, et, m0, n0, r * cpr, CHUNKS / S);
      }
      if (threadIdx.x == 64) TL(12);
    }
    // nobody leaves while a peer may still have stores to this CTA un-issued (arrive only after the epilogue:
    // a release-arrive in the middle would stall on the outbound stores)
    __syncwarp();
    asm volatile("barrier.cluster.arrive.relaxed.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
  } else if (cluster_mode) {
    // ---- split-K inside a thread-block cluster, PULL protocol (default):
    // every CTA has staged its partial tile in its own shared memory; after a cluster barrier
    // CTA r sums slice r of all partials through distributed shared memory (fixed rank order:
    // deterministic), applies the fused epilogue to that slice and stores it.  No global
    // workspace, no atomics, no memory fences.
    __syncwarp();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (threadIdx.x == 64) TL(10);
    if (warp >= 2) {
      const int et = threadIdx.x - 64;
      float* stage = reinterpret_cast<float*>(gen_base);
      const uint32_t stage_u32 = base;
      const int S = split_k, r = z;                        // split-K rank inside the cluster
      // cluster = (1, CY, S): CY = all M-tiles of the panel for the BatchNorm epilogue, else 1;
      // %cluster_ctarank of CTA (y', z') is y' + CY * z'
      const int CY = (EPI == EPI_BN) ? (int)gridDim.y : 1;
      const int ybase = (EPI == EPI_BN) ? (int)blockIdx.y : 0;
      const int log2S = 31 - __clz(S);
      const int cpr = (BM * BN / 4) / S;                   // chunks per rank
      float4 v[CHUNKS];
#pragma unroll
      for (int t = 0; t < CHUNKS; ++t) {
        const int j = t >> log2S, q = t & (S - 1);
        const int f = r * cpr + j * 128 + et;
        const uint32_t local = stage_u32 + (uint32_t)(((f >> 4) * SLD + (f & 15) * 4) * 4);
        v[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m0 + (f >> 4) < p.M) {                           // rows past M carry nothing: skip their DSMEM traffic
          uint32_t remote;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(ybase + CY * q));
          asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(v[t].x), "=f"(v[t].y), "=f"(v[t].z), "=f"(v[t].w) : "r"(remote) : "memory");
        }
      }
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int t = 0; t < CHUNKS; ++t) {
        const int j = t >> log2S, q = t & (S - 1);
        if (q == 0) acc = v[t];
        else { acc.x += v[t].x; acc.y += v[t].y; acc.z += v[t].z; acc.w += v[t].w; }
        if (q == S - 1) {
          const int f = r * cpr + j * 128 + et;
          *reinterpret_cast<float4*>(stage + (f >> 4) * SLD + (f & 15) * 4) = acc;     // own slice, own chunks
        }
      }
      if (threadIdx.x == 64) TL(11);
      if (EPI == EPI_BN) {
        asm volatile("bar.sync 1, 128;" ::: "memory");     // own slice fully reduced in shared memory
        bn_epilogue(p, bn, vec_ok, gen_base, base, et, m0, n0, r, S, CY, (int)blockIdx.y, nullptr, 0u, 0u);
      } else {
        epilogue_tile<EPI>(p, vec_ok, stage, et, m0, n0, r * cpr, CHUNKS / S);
      }
      if (threadIdx.x == 64) TL(12);
    } else if (EPI == EPI_BN) {
      // warps 0/1 take part in the statistics exchange barrier of bn_epilogue (.aligned: all threads)
      __syncwarp();
      asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    // peers may still be reading this CTA's shared memory: nobody leaves before everybody is done
    __syncwarp();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) TL(9);
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ------------------------------------------------------------------ persistent variant
// One CTA per SM walks the output tiles (m fastest, so neighbouring CTAs share a B panel in L2).  The
// TMA ring runs ahead across tile boundaries, the accumulator is double-buffered in TMEM (2 x 64
// columns): while the epilogue warps drain tile i from one half, the MMA issuer already fills the other
// half with tile i+1, and barrier init / TMEM allocation / descriptor prefetch are paid once per CTA
// instead of once per tile.  No split-K (multi-wave grids only).  The staged epilogue tile has its own
// shared-memory area here because the operand ring never goes idle.
// EXPERIMENTAL: opt-in (FTG_GEMM_PERSISTENT=1 / ftg_gemm_tc_set_persistent); the default launch path
// does not use it.
constexpr int PERSIST_STAGES = 5;
constexpr int PERSIST_STAGE_TILE_BYTES = BM * SLD * 4;                       // 34 816 B
constexpr int PERSIST_SMEM_BYTES = PERSIST_STAGES * STAGE_BYTES + PERSIST_STAGE_TILE_BYTES + 1024 + 256;

template <bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(NUM_THREADS, 1)
k_gemm_tc_persist(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const ftg_gemm_args p, int vec_ok_i, int tiles_m, int tiles_n) {
  constexpr int S = PERSIST_STAGES;
  const bool vec_ok = vec_ok_i != 0;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* gen_base = smem_raw + (base - raw);
  float* stage = reinterpret_cast<float*>(gen_base + S * STAGE_BYTES);         // dedicated epilogue staging tile
  const uint32_t bars = base + S * STAGE_BYTES + PERSIST_STAGE_TILE_BYTES;
  // full[S], empty[S], tmem_full[2], tmem_empty[2], tmem_slot
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (S + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * S + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * S + 2 + a); };
  volatile uint32_t* tmem_slot =
      reinterpret_cast<volatile uint32_t*>(gen_base + S * STAGE_BYTES + PERSIST_STAGE_TILE_BYTES + 8 * (2 * S + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = (p.K + BK - 1) / BK;
  const int n_tiles = tiles_m * tiles_n;

  asm volatile("griddepcontrol.launch_dependents;");
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }   // 4 epilogue warps
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(2 * TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer (runs ahead over tiles)
    int s = 0;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      const int m0 = (t % tiles_m) * BM, n0 = (t / tiles_m) * BN;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(empty_bar(s), phase ^ 1u);
        const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_STAGE_BYTES;
        mbar_expect_tx(full_bar(s), STAGE_BYTES);
        const int k = kb * BK;
        if (A_MN) {
#pragma unroll
          for (int c = 0; c < BM / 32; ++c) tma_load_2d(sa + c * (BK * 128), &map_a, full_bar(s), m0 + c * 32, k);
        } else {
          tma_load_2d(sa, &map_a, full_bar(s), k, m0);
        }
        if (B_MN) {
#pragma unroll
          for (int c = 0; c < BN / 32; ++c) tma_load_2d(sb + c * (BK * 128), &map_b, full_bar(s), n0 + c * 32, k);
        } else {
          tma_load_2d(sb, &map_b, full_bar(s), k, n0);
        }
        if (++s == S) { s = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // -------------------------------------------------------------- MMA issuer
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    int s = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (uint32_t)(it >> 1) & 1u;
      mbar_wait(tempty_bar(acc), acc_phase ^ 1u);            // epilogue has drained this accumulator half
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TMEM_COLS);
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(full_bar(s), phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_STAGE_BYTES;
#pragma unroll
        for (int ks = 0; ks < BK / 8; ++ks) {
          uint64_t ad = A_MN ? make_desc(sa + ks * 1024, BK * 128, 512, 1) : make_desc(sa + ks * 32, 16, 1024, 2);
          uint64_t bd = B_MN ? make_desc(sb + ks * 1024, BK * 128, 512, 1) : make_desc(sb + ks * 32, 16, 1024, 2);
          umma_tf32(d_tmem, ad, bd, idesc, (kb > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit(empty_bar(s));
        if (++s == S) { s = 0; phase ^= 1u; }
      }
      umma_commit(tfull_bar(acc));
    }
  } else if (warp >= 2) {
    // ---------------------------------------------------------------- epilogue warps
    const int q = warp & 3;
    const int row_in_tile = q * 32 + lane;
    const int et = threadIdx.x - 64;
    int it = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
      const int m0 = (t % tiles_m) * BM, n0 = (t / tiles_m) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (uint32_t)(it >> 1) & 1u;
      mbar_wait(tfull_bar(acc), acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint32_t r[64];
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * TMEM_COLS);
      tmem_ld32(taddr, r);
      tmem_ld32(taddr + 32, r + 32);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      // the accumulator half is in registers: hand it back to the MMA issuer before the slow part
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      asm volatile("bar.sync 1, 128;" ::: "memory");           // previous tile's staged data fully consumed
      {
        float4* myrow = reinterpret_cast<float4*>(stage + row_in_tile * SLD);
#pragma unroll
        for (int j = 0; j < BN / 4; ++j)
          myrow[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                 __uint_as_float(r[4 * j + 3]));
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");           // whole tile staged
      epilogue_tile<EPI>(p, vec_ok, stage, et, m0, n0, 0, CHUNKS);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * TMEM_COLS) : "memory");
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// L2 promotion of the TMA requests.  A box row is 128 B; the next 128 B of the same row is the next
// k-block (K-major) or the next 32-row chunk of the same stage (MN-major), so promoting a miss to
// 256 B turns every other request into an L2 hit and doubles the DRAM burst per opened page.
// FTG_TMA_L2_PROMO=128|256 overrides (experiments).
CUtensorMapL2promotion l2_promotion() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FTG_TMA_L2_PROMO");
    v = e ? atoi(e) : 128;
  }
  return v == 256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
}

// 2-D fp32 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld floats.
bool make_map(CUtensorMap* map, const float* ptr, int inner, int outer, int ld, int box_inner, int box_outer,
              bool mn_major) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   l2_promotion(),
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// Ring depth (measured: tools/gemm_stage_sweep.py, profiles/gemm_stage_sweep.json).
//  * A grid that fits the machine in one wave gets the deepest ring its k-loop can use: every k-block
//    in flight at once (the CTGAN iteration is latency bound).
//  * Multi-wave grids get 4 stages (99 KB) so that two CTAs are resident per SM and the SM overlaps
//    one tile's prologue / epilogue with its neighbour's main loop: wgrad of the wide first
//    discriminator layer (3936 tiles of 5 k-blocks) 165 -> 102 us, its dgrad (492 tiles) 70 -> 50 us.
//    Three CTAs per SM (3 stages) is no better than two.
//  * Clusters of 8 with one CTA per SM need 8 free SMs of one GPC each; 16 such clusters do not all
//    fit, so a 128-CTA split-8 grid ran in two waves (100 us); with 4 stages two CTAs of a cluster
//    share an SM and the grid is resident at once (67 us).
// FTG_GEMM_STAGES=<n> / ftg_gemm_tc_set_stages force a depth (experiments).
int g_forced_stages = -1;
int choose_stages(long long ctas, int kblocks_per_cta, int cluster_size, bool x3 = false) {
  if (g_forced_stages < 0) {
    const char* e = getenv("FTG_GEMM_STAGES");
    g_forced_stages = e ? atoi(e) : 0;
  }
  const int forced = g_forced_stages;
  int s = MAX_STAGES;
  if (forced >= 2 && forced <= MAX_STAGES) s = forced;
  else if (ctas > 148) s = 4;
  else if (cluster_size >= 8 && ctas > 96) s = 4;
  if (x3) {
    // a stage is 48 KB (hi + lo): 4 stages fill the SM; multi-wave grids / big clusters keep two CTAs per SM
    if (s > MAX_STAGES_X3) s = MAX_STAGES_X3;
    if (!(forced >= 2 && forced <= MAX_STAGES) && (ctas > 148 || (cluster_size >= 8 && ctas > 96))) s = 2;
  }
  if (s > kblocks_per_cta) s = kblocks_per_cta;
  return s < 2 ? 2 : s;
}

// Cluster split-K reduction protocol: pull (remote loads between two cluster barriers, default) or push
// (st.async into the owner's receive buffer, FTG_CLUSTER_PUSH=1).  Measured on B200 (profiles/gemm_cluster_protocols.md):
// a cluster barrier costs 1-2 k cycles and moving a 32 KB tile through DSMEM ~2 k cycles either way, so the two
// protocols finish within 2 % of each other per GEMM and the whole iteration is not faster with push.
int g_cluster_push = -1;
bool use_push() {
  if (g_cluster_push < 0) {
    const char* e = getenv("FTG_CLUSTER_PUSH");
    g_cluster_push = (e && atoi(e) != 0) ? 1 : 0;
  }
  return g_cluster_push == 1;
}

// Persistent-tile variant: opt-in (experimental).
int g_persistent = -1;
bool use_persistent() {
  if (g_persistent < 0) {
    const char* e = getenv("FTG_GEMM_PERSISTENT");
    g_persistent = (e && atoi(e) != 0) ? 1 : 0;
  }
  return g_persistent == 1;
}

template <bool A_MN, bool B_MN, int EPI, bool X3 = false>
int launch_epi(const CUtensorMap& ma, const CUtensorMap& mb, const ftg_gemm_args& p, int split_k, float* ws, int* ctr,
               int vec_ok, cudaStream_t stream, const ftg_bn_args* bnp = nullptr) {
  if (!X3 && p.x3) return launch_epi<A_MN, B_MN, EPI, true>(ma, mb, p, split_k, ws, ctr, vec_ok, stream, bnp);
  // no workspace given: reduce through DSMEM; the BatchNorm epilogue always runs as a cluster
  // 2 = push protocol (st.async into the slice owner's shared memory, default), 1 = pull (FTG_CLUSTER_PUSH=0)
  const int cluster_mode = ((ws == nullptr && split_k > 1) || EPI == EPI_BN) ? (use_push() ? 2 : 1) : 0;
  ftg_bn_args bn = {};
  if (bnp) bn = *bnp;
  // function attributes are per device: one flag per (instantiation, device)
  static std::atomic<bool> attr_done_dev[64];
  int dev_id = 0;
  if (cudaGetDevice(&dev_id) != cudaSuccess || dev_id < 0 || dev_id >= 64) return FTG_ERR_UNSUPPORTED;
  std::atomic<bool>& attr_done = attr_done_dev[dev_id];
  if (!attr_done.load(std::memory_order_acquire)) {
    cudaError_t e = cudaFuncSetAttribute(k_gemm_tc<A_MN, B_MN, EPI, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         X3 ? (smem_bytes(MAX_STAGES_X3, true) > smem_bytes(MAX_STAGES_X3 - 1, true, true)
                                                   ? smem_bytes(MAX_STAGES_X3, true)
                                                   : smem_bytes(MAX_STAGES_X3 - 1, true, true))
                                            : smem_bytes(MAX_STAGES, false, true));
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(k_gemm_tc<A_MN, B_MN, EPI, X3>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return (int)e;
    attr_done.store(true, std::memory_order_release);
  }
  dim3 grid(ftg_cdiv(p.N, BN), ftg_cdiv(p.M, BM), split_k);
  if (!X3 && EPI != EPI_BN && use_persistent() && split_k == 1 && (long long)grid.x * grid.y > 148) {
    static std::atomic<bool> pattr_done_dev[64];
    if (!pattr_done_dev[dev_id].load(std::memory_order_acquire)) {
      cudaError_t e = cudaFuncSetAttribute(k_gemm_tc_persist<A_MN, B_MN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           PERSIST_SMEM_BYTES);
      if (e != cudaSuccess) return (int)e;
      pattr_done_dev[dev_id].store(true, std::memory_order_release);
    }
    ++g_ftg_launch_count;
    cudaLaunchConfig_t pc = {};
    pc.gridDim = dim3(148);
    pc.blockDim = dim3(NUM_THREADS);
    pc.dynamicSmemBytes = PERSIST_SMEM_BYTES;
    pc.stream = stream;
    cudaLaunchAttribute pattr[1];
    pattr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pattr[0].val.programmaticStreamSerializationAllowed = 1;
    pc.attrs = pattr;
    pc.numAttrs = g_ftg_pdl ? 1 : 0;
    return (int)cudaLaunchKernelEx(&pc, k_gemm_tc_persist<A_MN, B_MN, EPI>, ma, mb, p, vec_ok, (int)grid.y, (int)grid.x);
  }
  const int cluster_ctas = cluster_mode ? split_k * (EPI == EPI_BN ? (int)grid.y : 1) : 1;
  int stages = choose_stages((long long)grid.x * grid.y * grid.z, (ftg_cdiv(p.K, BK) + split_k - 1) / split_k,
                             cluster_ctas, X3);
  if (X3 && cluster_mode == 2 && stages > MAX_STAGES_X3 - 1) stages = MAX_STAGES_X3 - 1;   // room for the receive buffers
  ++g_ftg_launch_count;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = smem_bytes(stages, X3, cluster_mode == 2);
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (cluster_mode) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 1;
    attr[na].val.clusterDim.y = (EPI == EPI_BN) ? grid.y : 1;
    attr[na].val.clusterDim.z = (unsigned)split_k;
    ++na;
  }
  if (g_ftg_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return (int)cudaLaunchKernelEx(&cfg, k_gemm_tc<A_MN, B_MN, EPI, X3>, ma, mb, p, split_k, ws, ctr, vec_ok,
                                 cluster_mode, stages, bn);
}

template <bool A_MN, bool B_MN>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const ftg_gemm_args& p, int split_k, float* ws, int* ctr,
           cudaStream_t stream, const ftg_bn_args* bn = nullptr) {
  auto al = [](const void* ptr, int ld) { return ptr == nullptr || (((reinterpret_cast<uintptr_t>(ptr) & 15) == 0) && (ld & 3) == 0); };
  const int vec_ok = al(p.d, p.ldd) && al(p.mask_in, p.ld_mask_in) && al(p.mask_out, p.ld_mask_out) &&
                     al(p.drop_u, p.ld_drop_u) && al(p.bias, 0);
  if (bn) {
    const int ok = al(p.d, p.ldd) && al(bn->out, bn->ldo) && al(p.bias, 0) && al(bn->gamma, 0) && al(bn->beta, 0);
    return launch_epi<A_MN, B_MN, EPI_BN>(ma, mb, p, split_k, nullptr, nullptr, ok, stream, bn);
  }
  // pick the epilogue specialisation that matches the flag combination exactly
  int epi = EPI_GENERIC;
  const bool b = p.bias != nullptr, a = p.act != 0, mi = p.mask_in != nullptr, acc = p.accumulate != 0;
  if (!b && !a && !mi && !acc) epi = EPI_PLAIN;
  else if (!b && !a && !mi && acc) epi = EPI_ACC;
  else if (b && !a && !mi && !acc) epi = EPI_BIAS;
  else if (b && a && !mi && !acc) epi = EPI_BIAS_ACT;
  else if (!b && !a && mi && !acc) epi = EPI_MASK;
  switch (epi) {
    case EPI_PLAIN: return launch_epi<A_MN, B_MN, EPI_PLAIN>(ma, mb, p, split_k, ws, ctr, vec_ok, stream);
    case EPI_ACC: return launch_epi<A_MN, B_MN, EPI_ACC>(ma, mb, p, split_k, ws, ctr, vec_ok, stream);
    case EPI_BIAS: return launch_epi<A_MN, B_MN, EPI_BIAS>(ma, mb, p, split_k, ws, ctr, vec_ok, stream);
    case EPI_BIAS_ACT: return launch_epi<A_MN, B_MN, EPI_BIAS_ACT>(ma, mb, p, split_k, ws, ctr, vec_ok, stream);
    case EPI_MASK: return launch_epi<A_MN, B_MN, EPI_MASK>(ma, mb, p, split_k, ws, ctr, vec_ok, stream);
    default: return launch_epi<A_MN, B_MN, EPI_GENERIC>(ma, mb, p, split_k, ws, ctr, vec_ok, stream);
  }
}

}  // namespace

// debug: enable / read the in-kernel timeline of CTA (0,0,0)
extern "C" int ftg_gemm_tc_timeline(int enable, long long* out16) {
  cudaError_t e = cudaMemcpyToSymbol(g_gemm_timeline_on, &enable, sizeof(int));
  if (e != cudaSuccess) return (int)e;
  if (out16) e = cudaMemcpyFromSymbol(out16, g_gemm_timeline, 16 * sizeof(long long));
  return (int)e;
}

extern "C" int ftg_gemm_tc_num_tiles(int M, int N) { return ftg_cdiv(M, BM) * ftg_cdiv(N, BN); }

extern "C" long long ftg_gemm_tc_workspace_bytes(int M, int N, int split_k) {
  if (split_k <= 1) return 0;
  return (long long)split_k * ftg_gemm_tc_num_tiles(M, N) * BM * BN * 4;
}

// Heuristic: split K until the grid covers ~1 wave of 148 SMs, keeping >= 4 k-blocks per CTA.
extern "C" int ftg_gemm_tc_choose_split(int M, int N, int K) {
  int tiles = ftg_gemm_tc_num_tiles(M, N);
  int nkb = ftg_cdiv(K, BK);
  int s = 148 / (tiles > 0 ? tiles : 1);
  int cap = nkb / 4;
  if (s > cap) s = cap;
  if (s > 32) s = 32;
  if (s < 1) s = 1;
  return s;
}

// Cluster split-K factor: largest power of two <= 8 that keeps >= 2 k-blocks per CTA and the
// grid within one wave of 148 SMs.
// Force the ring depth of subsequent launches (2..6; 0 = automatic).  Tuning / tests only.
extern "C" int ftg_gemm_tc_set_stages(int stages) {
  g_forced_stages = (stages >= 2 && stages <= MAX_STAGES) ? stages : 0;
  return 0;
}

// 1: multi-wave unsplit GEMMs use the persistent double-buffered-TMEM kernel (experimental); 0: never.
extern "C" int ftg_gemm_tc_set_persistent(int on) {
  g_persistent = on ? 1 : 0;
  return 0;
}

// 1: push protocol for the in-cluster split-K reduction, 0: pull protocol (default).
extern "C" int ftg_gemm_tc_set_cluster_push(int on) {
  g_cluster_push = on ? 1 : 0;
  return 0;
}

extern "C" int ftg_gemm_tc_choose_cluster_split(int M, int N, int K) {
  int tiles = ftg_gemm_tc_num_tiles(M, N);
  int nkb = ftg_cdiv(K, BK);
  int s = 1;
  while (s * 2 <= 8 && tiles * s * 2 <= 148 && nkb / (s * 2) >= 2) s *= 2;
  return s;
}

extern "C" int ftg_gemm_tc(const ftg_gemm_args* args, int split_k, float* workspace, int* tile_counters,
                           ftg_stream_t stream) {
  const ftg_gemm_args& p = *args;
  if (p.M <= 0 || p.N <= 0) return 0;
  if (p.K <= 0) return FTG_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(p.a) & 15) || (reinterpret_cast<uintptr_t>(p.b) & 15) || (p.lda & 3) || (p.ldb & 3))
    return FTG_ERR_UNSUPPORTED;
  const int nkb = ftg_cdiv(p.K, BK);
  if (split_k < 1) split_k = 1;
  if (split_k > nkb) split_k = nkb;
  if (split_k > 1 && (!workspace || !tile_counters)) {
    // cluster (DSMEM) split-K: power-of-two cluster of at most 8 CTAs along K
    if (split_k > 8 || (split_k & (split_k - 1))) return FTG_ERR_UNSUPPORTED;
    workspace = nullptr;
    tile_counters = nullptr;
  }
  CUtensorMap ma, mb;
  bool ok = p.a_t ? make_map(&ma, p.a, p.M, p.K, p.lda, 32, BK, true) : make_map(&ma, p.a, p.K, p.M, p.lda, BK, BM, false);
  ok = ok && (p.b_t ? make_map(&mb, p.b, p.N, p.K, p.ldb, 32, BK, true) : make_map(&mb, p.b, p.K, p.N, p.ldb, BK, BN, false));
  if (!ok) return FTG_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  if (p.a_t) return p.b_t ? launch<true, true>(ma, mb, p, split_k, workspace, tile_counters, s)
                          : launch<true, false>(ma, mb, p, split_k, workspace, tile_counters, s);
  return p.b_t ? launch<false, true>(ma, mb, p, split_k, workspace, tile_counters, s)
               : launch<false, false>(ma, mb, p, split_k, workspace, tile_counters, s);
}

// Linear + BatchNorm1d(train) + ReLU + concat-store in one launch (EPI_BN, see bn_epilogue).
// Returns FTG_ERR_UNSUPPORTED when the shape does not fit one cluster per column panel
// (more than 4 M-tiles, i.e. batch > 512) or an epilogue flag other than bias is set: the caller then
// runs ftg_gemm_tc + ftg_bn_relu_fwd.
extern "C" int ftg_gemm_tc_bn_plan(int M, int K) {
  const int cy = ftg_cdiv(M, BM), nkb = ftg_cdiv(K, BK);
  if (cy < 1 || cy > 4) return 0;
  static int cap = -1;
  if (cap < 0) {
    const char* e = getenv("FTG_BN_SPLIT");       // experiments: largest split-K factor of the fused-BN GEMM
    cap = e ? atoi(e) : 8;
    if (cap < 1) cap = 1;
  }
  int s = 1;
  while (cy * s * 2 <= 8 && s * 2 <= cap && nkb / (s * 2) >= 2) s *= 2;
  return s;
}

extern "C" int ftg_gemm_tc_bn(const ftg_gemm_args* args, const ftg_bn_args* bn, ftg_stream_t stream) {
  const ftg_gemm_args& p = *args;
  if (!bn || !bn->gamma || !bn->beta || !bn->running_mean || !bn->running_var || !bn->out) return FTG_ERR_UNSUPPORTED;
  if (p.M <= 0 || p.N <= 0 || p.K <= 0) return FTG_ERR_UNSUPPORTED;
  if (p.act || p.mask_in || p.mask_out || p.accumulate) return FTG_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(p.a) & 15) || (reinterpret_cast<uintptr_t>(p.b) & 15) || (p.lda & 3) || (p.ldb & 3))
    return FTG_ERR_UNSUPPORTED;
  const int split_k = ftg_gemm_tc_bn_plan(p.M, p.K);
  if (split_k < 1) return FTG_ERR_UNSUPPORTED;
  CUtensorMap ma, mb;
  bool ok = p.a_t ? make_map(&ma, p.a, p.M, p.K, p.lda, 32, BK, true) : make_map(&ma, p.a, p.K, p.M, p.lda, BK, BM, false);
  ok = ok && (p.b_t ? make_map(&mb, p.b, p.N, p.K, p.ldb, 32, BK, true) : make_map(&mb, p.b, p.K, p.N, p.ldb, BK, BN, false));
  if (!ok) return FTG_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  if (p.a_t) return p.b_t ? launch<true, true>(ma, mb, p, split_k, nullptr, nullptr, s, bn)
                          : launch<true, false>(ma, mb, p, split_k, nullptr, nullptr, s, bn);
  return p.b_t ? launch<false, true>(ma, mb, p, split_k, nullptr, nullptr, s, bn)
               : launch<false, false>(ma, mb, p, split_k, nullptr, nullptr, s, bn);
}

// ---------------------------------------------------------------------------------------
// ftg_gemm_tc_noalloc: the same GEMM for callers that must not allocate device memory.
//
// Some users of this library cannot call cudaMalloc at the point where they need a GEMM:
// a plugin running inside a host application that has handed the whole device heap to its
// own pool, a fail-safe path that has to keep working when the device is out of memory, a
// latency-critical loop that cannot afford cudaMalloc's implicit device synchronisation.
// ftg_gemm_tc() asks such a caller for a split-K workspace and tile counters; this variant
// takes them from a fixed, file-scope device array instead.  The caller still passes the
// operands and the output.
//
//  * The array costs FTG_NOALLOC_SCRATCH_BYTES (8 MiB + 4 KiB) of device memory on every
//    GPU that loads the library, from load until process exit.
//  * Requests that need more (split_k * tiles * 32 KiB of partial tiles, or more than
//    FTG_NOALLOC_MAX_TILES output tiles) are refused with FTG_ERR_SCRATCH_TOO_SMALL and a
//    message in ftg_last_error_message(); lower split_k (split_k = 1 needs no scratch).
//  * The first call on a device resolves the array's device address (cudaGetSymbolAddress,
//    once per device, with that device current).  Make that first call before any stream
//    capture begins on the device.
//  * Every call on a device shares the one array: the tile counters are cleared at the start
//    of each call and the partial tiles are fully rewritten before they are read, but two
//    calls whose kernels overlap on the GPU (different streams) race on it.  Serialise them.
//  * No cudaMalloc, cudaFree or allocating library routine is called on this path.
// ---------------------------------------------------------------------------------------
namespace {

constexpr long long NOALLOC_WS_BYTES = 8ll << 20;
constexpr int NOALLOC_MAX_TILES = 1024;
constexpr long long NOALLOC_COUNTER_BYTES = (long long)NOALLOC_MAX_TILES * sizeof(int);

// [ tile counters (4 KiB) | split-K partial tiles (8 MiB) ]
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_COUNTER_BYTES + NOALLOC_WS_BYTES];

constexpr int MAX_DEVICES = 64;
unsigned char* g_noalloc_ptr[MAX_DEVICES] = {nullptr};
std::mutex g_noalloc_mutex;
thread_local char g_last_error[256] = "";

// `g_noalloc_scratch` named in host code is only a host-side stand-in for the symbol; the
// address a kernel can use has to be asked for, per device.
int noalloc_device_ptr(unsigned char** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= MAX_DEVICES) return FTG_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess) {
      snprintf(g_last_error, sizeof(g_last_error), "ftg_gemm_tc_noalloc: cudaGetSymbolAddress failed on device %d: %s",
               dev, cudaGetErrorString(e));
      return (int)e;
    }
    g_noalloc_ptr[dev] = static_cast<unsigned char*>(ptr);
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

}  // namespace

extern "C" const char* ftg_last_error_message(void) { return g_last_error; }

extern "C" long long ftg_gemm_tc_noalloc_max_workspace_bytes(void) { return NOALLOC_WS_BYTES; }

extern "C" int ftg_gemm_tc_noalloc(const ftg_gemm_args* args, int split_k, ftg_stream_t stream) {
  g_last_error[0] = 0;
  if (!args) return FTG_ERR_UNSUPPORTED;
  if (split_k < 1) split_k = 1;
  const int nkb = ftg_cdiv(args->K > 0 ? args->K : 1, BK);
  if (split_k > nkb) split_k = nkb;
  const int tiles = ftg_gemm_tc_num_tiles(args->M, args->N);
  const long long need = ftg_gemm_tc_workspace_bytes(args->M, args->N, split_k);
  if (split_k > 1 && (need > NOALLOC_WS_BYTES || tiles > NOALLOC_MAX_TILES)) {
    snprintf(g_last_error, sizeof(g_last_error),
             "ftg_gemm_tc_noalloc: M=%d N=%d split_k=%d needs %lld bytes of split-K scratch in %d tiles; the built-in "
             "scratch holds %lld bytes and %d tiles (use a smaller split_k, or ftg_gemm_tc with your own workspace)",
             args->M, args->N, split_k, need, tiles, NOALLOC_WS_BYTES, NOALLOC_MAX_TILES);
    return FTG_ERR_SCRATCH_TOO_SMALL;
  }
  unsigned char* base = nullptr;
  int rc = noalloc_device_ptr(&base);
  if (rc) return rc;
  int* counters = reinterpret_cast<int*>(base);
  float* workspace = reinterpret_cast<float*>(base + NOALLOC_COUNTER_BYTES);
  if (split_k > 1) {
    // the counters must start at zero; an earlier call that died half-way could have left them dirty
    cudaError_t e = cudaMemsetAsync(counters, 0, (size_t)tiles * sizeof(int), (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
  }
  return ftg_gemm_tc(args, split_k, workspace, counters, stream);
}
