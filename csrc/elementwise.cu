// Fused element-wise / reduction / sampling kernels of the CTGAN step (sm_100a).
// Each kernel's fp32 PyTorch oracle is the function of the same name (without the ftg_
// prefix) in fed_tgan_b200/ops/reference.py.
#include "common.cuh"
#include <cstdlib>

std::atomic<long long> g_ftg_launch_count{0};
int g_ftg_pdl = [] {
  const char* e = getenv("FEDTGAN_PDL");
  return (e && e[0] == '0') ? 0 : 1;
}();
extern "C" void ftg_set_pdl(int on) { g_ftg_pdl = on ? 1 : 0; }

extern "C" const char* ftg_version(void) { return "fedtgan_b200 0.1.0 (sm_100a)"; }
extern "C" long long ftg_last_launch_count(void) { return g_ftg_launch_count.load(); }

// ------------------------------------------------------------------ rng tick
__global__ void k_tick(int64_t* state) {
  FTG_PDL_ENTER(); state[1] += 1; }

extern "C" int ftg_tick(int64_t* rng_state, ftg_stream_t stream) {
  ftg_launch(k_tick, dim3(1), dim3(1), 0, (cudaStream_t)stream, rng_state);
  FTG_LAUNCH_CHECK();
  return 0;
}

// --------------------------------------------------------------- fill_normal
__global__ void k_fill_normal(float* out, int rows, int cols, int ld, const int64_t* state, int stream_id) {
  long long total = (long long)rows * cols;
  long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // one Philox 4-vector per thread
  if (q * 4 >= total) return;
  Philox4 p = philox4x32_10((uint64_t)state[0], (uint64_t)state[1], (uint32_t)stream_id, (uint32_t)q);
  float u0 = fmaxf(u01(p.x), 5.96e-8f), u1 = u01(p.y), u2 = fmaxf(u01(p.z), 5.96e-8f), u3 = u01(p.w);
  float r0 = sqrtf(-2.f * __logf(u0)), r1 = sqrtf(-2.f * __logf(u2));
  float s0, c0, s1, c1;
  __sincosf(6.283185307179586f * u1, &s0, &c0);
  __sincosf(6.283185307179586f * u3, &s1, &c1);
  float z[4] = {r0 * c0, r0 * s0, r1 * c1, r1 * s1};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    long long e = q * 4 + j;
    if (e < total) out[(e / cols) * (long long)ld + (e % cols)] = z[j];
  }
}

extern "C" int ftg_fill_normal(float* out, int rows, int cols, int ld, const int64_t* rng_state, int rng_stream,
                               ftg_stream_t stream) {
  long long quads = ((long long)rows * cols + 3) / 4;
  if (quads == 0) return 0;
  k_fill_normal<<<ftg_cdiv(quads, 256), 256, 0, (cudaStream_t)stream>>>(out, rows, cols, ld, rng_state, rng_stream);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------ sample_condvec
__global__ void k_sample_condvec(const float* __restrict__ cum_p, int n_col, int max_w,
                                 const int* __restrict__ soft_width, int* col_out, int* opt_out, int B,
                                 const float* u_col, const float* u_opt, const int64_t* state, int stream_id) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  float uc = u_col ? u_col[i] : philox_uniform(state, stream_id, (uint32_t)i);
  float uo = u_opt ? u_opt[i] : philox_uniform(state, stream_id + 1, (uint32_t)i);
  int c = min((int)(uc * n_col), n_col - 1);
  int w = soft_width[c];
  const float* row = cum_p + (size_t)c * max_w;
  int o = w - 1;
  for (int j = 0; j < w; ++j) {
    if (row[j] > uo) { o = j; break; }
  }
  col_out[i] = c;
  opt_out[i] = o;
}

extern "C" int ftg_sample_condvec(const float* cum_p, int n_col, int max_w, const int* soft_width, int* col_out,
                                  int* opt_out, int B, const float* u_col, const float* u_opt,
                                  const int64_t* rng_state, int rng_stream, ftg_stream_t stream) {
  k_sample_condvec<<<ftg_cdiv(B, 128), 128, 0, (cudaStream_t)stream>>>(cum_p, n_col, max_w, soft_width, col_out,
                                                                       opt_out, B, u_col, u_opt, rng_state, rng_stream);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------- write_condvec
__global__ void k_write_condvec(const int* __restrict__ col, const int* __restrict__ opt,
                                const int* __restrict__ soft_cstart, float* c_out, int ldc, int n_opt, float* m_out,
                                int ldm, int n_col, const int* perm) {
  int i = blockIdx.x;
  int src = perm ? perm[i] : i;
  int c = col[src];
  int target = soft_cstart[c] + opt[src];
  for (int j = threadIdx.x; j < n_opt; j += blockDim.x) c_out[(size_t)i * ldc + j] = (j == target) ? 1.f : 0.f;
  if (m_out)
    for (int j = threadIdx.x; j < n_col; j += blockDim.x) m_out[(size_t)i * ldm + j] = (j == c) ? 1.f : 0.f;
}

extern "C" int ftg_write_condvec(const int* col, const int* opt, const int* soft_cstart, float* c_out, int ldc,
                                 int n_opt, float* m_out, int ldm, int n_col, const int* perm, int B,
                                 ftg_stream_t stream) {
  k_write_condvec<<<B, 128, 0, (cudaStream_t)stream>>>(col, opt, soft_cstart, c_out, ldc, n_opt, m_out, ldm, n_col, perm);
  FTG_LAUNCH_CHECK();
  return 0;
}

// -------------------------------------------------------------- argsort_perm
// One CTA: bitonic sort of (key, index) pairs in shared memory.  B <= 4096.
__global__ void k_argsort_perm(int* perm, int B, int P2, const float* u, const int64_t* state, int stream_id) {
  extern __shared__ unsigned char smem_raw[];
  float* key = reinterpret_cast<float*>(smem_raw);
  int* idx = reinterpret_cast<int*>(key + P2);
  for (int i = threadIdx.x; i < P2; i += blockDim.x) {
    key[i] = (i < B) ? (u ? u[i] : philox_uniform(state, stream_id, (uint32_t)i)) : 3.0e38f;
    idx[i] = i;
  }
  __syncthreads();
  for (int k = 2; k <= P2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < P2; i += blockDim.x) {
        int l = i ^ j;
        if (l > i) {
          bool up = ((i & k) == 0);
          float a = key[i], b = key[l];
          int ia = idx[i], ib = idx[l];
          bool gt = (a > b) || (a == b && ia > ib);
          if (gt == up) { key[i] = b; key[l] = a; idx[i] = ib; idx[l] = ia; }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < B; i += blockDim.x) perm[i] = idx[i];
}

extern "C" int ftg_argsort_perm(int* perm, int B, const float* u, const int64_t* rng_state, int rng_stream,
                                ftg_stream_t stream) {
  int P2 = 1;
  while (P2 < B) P2 <<= 1;
  if (P2 > 4096) return FTG_ERR_UNSUPPORTED;
  int threads = P2 < 1024 ? (P2 < 32 ? 32 : P2) : 1024;
  k_argsort_perm<<<1, threads, (size_t)P2 * 8, (cudaStream_t)stream>>>(perm, B, P2, u, rng_state, rng_stream);
  FTG_LAUNCH_CHECK();
  return 0;
}

// --------------------------------------------------------------- sample_rows
__global__ void k_sample_rows(const int* __restrict__ row_ptr, const int* __restrict__ row_ids,
                              const int* __restrict__ soft_cstart, const int* col, const int* opt, const int* perm,
                              int* rid_out, int B, const float* u, const int64_t* state, int stream_id) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  int src = perm ? perm[i] : i;
  int g = soft_cstart[col[src]] + opt[src];
  int lo = row_ptr[g], cnt = row_ptr[g + 1] - lo;
  float uu = u ? u[i] : philox_uniform(state, stream_id, (uint32_t)i);
  int k = min((int)(uu * cnt), max(cnt - 1, 0));
  rid_out[i] = row_ids[lo + k];
}

extern "C" int ftg_sample_rows(const int* row_ptr, const int* row_ids, const int* soft_cstart, const int* col,
                               const int* opt, const int* perm, int* rid_out, int B, const float* u,
                               const int64_t* rng_state, int rng_stream, ftg_stream_t stream) {
  k_sample_rows<<<ftg_cdiv(B, 128), 128, 0, (cudaStream_t)stream>>>(row_ptr, row_ids, soft_cstart, col, opt, perm,
                                                                    rid_out, B, u, rng_state, rng_stream);
  FTG_LAUNCH_CHECK();
  return 0;
}

// --------------------------------------------------------------- expand_rows
__global__ void k_expand_rows(const float* __restrict__ vals, int n_tanh, const int* __restrict__ codes, int n_col,
                              const int* __restrict__ rid, const int* __restrict__ tanh_start,
                              const int* __restrict__ soft_start, float* out, int ld, int dd) {
  int i = blockIdx.x;
  int r = rid[i];
  float* o = out + (size_t)i * ld;
  for (int j = threadIdx.x; j < dd; j += blockDim.x) o[j] = 0.f;
  __syncthreads();
  for (int t = threadIdx.x; t < n_tanh; t += blockDim.x) o[tanh_start[t]] = vals[(size_t)r * n_tanh + t];
  for (int s = threadIdx.x; s < n_col; s += blockDim.x) o[soft_start[s] + codes[(size_t)r * n_col + s]] = 1.f;
}

extern "C" int ftg_expand_rows(const float* vals, int n_tanh, const int* codes, int n_col, const int* rid,
                               const int* tanh_start, const int* soft_start, float* out, int ld, int dd, int B,
                               ftg_stream_t stream) {
  k_expand_rows<<<B, 128, 0, (cudaStream_t)stream>>>(vals, n_tanh, codes, n_col, rid, tanh_start, soft_start, out, ld, dd);
  FTG_LAUNCH_CHECK();
  return 0;
}

// -------------------------------------------------------------------- rowdot
// out[m] = a[m,:].w + bias ; optionally loss_out = sum_m rowscale[m] * out[m] (fused WGAN loss value).
// M is a few hundred at most: one CTA, one warp per row, deterministic block reduction.
__global__ void __launch_bounds__(1024) k_rowdot(const float* __restrict__ a, int lda, int M, int K,
                                                 const float* __restrict__ w, const float* bias, float* out,
                                                 const float* __restrict__ rowscale, float* loss_out) {
  // 8 lanes per row (128 rows in flight per CTA), strided 8 apart so each lane keeps
  // several independent loads in flight
  __shared__ float red[32];
  const int sub = threadIdx.x & 7, rloc = threadIdx.x >> 3, rows_per_pass = blockDim.x >> 3;
  float contrib = 0.f;
  for (int row0 = blockIdx.x * rows_per_pass; row0 < M; row0 += gridDim.x * rows_per_pass) {
    const int row = row0 + rloc;
    float s = 0.f;
    if (row < M) {
      const float* ar = a + (size_t)row * lda;
#pragma unroll 4
      for (int k = sub; k < K; k += 8) s += ar[k] * __ldg(w + k);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (row < M && sub == 0) {
      s += bias ? bias[0] : 0.f;
      out[row] = s;
      if (rowscale) contrib += rowscale[row] * s;
    }
  }
  if (loss_out) {
    float tot = block_sum(contrib, red);
    if (threadIdx.x == 0) loss_out[0] = tot;
  }
}

extern "C" int ftg_rowdot(const float* a, int lda, int M, int K, const float* w, const float* bias, float* out,
                          ftg_stream_t stream) {
  k_rowdot<<<ftg_cdiv(M, 128), 1024, 0, (cudaStream_t)stream>>>(a, lda, M, K, w, bias, out, nullptr, nullptr);
  FTG_LAUNCH_CHECK();
  return 0;
}

// Discriminator head, forward + first backward signal in ONE launch:
//   block 0      : y[m] = a[m,:].w + bias ; loss = sum_m rs_loss[m] * y[m]
//   blocks 1..   : Z[m,n] = rs_sig[m] * w[n] * mask[m,n]      (d loss / d pre-activation of the last hidden layer)
__global__ void __launch_bounds__(1024) k_d_head(const float* __restrict__ a, int lda, int M, int K,
                                                 const float* __restrict__ w, const float* bias, float* y,
                                                 const float* __restrict__ rs_loss, float* loss_out,
                                                 const float* __restrict__ rs_sig, const float* __restrict__ mask,
                                                 int ldm, float* Z, int ldz) {
  FTG_PDL_ENTER();
  if (blockIdx.x == 0) {
    __shared__ float red[32];
    const int sub = threadIdx.x & 7, rloc = threadIdx.x >> 3, rpp = blockDim.x >> 3;
    float contrib = 0.f;
    for (int row0 = 0; row0 < M; row0 += rpp) {
      const int row = row0 + rloc;
      float s = 0.f;
      if (row < M) {
        const float* ar = a + (size_t)row * lda;
#pragma unroll 4
        for (int k = sub; k < K; k += 8) s += ar[k] * __ldg(w + k);
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      if (row < M && sub == 0) {
        s += bias ? bias[0] : 0.f;
        y[row] = s;
        contrib += rs_loss[row] * s;
      }
    }
    const float tot = block_sum(contrib, red);
    if (threadIdx.x == 0) loss_out[0] = tot;
  } else {
    const long long total = (long long)M * K;
    for (long long e = (long long)(blockIdx.x - 1) * blockDim.x + threadIdx.x; e < total;
         e += (long long)(gridDim.x - 1) * blockDim.x) {
      const int m = (int)(e / K), n = (int)(e % K);
      Z[(size_t)m * ldz + n] = rs_sig[m] * __ldg(w + n) * mask[(size_t)m * ldm + n];
    }
  }
}

extern "C" int ftg_d_head(const float* a, int lda, int M, int K, const float* w, const float* bias, float* y,
                          const float* rs_loss, float* loss_out, const float* rs_sig, const float* mask, int ldm,
                          float* Z, int ldz, ftg_stream_t stream) {
  int extra = ftg_cdiv((long long)M * K, 1024 * 4);
  if (extra < 1) extra = 1;
  ftg_launch(k_d_head, dim3(1 + extra), dim3(1024), 0, (cudaStream_t)stream, a, lda, M, K, w, bias, y, rs_loss, loss_out, rs_sig, mask, ldm,
                                                        Z, ldz);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_rowdot_loss(const float* a, int lda, int M, int K, const float* w, const float* bias, float* out,
                               const float* rowscale, float* loss_out, ftg_stream_t stream) {
  k_rowdot<<<1, 1024, 0, (cudaStream_t)stream>>>(a, lda, M, K, w, bias, out, rowscale, loss_out);
  FTG_LAUNCH_CHECK();
  return 0;
}

// -------------------------------------------------------------------- colsum
// blockDim = (32, 32): 32 columns per block, rows strided over y; deterministic tree.
__global__ void __launch_bounds__(1024) k_colsum(const float* __restrict__ a, int lda, int M, int N,
                                                 const float* __restrict__ rowscale, float* out, int accumulate) {
  FTG_PDL_ENTER();
  __shared__ float part[32][33];
  int n = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (n < N)
    for (int m = threadIdx.y; m < M; m += 32) s += a[(size_t)m * lda + n] * (rowscale ? rowscale[m] : 1.f);
  part[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) t += part[j][threadIdx.x];
    out[n] = accumulate ? out[n] + t : t;
  }
}

extern "C" int ftg_colsum(const float* a, int lda, int M, int N, const float* rowscale, float* out, int accumulate,
                          ftg_stream_t stream) {
  ftg_launch(k_colsum, dim3(ftg_cdiv(N, 32)), dim3(dim3(32, 32)), 0, (cudaStream_t)stream, a, lda, M, N, rowscale, out, accumulate);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------- scale_outer_mask
__global__ void k_scale_outer_mask(const float* __restrict__ rowscale, const float* __restrict__ v,
                                   const float* __restrict__ mask, int ldm, float* out, int ldo, int M, int N) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  int m = blockIdx.y;
  if (n < N) out[(size_t)m * ldo + n] = rowscale[m] * v[n] * mask[(size_t)m * ldm + n];
}

extern "C" int ftg_scale_outer_mask(const float* rowscale, const float* v, const float* mask, int ldm, float* out,
                                    int ldo, int M, int N, ftg_stream_t stream) {
  k_scale_outer_mask<<<dim3(ftg_cdiv(N, 128), M), 128, 0, (cudaStream_t)stream>>>(rowscale, v, mask, ldm, out, ldo, M, N);
  FTG_LAUNCH_CHECK();
  return 0;
}

// --------------------------------------------------------------- bn_relu_fwd
// blockDim = (8, 128): 8 features per block (one 32 B sector per row), 128 row lanes, so a
// 256-wide layer fills 32 CTAs with 4 rows per thread; two-pass batch statistics in fp32.
#define BN_F 8
#define BN_R 128
__device__ __forceinline__ float bn_block_reduce(float v, float (*red)[BN_F], int tx, int ty) {
  // reduce over ty (128 lanes = 4 warps of 32 threads; a warp holds 4 rows x 8 features)
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  __syncthreads();
  if ((ty & 3) == 0) red[ty >> 2][tx] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int j = 0; j < BN_R / 4; ++j) t += red[j][tx];
  return t;
}

__global__ void __launch_bounds__(BN_F * BN_R)
k_bn_relu_fwd(const float* __restrict__ a, int lda, int B, int H, const float* __restrict__ gamma,
              const float* __restrict__ beta, float* running_mean, float* running_var, float* out, int ldo,
              float* save_mean, float* save_invstd, float eps, float momentum, float* nbt) {
  FTG_PDL_ENTER();
  __shared__ float red[BN_R / 4][BN_F];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int h = blockIdx.x * BN_F + tx;
  const bool ok = h < H;
  // rows ty, ty+128, ... : the first 8 per thread stay in registers (B <= 1024 never re-reads)
  float x[8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = ty + i * BN_R;
    x[i] = (ok && m < B) ? a[(size_t)m * lda + h] : 0.f;
    s += x[i];
  }
  for (int m = ty + 8 * BN_R; m < B; m += BN_R) s += ok ? a[(size_t)m * lda + h] : 0.f;
  const float mu = bn_block_reduce(s, red, tx, ty) / (float)B;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float d = x[i] - mu;
    if (ty + i * BN_R < B) q += d * d;
  }
  for (int m = ty + 8 * BN_R; m < B; m += BN_R) {
    const float d = (ok ? a[(size_t)m * lda + h] : 0.f) - mu;
    q += d * d;
  }
  const float var = bn_block_reduce(q, red, tx, ty) / (float)B;
  const float inv = rsqrtf(var + eps);
  if (ty == 0 && ok) {
    running_mean[h] = (1.f - momentum) * running_mean[h] + momentum * mu;
    running_var[h] = (1.f - momentum) * running_var[h] + momentum * var * ((float)B / (float)max(B - 1, 1));
    if (save_mean) { save_mean[h] = mu; save_invstd[h] = inv; }
  }
  if (blockIdx.x == 0 && tx == 0 && ty == 0 && nbt) nbt[0] += 1.f;
  if (!ok) return;
  const float g = gamma[h], b = beta[h];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = ty + i * BN_R;
    if (m < B) out[(size_t)m * ldo + h] = fmaxf((x[i] - mu) * inv * g + b, 0.f);
  }
  for (int m = ty + 8 * BN_R; m < B; m += BN_R)
    out[(size_t)m * ldo + h] = fmaxf((a[(size_t)m * lda + h] - mu) * inv * g + b, 0.f);
}

// eval mode is row independent: tile rows over blockIdx.y so 40 000-row generation fills the GPU
__global__ void k_bn_relu_eval(const float* __restrict__ a, int lda, int B, int H, const float* __restrict__ gamma,
                               const float* __restrict__ beta, const float* __restrict__ running_mean,
                               const float* __restrict__ running_var, float* out, int ldo, float eps) {
  int h = blockIdx.x * 32 + threadIdx.x;
  if (h >= H) return;
  float mu = running_mean[h], inv = rsqrtf(running_var[h] + eps), g = gamma[h], b = beta[h];
  int r0 = blockIdx.y * 128;
  int r1 = min(r0 + 128, B);
  for (int m = r0 + threadIdx.y; m < r1; m += 8) {
    float y = (a[(size_t)m * lda + h] - mu) * inv * g + b;
    out[(size_t)m * ldo + h] = fmaxf(y, 0.f);
  }
}

extern "C" int ftg_bn_relu_fwd(const float* a, int lda, int B, int H, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, float* out, int ldo, float* save_mean,
                               float* save_invstd, int training, float eps, float momentum,
                               float* num_batches_tracked, ftg_stream_t stream) {
  if (training) {
    ftg_launch(k_bn_relu_fwd, dim3(ftg_cdiv(H, BN_F)), dim3(dim3(BN_F, BN_R)), 0, (cudaStream_t)stream, a, lda, B, H, gamma, beta, running_mean, running_var, out, ldo, save_mean, save_invstd, eps, momentum,
        num_batches_tracked);
  } else {
    k_bn_relu_eval<<<dim3(ftg_cdiv(H, 32), ftg_cdiv(B, 128)), dim3(32, 8), 0, (cudaStream_t)stream>>>(
        a, lda, B, H, gamma, beta, running_mean, running_var, out, ldo, eps);
  }
  FTG_LAUNCH_CHECK();
  return 0;
}

// --------------------------------------------------------------- bn_relu_bwd
// Also emits d(fc bias) = column sum of `da` (the Linear in front of the BatchNorm), so the
// generator backward needs no separate reduction kernel.
__global__ void __launch_bounds__(BN_F * BN_R)
k_bn_relu_bwd(const float* __restrict__ dout, int lddo, const float* __restrict__ a, int lda, int B, int H,
              const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
              const float* __restrict__ gamma, const float* __restrict__ beta, float* da, int ldda, float* dgamma,
              float* dbeta, float* dbias_fc) {
  FTG_PDL_ENTER();
  __shared__ float red[BN_R / 4][BN_F];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int h = blockIdx.x * BN_F + tx;
  const bool ok = h < H;
  const float mu = ok ? save_mean[h] : 0.f, inv = ok ? save_invstd[h] : 0.f, g = ok ? gamma[h] : 0.f,
              b = ok ? beta[h] : 0.f;
  float xh[8], dy[8];
  float sb = 0.f, sg = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = ty + i * BN_R;
    const bool in = ok && m < B;
    xh[i] = in ? (a[(size_t)m * lda + h] - mu) * inv : 0.f;
    dy[i] = (in && xh[i] * g + b > 0.f) ? dout[(size_t)m * lddo + h] : 0.f;
    sb += dy[i];
    sg += dy[i] * xh[i];
  }
  for (int m = ty + 8 * BN_R; m < B; m += BN_R) {
    const float x = ok ? (a[(size_t)m * lda + h] - mu) * inv : 0.f;
    const float d = (ok && x * g + b > 0.f) ? dout[(size_t)m * lddo + h] : 0.f;
    sb += d;
    sg += d * x;
  }
  const float db = bn_block_reduce(sb, red, tx, ty);
  const float dg = bn_block_reduce(sg, red, tx, ty);
  if (ty == 0 && ok) { dbeta[h] = db; dgamma[h] = dg; }
  const float scale = g * inv / (float)B;
  float sda = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = ty + i * BN_R;
    if (ok && m < B) {
      const float v = scale * ((float)B * dy[i] - db - xh[i] * dg);
      da[(size_t)m * ldda + h] = v;
      sda += v;
    }
  }
  for (int m = ty + 8 * BN_R; m < B; m += BN_R) {
    if (!ok) continue;
    const float x = (a[(size_t)m * lda + h] - mu) * inv;
    const float d = (x * g + b > 0.f) ? dout[(size_t)m * lddo + h] : 0.f;
    const float v = scale * ((float)B * d - db - x * dg);
    da[(size_t)m * ldda + h] = v;
    sda += v;
  }
  if (dbias_fc) {
    const float t = bn_block_reduce(sda, red, tx, ty);
    if (ty == 0 && ok) dbias_fc[h] = t;
  }
}

extern "C" int ftg_bn_relu_bwd(const float* dout, int lddo, const float* a, int lda, int B, int H,
                               const float* save_mean, const float* save_invstd, const float* gamma,
                               const float* beta, float* da, int ldda, float* dgamma, float* dbeta,
                               ftg_stream_t stream) {
  ftg_launch(k_bn_relu_bwd, dim3(ftg_cdiv(H, BN_F)), dim3(dim3(BN_F, BN_R)), 0, (cudaStream_t)stream, dout, lddo, a, lda, B, H, save_mean, save_invstd, gamma, beta, da, ldda, dgamma, dbeta, nullptr);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_bn_relu_bwd_bias(const float* dout, int lddo, const float* a, int lda, int B, int H,
                                    const float* save_mean, const float* save_invstd, const float* gamma,
                                    const float* beta, float* da, int ldda, float* dgamma, float* dbeta,
                                    float* dbias_fc, ftg_stream_t stream) {
  ftg_launch(k_bn_relu_bwd, dim3(ftg_cdiv(H, BN_F)), dim3(dim3(BN_F, BN_R)), 0, (cudaStream_t)stream, dout, lddo, a, lda, B, H, save_mean, save_invstd, gamma, beta, da, ldda, dgamma, dbeta, dbias_fc);
  FTG_LAUNCH_CHECK();
  return 0;
}

// -------------------------------------------------------------- activate_fwd
// One CTA per row.  Phase 1: every column computes tanh or the tempered, gumbel-perturbed
// logit into shared memory.  Phase 2: one thread per softmax span finds max / sum.
// Phase 3: every column normalises.  The conditional vector is appended by the same CTA
// (this is the torch.cat([fakeact, c1]) of the reference written in place).
__global__ void __launch_bounds__(128)
k_activate_fwd(const float* __restrict__ fake, int ldf, int dd, int n_spans, const int* __restrict__ span_start,
               const int* __restrict__ span_width, const int* __restrict__ span_kind,
               const int* __restrict__ col_span, float* out, int ldo, float inv_tau, const float* u, int ldu,
               const int64_t* state, int stream_id, const float* cond, int ldc, int n_opt, int use_noise,
               const float* __restrict__ real, int ldr, float* interp, int ldi, const float* alpha, int alpha_stream) {
  FTG_PDL_ENTER();
  extern __shared__ float sm[];
  __shared__ float red[32];
  float* val = sm;                 // [dd]
  float* smax = sm + dd;           // [n_spans]
  float* sinv = smax + n_spans;    // [n_spans]
  float* rowbuf = sinv + n_spans;  // [dd + n_opt]  the finished fake_cat row (only used with slerp)
  const int row = blockIdx.x;
  const float* x = fake + (size_t)row * ldf;
  float* o = out + (size_t)row * ldo;
  for (int j = threadIdx.x; j < dd; j += blockDim.x) {
    const int s = col_span[j];
    if (span_kind[s] == 0) {
      val[j] = tanhf(x[j]);
    } else {
      float g = 0.f;
      if (use_noise) {
        float uu = u ? u[(size_t)row * ldu + j] : philox_uniform(state, stream_id, (uint32_t)(row * dd + j));
        uu = fminf(fmaxf(uu, 1e-10f), 1.0f - 1e-7f);
        g = -logf(-logf(uu));
      }
      val[j] = (x[j] + g) * inv_tau;
    }
  }
  __syncthreads();
  for (int s = threadIdx.x; s < n_spans; s += blockDim.x) {
    if (span_kind[s] == 0) continue;
    const int st = span_start[s], w = span_width[s];
    float mx = -3.0e38f;
    for (int k = 0; k < w; ++k) mx = fmaxf(mx, val[st + k]);
    float sum = 0.f;
    for (int k = 0; k < w; ++k) sum += expf(val[st + k] - mx);
    smax[s] = mx;
    sinv[s] = 1.f / sum;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < dd; j += blockDim.x) {
    const int s = col_span[j];
    const float v = (span_kind[s] == 0) ? val[j] : expf(val[j] - smax[s]) * sinv[s];
    o[j] = v;
    if (interp) rowbuf[j] = v;
  }
  if (cond)
    for (int j = threadIdx.x; j < n_opt; j += blockDim.x) {
      const float c = cond[(size_t)row * ldc + j];
      o[dd + j] = c;
      if (interp) rowbuf[dd + j] = c;
    }
  if (!interp) return;
  // fused slerp (ctgan.py:231-237): interp = slerp(alpha, low = real_cat row, high = this fake_cat row)
  __syncthreads();
  const int din = dd + n_opt;
  const float* l = real + (size_t)row * ldr;
  float ll = 0.f, hh = 0.f, lh = 0.f;
  for (int j = threadIdx.x; j < din; j += blockDim.x) {
    const float a = l[j], b = rowbuf[j];
    ll += a * a; hh += b * b; lh += a * b;
  }
  ll = block_sum(ll, red);
  hh = block_sum(hh, red);
  lh = block_sum(lh, red);
  const float omega = acosf(lh / (sqrtf(ll) * sqrtf(hh)));
  const float so = sinf(omega);
  const float a = alpha ? alpha[row] : philox_uniform(state, alpha_stream, (uint32_t)row);
  const float wl = sinf((1.f - a) * omega) / so, wh = sinf(a * omega) / so;
  float* io = interp + (size_t)row * ldi;
  for (int j = threadIdx.x; j < din; j += blockDim.x) io[j] = wl * l[j] + wh * rowbuf[j];
}

static int activate_fwd_launch(const float* fake, int ldf, int B, int dd, int n_spans, const int* span_start,
                               const int* span_width, const int* span_kind, const int* col_span, float* out, int ldo,
                               float tau, const float* u, int ldu, const int64_t* rng_state, int rng_stream,
                               const float* cond, int ldc, int n_opt, int use_noise, const float* real, int ldr,
                               float* interp, int ldi, const float* alpha, int alpha_stream, ftg_stream_t stream) {
  size_t smem = (size_t)(dd + 2 * n_spans + (interp ? dd + n_opt : 0)) * sizeof(float);
  if (smem > 48 * 1024) return FTG_ERR_UNSUPPORTED;
  ftg_launch(k_activate_fwd, dim3(B), dim3(128), smem, (cudaStream_t)stream, fake, ldf, dd, n_spans, span_start,
             span_width, span_kind, col_span, out, ldo, 1.f / tau, u, ldu, rng_state, rng_stream, cond, ldc, n_opt,
             use_noise, real, ldr, interp, ldi, alpha, alpha_stream);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_activate_fwd(const float* fake, int ldf, int B, int dd, int n_spans, const int* span_start,
                                const int* span_width, const int* span_kind, const int* col_span, float* out,
                                int ldo, float tau, const float* u, int ldu, const int64_t* rng_state,
                                int rng_stream, const float* cond, int ldc, int n_opt, int use_noise,
                                ftg_stream_t stream) {
  return activate_fwd_launch(fake, ldf, B, dd, n_spans, span_start, span_width, span_kind, col_span, out, ldo, tau, u,
                             ldu, rng_state, rng_stream, cond, ldc, n_opt, use_noise, nullptr, 0, nullptr, 0, nullptr,
                             0, stream);
}

// activate + concat + slerp in one launch: also writes interp = slerp(alpha, real, out) row by row
extern "C" int ftg_activate_fwd_slerp(const float* fake, int ldf, int B, int dd, int n_spans, const int* span_start,
                                      const int* span_width, const int* span_kind, const int* col_span, float* out,
                                      int ldo, float tau, const float* u, int ldu, const int64_t* rng_state,
                                      int rng_stream, const float* cond, int ldc, int n_opt, int use_noise,
                                      const float* real, int ldr, float* interp, int ldi, const float* alpha,
                                      int alpha_stream, ftg_stream_t stream) {
  return activate_fwd_launch(fake, ldf, B, dd, n_spans, span_start, span_width, span_kind, col_span, out, ldo, tau, u,
                             ldu, rng_state, rng_stream, cond, ldc, n_opt, use_noise, real, ldr, interp, ldi, alpha,
                             alpha_stream, stream);
}

// -------------------------------------------------------------- activate_bwd
// Optionally fused with cond_loss (ctgan.py:174-194): the thread that owns (row, softmax span j)
// also adds m[row,j]/B * (softmax(fake span) - onehot(target)) and the span's CE to the loss.
__global__ void k_activate_bwd(const float* __restrict__ dout, int lddo, const float* __restrict__ act, int lda, int B,
                               int n_spans, const int* __restrict__ span_start, const int* __restrict__ span_width,
                               const int* __restrict__ span_kind, float* dfake, int ldf, float inv_tau, int accumulate,
                               const float* __restrict__ fake, int ldfk, const float* __restrict__ cond, int ldc,
                               const float* __restrict__ mask, int ldm, const int* __restrict__ col_soft,
                               const int* __restrict__ soft_cstart, float* loss_out) {
  FTG_PDL_ENTER();
  __shared__ float red[32];
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float contrib = 0.f;
  if (t < (long long)B * n_spans) {
    int row = (int)(t / n_spans), s = (int)(t % n_spans);
    int st = span_start[s], w = span_width[s];
    const float* d = dout + (size_t)row * lddo + st;
    const float* y = act + (size_t)row * lda + st;
    float* o = dfake + (size_t)row * ldf + st;
    if (span_kind[s] == 0) {
      for (int j = 0; j < w; ++j) {
        float g = d[j] * (1.f - y[j] * y[j]);
        o[j] = accumulate ? o[j] + g : g;
      }
    } else {
      float dot = 0.f;
      for (int j = 0; j < w; ++j) dot += d[j] * y[j];
      float mj = 0.f;
      int target = 0;
      float lse = 0.f;
      const float* x = nullptr;
      if (loss_out) {
        const int js = col_soft[st];
        mj = mask[(size_t)row * ldm + js];
        if (mj != 0.f) {
          x = fake + (size_t)row * ldfk + st;
          const float* c = cond + (size_t)row * ldc + soft_cstart[js];
          float best = c[0], mx = x[0];
          for (int k = 1; k < w; ++k) {
            if (c[k] > best) { best = c[k]; target = k; }
            mx = fmaxf(mx, x[k]);
          }
          float sum = 0.f;
          for (int k = 0; k < w; ++k) sum += expf(x[k] - mx);
          lse = mx + logf(sum);
          contrib = (lse - x[target]) * mj;
        }
      }
      const float scale = mj / (float)B;
      for (int j = 0; j < w; ++j) {
        float g = y[j] * (d[j] - dot) * inv_tau;
        if (mj != 0.f) g += scale * (expf(x[j] - lse) - (j == target ? 1.f : 0.f));
        o[j] = accumulate ? o[j] + g : g;
      }
    }
  }
  if (loss_out) {
    float tot = block_sum(contrib, red);
    if (threadIdx.x == 0 && tot != 0.f) atomicAdd(loss_out, tot / (float)B);
  }
}

extern "C" int ftg_activate_bwd(const float* dout, int lddo, const float* act, int lda, int B, int n_spans,
                                const int* span_start, const int* span_width, const int* span_kind, float* dfake,
                                int ldf, float tau, int accumulate, ftg_stream_t stream) {
  long long n_work = (long long)B * n_spans;
  ftg_launch(k_activate_bwd, dim3(ftg_cdiv(n_work, 128)), dim3(128), 0, (cudaStream_t)stream, dout, lddo, act, lda, B,
             n_spans, span_start, span_width, span_kind, dfake, ldf, 1.f / tau, accumulate, nullptr, 0, nullptr, 0, nullptr,
             0, nullptr, nullptr, nullptr);
  FTG_LAUNCH_CHECK();
  return 0;
}

// backward of apply_activate + cond_loss value and gradient, one launch (+ one 4-byte memset)
extern "C" int ftg_gen_output_bwd(const float* dout, int lddo, const float* act, int lda, int B, int n_spans,
                                  const int* span_start, const int* span_width, const int* span_kind, float* dfake,
                                  int ldf, float tau, const float* fake, int ldfk, const float* cond, int ldc,
                                  const float* mask, int ldm, const int* col_soft, const int* soft_cstart,
                                  float* loss_out, ftg_stream_t stream) {
  cudaError_t e = cudaMemsetAsync(loss_out, 0, sizeof(float), (cudaStream_t)stream);
  if (e != cudaSuccess) return (int)e;
  long long n_work = (long long)B * n_spans;
  ftg_launch(k_activate_bwd, dim3(ftg_cdiv(n_work, 128)), dim3(128), 0, (cudaStream_t)stream, dout, lddo, act, lda, B,
             n_spans, span_start, span_width, span_kind, dfake, ldf, 1.f / tau, 0, fake, ldfk, cond, ldc, mask, ldm,
             col_soft, soft_cstart, loss_out);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ----------------------------------------------------------------- cond_loss
__global__ void k_zero2d(float* p, int ld, int rows, int cols) {
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e < (long long)rows * cols) p[(e / cols) * (long long)ld + (e % cols)] = 0.f;
}

__global__ void k_cond_loss(const float* __restrict__ fake, int ldf, const float* __restrict__ cond, int ldc,
                            const float* __restrict__ mask, int ldm, int B, int n_col,
                            const int* __restrict__ soft_start, const int* __restrict__ soft_width,
                            const int* __restrict__ soft_cstart, float* dfake, int lddf, float* loss_out) {
  FTG_PDL_ENTER();
  __shared__ float red[32];
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float contrib = 0.f;
  if (t < (long long)B * n_col) {
    int row = (int)(t / n_col), j = (int)(t % n_col);
    float mj = mask[(size_t)row * ldm + j];
    if (mj != 0.f) {
      int st = soft_start[j], w = soft_width[j], cst = soft_cstart[j];
      const float* x = fake + (size_t)row * ldf + st;
      const float* c = cond + (size_t)row * ldc + cst;
      int target = 0;
      float best = c[0], mx = x[0];
      for (int k = 1; k < w; ++k) {
        if (c[k] > best) { best = c[k]; target = k; }
        mx = fmaxf(mx, x[k]);
      }
      float sum = 0.f;
      for (int k = 0; k < w; ++k) sum += expf(x[k] - mx);
      float lse = mx + logf(sum);
      contrib = (lse - x[target]) * mj;
      float scale = mj / (float)B;
      float* o = dfake + (size_t)row * lddf + st;
      for (int k = 0; k < w; ++k) o[k] += scale * (expf(x[k] - lse) - (k == target ? 1.f : 0.f));
    }
  }
  float tot = block_sum(contrib, red);
  if (threadIdx.x == 0 && tot != 0.f) atomicAdd(loss_out, tot / (float)B);
}

extern "C" int ftg_cond_loss(const float* fake, int ldf, const float* cond, int ldc, const float* mask, int ldm,
                             int B, int dd, int n_col, const int* soft_start, const int* soft_width,
                             const int* soft_cstart, float* dfake, int lddf, float* loss_out, int accumulate,
                             ftg_stream_t stream) {
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(loss_out, 0, sizeof(float), s);
  if (e != cudaSuccess) return (int)e;
  if (!accumulate) {
    k_zero2d<<<ftg_cdiv((long long)B * dd, 256), 256, 0, s>>>(dfake, lddf, B, dd);
    ++g_ftg_launch_count;
  }
  long long n_work = (long long)B * n_col;
  ftg_launch(k_cond_loss, dim3(ftg_cdiv(n_work, 128)), dim3(128), 0, s, fake, ldf, cond, ldc, mask, ldm, B, n_col, soft_start, soft_width,
                                                    soft_cstart, dfake, lddf, loss_out);
  FTG_LAUNCH_CHECK();
  return 0;
}

// --------------------------------------------------------------------- slerp
// One warp per row.
__global__ void k_slerp(const float* __restrict__ low, int ldl, const float* __restrict__ high, int ldh, float* out,
                        int ldo, int B, int din, const float* alpha, const int64_t* state, int stream_id) {
  FTG_PDL_ENTER();
  int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (row >= B) return;
  const float* l = low + (size_t)row * ldl;
  const float* h = high + (size_t)row * ldh;
  float ll = 0.f, hh = 0.f, lh = 0.f;
  for (int j = lane; j < din; j += 32) {
    float a = l[j], b = h[j];
    ll += a * a; hh += b * b; lh += a * b;
  }
  ll = warp_sum(ll); hh = warp_sum(hh); lh = warp_sum(lh);
  float cosv = lh / (sqrtf(ll) * sqrtf(hh));
  float omega = acosf(cosv);
  float so = sinf(omega);
  float a = alpha ? alpha[row] : philox_uniform(state, stream_id, (uint32_t)row);
  float wl = sinf((1.f - a) * omega) / so, wh = sinf(a * omega) / so;
  float* o = out + (size_t)row * ldo;
  for (int j = lane; j < din; j += 32) o[j] = wl * l[j] + wh * h[j];
}

extern "C" int ftg_slerp(const float* low, int ldl, const float* high, int ldh, float* out, int ldo, int B, int din,
                         const float* alpha, const int64_t* rng_state, int rng_stream, ftg_stream_t stream) {
  ftg_launch(k_slerp, dim3(ftg_cdiv(B, 4)), dim3(128), 0, (cudaStream_t)stream, low, ldl, high, ldh, out, ldo, B, din, alpha, rng_state,
                                                            rng_stream);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------- gp_penalty
// One CTA per pack row: ||g||, penalty contribution, and in-place d pen / d g.
__global__ void k_gp_penalty(float* g, int ldg, int Bp, int Din, float* pen_out, float lambda_) {
  FTG_PDL_ENTER();
  __shared__ float red[32];
  float* row = g + (size_t)blockIdx.x * ldg;
  float s = 0.f;
  for (int j = threadIdx.x; j < Din; j += blockDim.x) {
    float v = row[j];
    s += v * v;
  }
  float nrm = sqrtf(block_sum(s, red));
  float coef = (2.f * lambda_ / (float)Bp) * (nrm - 1.f) / fmaxf(nrm, 1e-30f);
  for (int j = threadIdx.x; j < Din; j += blockDim.x) row[j] *= coef;
  if (threadIdx.x == 0) atomicAdd(pen_out, lambda_ * (nrm - 1.f) * (nrm - 1.f) / (float)Bp);
}

extern "C" int ftg_gp_penalty(float* g, int ldg, int Bp, int Din, float* pen_out, float lambda_, ftg_stream_t stream) {
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(pen_out, 0, sizeof(float), s);
  if (e != cudaSuccess) return (int)e;
  ftg_launch(k_gp_penalty, dim3(Bp), dim3(256), 0, s, g, ldg, Bp, Din, pen_out, lambda_);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------- adam
__global__ void k_adam(float* __restrict__ param, const float* __restrict__ grad, float* __restrict__ m,
                       float* __restrict__ v, long long n, const int64_t* state, float lr, float b1, float b2,
                       float eps, float wd) {
  FTG_PDL_ENTER();
  float t = (float)state[1];
  float bc1 = 1.f - powf(b1, t), bc2 = 1.f - powf(b2, t);
  float step = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    float4 p = *reinterpret_cast<float4*>(param + i);
    float4 g = *reinterpret_cast<const float4*>(grad + i);
    float4 mm = *reinterpret_cast<float4*>(m + i);
    float4 vv = *reinterpret_cast<float4*>(v + i);
    float* pp = &p.x; float* gg = &g.x; float* pm = &mm.x; float* pv = &vv.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float gj = gg[j] + wd * pp[j];
      pm[j] = b1 * pm[j] + (1.f - b1) * gj;
      pv[j] = b2 * pv[j] + (1.f - b2) * gj * gj;
      pp[j] -= step * pm[j] / (sqrtf(pv[j]) * inv_sqrt_bc2 + eps);
    }
    *reinterpret_cast<float4*>(param + i) = p;
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
  } else {
    for (long long e = i; e < n; ++e) {
      float gj = grad[e] + wd * param[e];
      m[e] = b1 * m[e] + (1.f - b1) * gj;
      v[e] = b2 * v[e] + (1.f - b2) * gj * gj;
      param[e] -= step * m[e] / (sqrtf(v[e]) * inv_sqrt_bc2 + eps);
    }
  }
}

extern "C" int ftg_adam(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n,
                        const int64_t* rng_state, float lr, float beta1, float beta2, float eps, float weight_decay,
                        ftg_stream_t stream) {
  if (n <= 0) return 0;
  ftg_launch(k_adam, dim3(ftg_cdiv((n + 3) / 4, 256)), dim3(256), 0, (cudaStream_t)stream, param, grad, exp_avg, exp_avg_sq, n, rng_state,
                                                                       lr, beta1, beta2, eps, weight_decay);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------- vgm_decode
// One thread per (row, table column).
__global__ void k_vgm_decode(const float* __restrict__ act, int lda, int rows, int n_columns,
                             const int* __restrict__ col_is_cont, const int* __restrict__ col_start,
                             const int* __restrict__ col_width, const int* __restrict__ col_aux,
                             const float* __restrict__ mu, const float* __restrict__ sd, int n_clusters,
                             const float* __restrict__ cat_ids, float* out, int ldo) {
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)rows * n_columns) return;
  int row = (int)(t / n_columns), c = (int)(t % n_columns);
  const float* x = act + (size_t)row * lda + col_start[c];
  int w = col_width[c];
  float res;
  if (col_is_cont[c]) {
    float u = fminf(fmaxf(x[0], -1.f), 1.f);
    int pick = 0;
    float best = x[1];
    for (int k = 1; k < w; ++k)
      if (x[1 + k] > best) { best = x[1 + k]; pick = k; }
    int ci = col_aux[c];
    res = u * 4.f * sd[ci * n_clusters + pick] + mu[ci * n_clusters + pick];
  } else {
    int pick = 0;
    float best = x[0];
    for (int k = 1; k < w; ++k)
      if (x[k] > best) { best = x[k]; pick = k; }
    res = cat_ids[col_aux[c] + pick];
  }
  out[(size_t)row * ldo + c] = res;
}

extern "C" int ftg_vgm_decode(const float* act, int lda, int rows, int n_columns, const int* col_is_cont,
                              const int* col_start, const int* col_width, const int* col_aux, const float* mu,
                              const float* sd, int n_clusters, const float* cat_ids, float* out, int ldo,
                              ftg_stream_t stream) {
  long long n_work = (long long)rows * n_columns;
  k_vgm_decode<<<ftg_cdiv(n_work, 256), 256, 0, (cudaStream_t)stream>>>(act, lda, rows, n_columns, col_is_cont,
                                                                        col_start, col_width, col_aux, mu, sd,
                                                                        n_clusters, cat_ids, out, ldo);
  FTG_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------- vgm_encode
__global__ void k_vgm_encode(const float* __restrict__ x, int rows, const float* __restrict__ logit_const,
                             const float* __restrict__ mu, const float* __restrict__ var, int k_active, const float* u,
                             const int64_t* state, int stream_id, float* val_out, int val_stride, int* code_out,
                             int code_stride) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows) return;
  float xi = x[i];
  float lg[16];
  float mx = -3.0e38f;
  for (int k = 0; k < k_active; ++k) {
    float d = xi - mu[k];
    lg[k] = logit_const[k] - 0.5f * d * d / var[k];
    mx = fmaxf(mx, lg[k]);
  }
  float sum = 0.f;
  for (int k = 0; k < k_active; ++k) { lg[k] = expf(lg[k] - mx); sum += lg[k]; }
  float tot = 0.f;
  for (int k = 0; k < k_active; ++k) { lg[k] = lg[k] / sum + 1e-6f; tot += lg[k]; }
  float uu = u ? u[i] : philox_uniform(state, stream_id, (uint32_t)i);
  float acc = 0.f;
  int sel = k_active - 1;
  for (int k = 0; k < k_active; ++k) {
    acc += lg[k] / tot;
    if (acc > uu) { sel = k; break; }
  }
  float v = (xi - mu[sel]) / (4.f * sqrtf(var[sel]));
  val_out[(size_t)i * val_stride] = fminf(fmaxf(v, -0.99f), 0.99f);
  code_out[(size_t)i * code_stride] = sel;
}

extern "C" int ftg_vgm_encode(const float* x, int rows, const float* logit_const, const float* mu, const float* var,
                              int k_active, const float* u, const int64_t* rng_state, int rng_stream, float* val_out,
                              int val_stride, int* code_out, int code_stride, ftg_stream_t stream) {
  if (k_active > 16) return FTG_ERR_UNSUPPORTED;
  k_vgm_encode<<<ftg_cdiv(rows, 256), 256, 0, (cudaStream_t)stream>>>(x, rows, logit_const, mu, var, k_active, u,
                                                                      rng_state, rng_stream, val_out, val_stride,
                                                                      code_out, code_stride);
  FTG_LAUNCH_CHECK();
  return 0;
}
