// Fused batch preparation: everything the reference does on the host
// with numpy + Python loops before each D / G step (Cond.sample, shuffle, Sampler.sample,
// torch.normal, torch.cat, H2D copies; Server/dtds/distributed.py:335-352,381-390,
// ctgan.py:147-161,221-228) in at most two launches, with Philox randomness keyed by the
// device-resident (seed, step).
//
//   k_prep_select  one CTA: conditional (column, option) per row, a random permutation of
//                  the batch (bitonic argsort of iid uniforms) and, for the permuted
//                  conditions, a uniformly random matching training row from the CSR index
//   k_prep_rows    one CTA per row: noise z ~ N(0,1), the one-hot cond vector c1 written
//                  straight into the generator input window, the column mask m1 (G step),
//                  and (D step) the real row expanded from the compact table into the
//                  discriminator input with its permuted cond vector c2 appended
#include "common.cuh"

namespace {

__device__ __forceinline__ void sample_col_opt(const float* __restrict__ cum_p, int n_col, int max_w,
                                               const int* __restrict__ soft_width, const int64_t* state, int stream,
                                               int i, int& c, int& o) {
  float uc = philox_uniform(state, stream, (uint32_t)i);
  float uo = philox_uniform(state, stream + 1, (uint32_t)i);
  c = min((int)(uc * n_col), n_col - 1);
  int w = soft_width[c];
  const float* row = cum_p + (size_t)c * max_w;
  o = w - 1;
  for (int j = 0; j < w; ++j)
    if (row[j] > uo) { o = j; break; }
}

__global__ void __launch_bounds__(1024)
k_prep_select(const float* __restrict__ cum_p, int n_col, int max_w, const int* __restrict__ soft_width,
              const int* __restrict__ soft_cstart, const int* __restrict__ row_ptr, const int* __restrict__ row_ids,
              int* col, int* opt, int* perm, int* rid, int B, int P2, int64_t* state, int s_col, int s_perm,
              int s_row, int do_tick) {
  FTG_PDL_ENTER();
  if (do_tick) {                       // first kernel of an iteration: advance the (seed, step) state
    if (threadIdx.x == 0) state[1] += 1;
    __syncthreads();
  }
  extern __shared__ unsigned char smem_raw[];
  float* key = reinterpret_cast<float*>(smem_raw);
  int* idx = reinterpret_cast<int*>(key + P2);
  int* gopt = idx + P2;                               // global option id per (unpermuted) row
  for (int i = threadIdx.x; i < P2; i += blockDim.x) {
    if (i < B) {
      int c, o;
      sample_col_opt(cum_p, n_col, max_w, soft_width, state, s_col, i, c, o);
      col[i] = c;
      opt[i] = o;
      gopt[i] = soft_cstart[c] + o;
      key[i] = philox_uniform(state, s_perm, (uint32_t)i);
    } else {
      key[i] = 3.0e38f;
    }
    idx[i] = i;
  }
  __syncthreads();
  // random permutation = argsort of the iid keys, by rank counting: rank(i) = #{j : key_j < key_i}
  // (B^2 shared-memory broadcast reads, no barriers -- much cheaper than a bitonic network here)
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float ki = key[i];
    int rank = 0;
    for (int j = 0; j < B; ++j) {
      const float kj = key[j];
      rank += (kj < ki) || (kj == ki && j < i);
    }
    idx[rank] = i;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    int src = idx[i];
    perm[i] = src;
    int g = gopt[src];
    int lo = row_ptr[g], cnt = row_ptr[g + 1] - lo;
    float uu = philox_uniform(state, s_row, (uint32_t)i);
    int k = min((int)(uu * cnt), max(cnt - 1, 0));
    rid[i] = row_ids[lo + k];
  }
}

struct PrepRows {
  int mode;                      // 0 = G step (sample col/opt here, write m1), 1 = D step (use select's output)
  int* col; int* opt; const int* perm; const int* rid;
  const int* soft_cstart;
  float* z_out; int ldz; int E;
  float* c_out; int ldc; int n_opt;
  float* m_out; int ldm; int n_col;
  const float* vals; int n_tanh; const int* codes;
  const int* tanh_start; const int* soft_start;
  float* real_out; int ldr; int dd;
  const float* cum_p; int max_w; const int* soft_width;
  const int64_t* state; int s_z; int s_col;
};

__global__ void __launch_bounds__(128) k_prep_rows(const PrepRows P) {
  FTG_PDL_ENTER();
  __shared__ int s_c, s_o;
  const int i = blockIdx.x;
  if (threadIdx.x == 0) {
    int c, o;
    if (P.mode == 0) {
      sample_col_opt(P.cum_p, P.n_col, P.max_w, P.soft_width, P.state, P.s_col, i, c, o);
      P.col[i] = c;
      P.opt[i] = o;
    } else {
      c = P.col[i];
      o = P.opt[i];
    }
    s_c = c;
    s_o = o;
  }
  __syncthreads();
  const int c = s_c, target = P.soft_cstart[c] + s_o;
  // noise: one Philox 4-vector -> 4 normals (Box-Muller)
  for (int q = threadIdx.x; q * 4 < P.E; q += blockDim.x) {
    Philox4 p = philox4x32_10((uint64_t)P.state[0], (uint64_t)P.state[1], (uint32_t)P.s_z, (uint32_t)(i * ((P.E + 3) / 4) + q));
    float u0 = fmaxf(u01(p.x), 5.96e-8f), u1 = u01(p.y), u2 = fmaxf(u01(p.z), 5.96e-8f), u3 = u01(p.w);
    float r0 = sqrtf(-2.f * __logf(u0)), r1 = sqrtf(-2.f * __logf(u2));
    float s0, c0, s1, c1;
    __sincosf(6.283185307179586f * u1, &s0, &c0);
    __sincosf(6.283185307179586f * u3, &s1, &c1);
    float zz[4] = {r0 * c0, r0 * s0, r1 * c1, r1 * s1};
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (q * 4 + j < P.E) P.z_out[(size_t)i * P.ldz + q * 4 + j] = zz[j];
  }
  for (int j = threadIdx.x; j < P.n_opt; j += blockDim.x) P.c_out[(size_t)i * P.ldc + j] = (j == target) ? 1.f : 0.f;
  if (P.mode == 0) {
    for (int j = threadIdx.x; j < P.n_col; j += blockDim.x) P.m_out[(size_t)i * P.ldm + j] = (j == c) ? 1.f : 0.f;
    return;
  }
  // D step: real row i is conditioned on the cond vector of fake row perm[i]
  const int src = P.perm[i];
  const int target2 = P.soft_cstart[P.col[src]] + P.opt[src];
  const int r = P.rid[i];
  float* o = P.real_out + (size_t)i * P.ldr;
  for (int j = threadIdx.x; j < P.dd; j += blockDim.x) o[j] = 0.f;
  for (int j = threadIdx.x; j < P.n_opt; j += blockDim.x) o[P.dd + j] = (j == target2) ? 1.f : 0.f;
  __syncthreads();
  for (int t = threadIdx.x; t < P.n_tanh; t += blockDim.x) o[P.tanh_start[t]] = P.vals[(size_t)r * P.n_tanh + t];
  for (int s = threadIdx.x; s < P.n_col; s += blockDim.x) o[P.soft_start[s] + P.codes[(size_t)r * P.n_col + s]] = 1.f;
}

}  // namespace

extern "C" int ftg_prep_select(const float* cum_p, int n_col, int max_w, const int* soft_width,
                               const int* soft_cstart, const int* row_ptr, const int* row_ids, int* col, int* opt,
                               int* perm, int* rid, int B, int64_t* rng_state, int s_col, int s_perm, int s_row,
                               int do_tick, ftg_stream_t stream) {
  int P2 = 1;
  while (P2 < B) P2 <<= 1;
  if (P2 > 4096) return FTG_ERR_UNSUPPORTED;
  int threads = P2 < 1024 ? (P2 < 32 ? 32 : P2) : 1024;
  ftg_launch(k_prep_select, dim3(1), dim3(threads), (size_t)P2 * 12, (cudaStream_t)stream, cum_p, n_col, max_w, soft_width, soft_cstart,
                                                                      row_ptr, row_ids, col, opt, perm, rid, B, P2,
                                                                      rng_state, s_col, s_perm, s_row, do_tick);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_prep_rows(int mode, int* col, int* opt, const int* perm, const int* rid, const int* soft_cstart,
                             float* z_out, int ldz, int E, float* c_out, int ldc, int n_opt, float* m_out, int ldm,
                             int n_col, const float* vals, int n_tanh, const int* codes, const int* tanh_start,
                             const int* soft_start, float* real_out, int ldr, int dd, const float* cum_p, int max_w,
                             const int* soft_width, int B, const int64_t* rng_state, int s_z, int s_col,
                             ftg_stream_t stream) {
  PrepRows P;
  P.mode = mode; P.col = col; P.opt = opt; P.perm = perm; P.rid = rid; P.soft_cstart = soft_cstart;
  P.z_out = z_out; P.ldz = ldz; P.E = E; P.c_out = c_out; P.ldc = ldc; P.n_opt = n_opt;
  P.m_out = m_out; P.ldm = ldm; P.n_col = n_col; P.vals = vals; P.n_tanh = n_tanh; P.codes = codes;
  P.tanh_start = tanh_start; P.soft_start = soft_start; P.real_out = real_out; P.ldr = ldr; P.dd = dd;
  P.cum_p = cum_p; P.max_w = max_w; P.soft_width = soft_width; P.state = rng_state; P.s_z = s_z; P.s_col = s_col;
  ftg_launch(k_prep_rows, dim3(B), dim3(128), 0, (cudaStream_t)stream, P);
  FTG_LAUNCH_CHECK();
  return 0;
}
