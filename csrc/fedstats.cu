// Federated initialisation statistics on the device (SURVEY K14 second half + K15).
//
// Reference: MDGANServer.uniform_meta_category (JSD), uniform_continuous_gmm (GMM sampling, pooled sample,
// Wasserstein distance) and calculate_final_weights_for_aggregation
// (Server/dtds/distributed.py:592-686, 689-765, 767-783).  There the server pulls pickled sklearn mixtures
// and count dictionaries over RPC, calls `client_gmm.sample(int(sum_rows * frac_i))`, concatenates, and
// runs scipy's sort-based wasserstein_distance / jensenshannon per (client, column) on the CPU.
//
// Here every client keeps, in peer-visible device memory (CUDA-IPC block, parallel/symm.py):
//   * its category histogram in global label order            (floats)
//   * one SORTED sample of each of its continuous columns' mixtures, drawn on the device
//     (k_gmm_sample_sort: Philox draw + in-kernel bitonic sort, one CTA per column)
// and ONE pair of kernels per rank reads all peers' histograms and sorted samples straight over NVLink
// (plain ld.global on peer pointers) and leaves the normalised distances and the soft-maxed aggregation
// weights w[N] in device memory, where the fedavg kernels read them (no host round trip):
//   * JSD: as scipy.spatial.distance.jensenshannon (sqrt, natural log), fp64 accumulation
//   * WD : sum over consecutive pooled points of |F_pool - F_i| dx.  The pooled order is never
//          materialised: for a sample x of client j, upper_bound(x) in every client's sorted sample gives
//          F_i(x), the pooled rank (sum of the bounds) and the next pooled point (min over the successors).
//          The same ranks select the pooled quantiles that seed the global mixture refit.
// All reductions run in a fixed order: every rank computes bit-identical weights.
#include "common.cuh"

namespace {

constexpr int FS_MAX_WORLD = 16;
constexpr int SORT_TILE = 4096;          // elements per shared-memory sorting tile (16 KB)
constexpr int SORT_THREADS = 1024;
constexpr int WD_THREADS = 256;

struct FedPtrs {
  const float* hist[FS_MAX_WORLD];       // [hist_len] per client
  const float* seg[FS_MAX_WORLD];        // [n_cont][n_pad_i] sorted samples per client
  int n[FS_MAX_WORLD];                   // samples per column of client i
  int n_pad[FS_MAX_WORLD];               // row pitch of seg[i]
  float rows[FS_MAX_WORLD];
};

// ---------------------------------------------------------------- K14: sample + sort
__device__ __forceinline__ void cmpx(float& a, float& b, bool asc) {
  if ((a > b) == asc) { const float t = a; a = b; b = t; }
}

// bitonic steps j = j_hi, j_hi/2, ..., 1 of merge size k on one shared-memory tile whose first element
// has global index `g0` (the sort direction of a pair depends on the GLOBAL index)
__device__ __forceinline__ void tile_steps(float* tile, int g0, int k, int j_hi) {
  for (int j = j_hi; j > 0; j >>= 1) {
    for (int pidx = threadIdx.x; pidx < SORT_TILE / 2; pidx += SORT_THREADS) {
      const int i = 2 * j * (pidx / j) + (pidx % j);
      cmpx(tile[i], tile[i + j], ((g0 + i) & k) == 0);
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(SORT_THREADS)
k_gmm_sample_sort(const double* __restrict__ gmm, int K, int n, int n_pad, unsigned long long seed, int client_id,
                  float* __restrict__ out, const float* __restrict__ src) {
  __shared__ float tile[SORT_TILE];
  __shared__ float cw[16], mu[16], sd[16];
  const int c = blockIdx.x;
  const double* g = gmm + (size_t)c * 7 * K;            // rows: means, covs, weights, ...
  if (threadIdx.x == 0 && src == nullptr) {
    double tot = 0.0;
    for (int k = 0; k < K; ++k) tot += g[2 * K + k];
    double acc = 0.0;
    for (int k = 0; k < K; ++k) {
      acc += g[2 * K + k] / tot;
      cw[k] = (float)acc;
      mu[k] = (float)g[k];
      sd[k] = (float)sqrt(g[K + k]);
    }
    cw[K - 1] = 2.f;                                     // guard against round-off in the last bucket
  }
  __syncthreads();
  float* seg = out + (size_t)c * n_pad;
  // ---- draw: component by inverse CDF, value by Box-Muller (sklearn BaseMixture.sample semantics:
  //      multinomial component counts emerge from iid component draws)
  for (int idx = threadIdx.x; idx < n_pad; idx += SORT_THREADS) {
    float v = INFINITY;                                  // padding sorts to the end
    if (idx < n && src != nullptr) {
      v = src[(size_t)c * n + idx];                      // plain row sort (ftg_sort_rows)
    } else if (idx < n) {
      const Philox4 r = philox4x32_10(seed, (unsigned long long)client_id, 0x51u + (uint32_t)c, (uint32_t)idx);
      const float u = u01(r.x);
      int k = 0;
      while (k < K - 1 && u >= cw[k]) ++k;
      const float u1 = fmaxf(u01(r.y), 1e-12f), u2 = u01(r.z);
      const float z = sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);
      v = mu[k] + sd[k] * z;
    }
    seg[idx] = v;
  }
  __syncthreads();
  // ---- bitonic sort of n_pad (power of two, >= SORT_TILE) elements by this one CTA
  for (int t0 = 0; t0 < n_pad; t0 += SORT_TILE) {        // every tile fully sorted in shared memory
    for (int i = threadIdx.x; i < SORT_TILE; i += SORT_THREADS) tile[i] = seg[t0 + i];
    __syncthreads();
    for (int k = 2; k <= SORT_TILE; k <<= 1) tile_steps(tile, t0, k, k >> 1);
    for (int i = threadIdx.x; i < SORT_TILE; i += SORT_THREADS) seg[t0 + i] = tile[i];
    __syncthreads();
  }
  for (int k = 2 * SORT_TILE; k <= n_pad; k <<= 1) {
    for (int j = k >> 1; j >= SORT_TILE; j >>= 1) {      // long-distance steps in global memory
      for (int pidx = threadIdx.x; pidx < n_pad / 2; pidx += SORT_THREADS) {
        const int i = 2 * j * (pidx / j) + (pidx % j);
        float a = seg[i], b = seg[i + j];
        const bool asc = (i & k) == 0;
        if ((a > b) == asc) { seg[i] = b; seg[i + j] = a; }
      }
      __syncthreads();
    }
    for (int t0 = 0; t0 < n_pad; t0 += SORT_TILE) {      // remaining steps tile by tile in shared memory
      for (int i = threadIdx.x; i < SORT_TILE; i += SORT_THREADS) tile[i] = seg[t0 + i];
      __syncthreads();
      tile_steps(tile, t0, k, SORT_TILE >> 1);
      for (int i = threadIdx.x; i < SORT_TILE; i += SORT_THREADS) seg[t0 + i] = tile[i];
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------- K15: distances over peer memory
__device__ __forceinline__ double block_sum_d(double v, double* dred) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) dred[wid] = v;
  __syncthreads();
  double t = 0.0;
  for (int w2 = 0; w2 < nw; ++w2) t += dred[w2];
  return t;
}

// grid (world, n_cat): Jensen-Shannon distance between the pooled histogram and client i's
__global__ void __launch_bounds__(256)
k_fed_jsd(FedPtrs P, int world, int n_cat, const int* __restrict__ cat_off, float* jsd) {
  __shared__ double dred[32];
  const int i = blockIdx.x, c = blockIdx.y;
  const int lo = cat_off[c], w = cat_off[c + 1] - lo;
  double tp = 0.0, tq = 0.0;
  for (int k = threadIdx.x; k < w; k += blockDim.x) {
    double g = 0.0;
    for (int r = 0; r < world; ++r) g += P.hist[r][lo + k];           // peer loads
    tq += g;
    tp += P.hist[i][lo + k];
  }
  tp = block_sum_d(tp, dred);
  tq = block_sum_d(tq, dred);
  double acc = 0.0;
  for (int k = threadIdx.x; k < w; k += blockDim.x) {
    double g = 0.0;
    for (int r = 0; r < world; ++r) g += P.hist[r][lo + k];
    const double q = g / tq, p = (double)P.hist[i][lo + k] / tp;
    const double m = 0.5 * (p + q);
    if (p > 0.0) acc += p * log(p / m);
    if (q > 0.0) acc += q * log(q / m);
  }
  acc = block_sum_d(acc, dred);
  if (threadIdx.x == 0) jsd[(size_t)i * n_cat + c] = (float)sqrt(fmax(0.5 * acc, 0.0));
}

__device__ __forceinline__ int upper_bound(const float* __restrict__ s, int n, float x) {
  int a = 0, b = n;
  while (a < b) {
    const int mid = (a + b) >> 1;
    if (s[mid] <= x) a = mid + 1; else b = mid;
  }
  return a;
}

// grid (chunks, n_cont).  Pooled element e = (client j, position t).  partial: [n_cont][chunks][world].
__global__ void __launch_bounds__(WD_THREADS)
k_fed_wd(FedPtrs P, int world, int n_total, int chunks, float* __restrict__ partial, float* __restrict__ centers,
         int K, float* __restrict__ pool_out) {
  __shared__ float red[32];
  const int c = blockIdx.y, chunk = blockIdx.x;
  const int per = (n_total + chunks - 1) / chunks;
  const int e0 = chunk * per, e1 = min(n_total, e0 + per);
  const float* seg[FS_MAX_WORLD];
  int start[FS_MAX_WORLD + 1];
  start[0] = 0;
#pragma unroll
  for (int i = 0; i < FS_MAX_WORLD; ++i) {
    seg[i] = i < world ? P.seg[i] + (size_t)c * P.n_pad[i] : nullptr;
    start[i + 1] = start[i] + (i < world ? P.n[i] : 0);
  }
  float acc[FS_MAX_WORLD];
#pragma unroll
  for (int i = 0; i < FS_MAX_WORLD; ++i) acc[i] = 0.f;
  const float invN = 1.f / (float)n_total;
  for (int e = e0 + threadIdx.x; e < e1; e += WD_THREADS) {
    int j = 0;
    while (j + 1 < world && e >= start[j + 1]) ++j;
    const int t = e - start[j];
    const float x = seg[j][t];
    if (pool_out) pool_out[(size_t)c * n_total + e] = x;               // pooled sample (client order) for the global refit
    int ub[FS_MAX_WORLD];
    int rank = 0;
    float next = INFINITY;
    bool rep = true;
#pragma unroll
    for (int i = 0; i < FS_MAX_WORLD; ++i) {
      ub[i] = 0;
      if (i < world) {
        const int a = upper_bound(seg[i], P.n[i], x);
        ub[i] = a;
        rank += a;
        if (a < P.n[i]) next = fminf(next, seg[i][a]);
        const bool has = a > 0 && seg[i][a - 1] == x;
        if (i == j) { if (t != a - 1) rep = false; }                   // not the last copy in its own sample
        else if (i > j && has) rep = false;                            // a later client holds the same value
      }
    }
    if (!rep) continue;                                                // every distinct value is counted once
    if (next < INFINITY) {
      const float dx = next - x, fp = (float)rank * invN;
#pragma unroll
      for (int i = 0; i < FS_MAX_WORLD; ++i)
        if (i < world) acc[i] += fabsf(fp - (float)ub[i] / (float)P.n[i]) * dx;
    }
    if (centers) {
      // pooled positions [rank - copies, rank) hold this value: seed the refit with the pooled quantiles
      int copies = 0;
      for (int i = 0; i < world; ++i) {
        int lb = ub[i];
        while (lb > 0 && seg[i][lb - 1] == x) --lb;
        copies += ub[i] - lb;
      }
      for (int k = 0; k < K; ++k) {
        const int q = (int)llrint(((double)k + 0.5) / (double)K * (double)(n_total - 1));
        if (rank - copies <= q && q < rank) centers[(size_t)c * K + k] = x;
      }
    }
  }
  for (int i = 0; i < world; ++i) {
    const float s = block_sum(acc[i], red);
    if (threadIdx.x == 0) partial[((size_t)c * chunks + chunk) * world + i] = s;
  }
}

// one CTA: fold the partials (fixed order), normalise columns over clients, soft-max
__global__ void k_fed_finish(int world, int n_cat, int n_cont, int chunks, FedPtrs P, const float* __restrict__ partial,
                             float* jsd, float* wd, float* weights) {
  __shared__ float sd[64];
  for (int idx = threadIdx.x; idx < n_cont * world; idx += blockDim.x) {
    const int c = idx / world, i = idx % world;
    float s = 0.f;
    for (int ch = 0; ch < chunks; ++ch) s += partial[((size_t)c * chunks + ch) * world + i];
    wd[(size_t)i * n_cont + c] = s;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < n_cat; c += blockDim.x) {
    float t = 0.f;
    for (int i = 0; i < world; ++i) t += jsd[(size_t)i * n_cat + c];
    for (int i = 0; i < world; ++i) jsd[(size_t)i * n_cat + c] = (t != 0.f) ? jsd[(size_t)i * n_cat + c] / t : 1.f / world;
  }
  for (int c = threadIdx.x; c < n_cont; c += blockDim.x) {
    float t = 0.f;
    for (int i = 0; i < world; ++i) t += wd[(size_t)i * n_cont + c];
    if (t != 0.f)
      for (int i = 0; i < world; ++i) wd[(size_t)i * n_cont + c] /= t;
  }
  __syncthreads();
  if (threadIdx.x < world) {
    float s = 0.f;
    for (int c = 0; c < n_cat; ++c) s += jsd[(size_t)threadIdx.x * n_cat + c];
    for (int c = 0; c < n_cont; ++c) s += wd[(size_t)threadIdx.x * n_cont + c];
    sd[threadIdx.x] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f, rt = 0.f, ex[FS_MAX_WORLD], es = 0.f;
    for (int i = 0; i < world; ++i) { tot += sd[i]; rt += P.rows[i]; }
    for (int i = 0; i < world; ++i) {
      const float w = (1.f - (tot != 0.f ? sd[i] / tot : 0.f)) * P.rows[i] / rt;
      ex[i] = expf(w);
      es += ex[i];
    }
    for (int i = 0; i < world; ++i) weights[i] = ex[i] / es;
  }
}

}  // namespace

extern "C" int ftg_sort_pad(int n) {
  int p = SORT_TILE;
  while (p < n) p <<= 1;
  return p;
}

extern "C" int ftg_gmm_sample_sort(const double* gmm, int C, int K, int n, unsigned long long seed, int client_id,
                                   float* out, ftg_stream_t stream) {
  if (C <= 0) return 0;
  if (K < 1 || K > 16 || n < 1) return FTG_ERR_UNSUPPORTED;
  k_gmm_sample_sort<<<C, SORT_THREADS, 0, (cudaStream_t)stream>>>(gmm, K, n, ftg_sort_pad(n), seed, client_id, out,
                                                                  nullptr);
  FTG_LAUNCH_CHECK();
  return 0;
}

// Sort every row of x [C][n] ascending into out [C][ftg_sort_pad(n)] (padding +inf): the same in-kernel
// bitonic network, used for the quantile seeds of the per-client mixture fits.
extern "C" int ftg_sort_rows(const float* x, int C, int n, float* out, ftg_stream_t stream) {
  if (C <= 0) return 0;
  if (n < 1) return FTG_ERR_UNSUPPORTED;
  k_gmm_sample_sort<<<C, SORT_THREADS, 0, (cudaStream_t)stream>>>(nullptr, 1, n, ftg_sort_pad(n), 0ull, 0, out, x);
  FTG_LAUNCH_CHECK();
  return 0;
}

extern "C" int ftg_fed_stats_to_weights(const float* const* hist, const float* const* seg, const int* n_i,
                                        const float* rows, int world, int n_cat, const int* cat_off, int n_cont,
                                        float* partial, int chunks, float* centers, int K, float* pool_out,
                                        float* jsd_out, float* wd_out, float* weights_out, ftg_stream_t stream) {
  if (world < 1 || world > FS_MAX_WORLD) return FTG_ERR_UNSUPPORTED;
  FedPtrs P;
  long long n_total = 0;
  for (int i = 0; i < FS_MAX_WORLD; ++i) {
    P.hist[i] = i < world && hist ? hist[i] : nullptr;
    P.seg[i] = i < world && seg ? seg[i] : nullptr;
    P.n[i] = i < world && n_i ? n_i[i] : 0;
    P.n_pad[i] = i < world && n_i ? ftg_sort_pad(n_i[i]) : 0;
    P.rows[i] = i < world ? rows[i] : 0.f;
    n_total += P.n[i];
  }
  if (n_total > 0x7fffffff) return FTG_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  if (n_cat > 0) {
    k_fed_jsd<<<dim3(world, n_cat), 256, 0, s>>>(P, world, n_cat, cat_off, jsd_out);
    ++g_ftg_launch_count;
  }
  if (n_cont > 0) {
    if (chunks < 1 || !partial) return FTG_ERR_UNSUPPORTED;
    k_fed_wd<<<dim3(chunks, n_cont), WD_THREADS, 0, s>>>(P, world, (int)n_total, chunks, partial, centers, K, pool_out);
    ++g_ftg_launch_count;
  }
  k_fed_finish<<<1, 128, 0, s>>>(world, n_cat, n_cont, chunks, P, partial, jsd_out, wd_out, weights_out);
  FTG_LAUNCH_CHECK();
  return 0;
}
