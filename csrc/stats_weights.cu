// Similarity statistics -> aggregation weights, on the device.
//
// Restates MDGANServer.uniform_meta_category (JSD part), uniform_continuous_gmm (WD part)
// and calculate_final_weights_for_aggregation (Server/dtds/distributed.py:592-783):
//   JSD[i,c]  Jensen-Shannon distance (sqrt, natural log) between the pooled category
//             histogram of column c and client i's histogram
//   WD[i,c]   1-D Wasserstein distance between the pooled GMM sample of column c and
//             client i's sample (both sorted):  sum_k |F_pool(x_k) - F_i(x_k)| (x_{k+1}-x_k)
//   per-column normalisation over clients (zero-sum columns: 1/N categorical, 0 continuous)
//   w_i = (1 - SD_i / sum SD) * rows_i / sum rows ;  weights = softmax(w)
// Kernel 1: one CTA per (client, column).  Kernel 2: one CTA finishes the weights, so the
// result lands in device memory that the fedavg kernel reads directly.
#include "common.cuh"

namespace {

__global__ void k_distances(const float* __restrict__ hist, int world, int n_cat, const int* __restrict__ cat_off,
                            const float* __restrict__ pool_sorted, int n_total, const float* __restrict__ client_sorted,
                            const int* __restrict__ smp_off, int n_cont, float* jsd, float* wd) {
  __shared__ float red[32];
  const int i = blockIdx.x, c = blockIdx.y;
  if (c < n_cat) {
    const int lo = cat_off[c], w = cat_off[c + 1] - lo;
    const int hist_len = cat_off[n_cat];
    // totals (fp64: near-identical client distributions make the divergence a difference of
    // almost equal terms, fp32 loses it and the per-column normalisation amplifies the error)
    __shared__ double dred[32];
    auto block_sum_d = [&](double v) {
      const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      __syncthreads();
      if (lane == 0) dred[wid] = v;
      __syncthreads();
      double t = 0.0;
      for (int w2 = 0; w2 < nw; ++w2) t += dred[w2];
      return t;
    };
    double tp = 0.0, tq = 0.0;
    for (int k = threadIdx.x; k < w; k += blockDim.x) {
      double g = 0.0;
      for (int r = 0; r < world; ++r) g += hist[(size_t)r * hist_len + lo + k];
      tq += g;
      tp += hist[(size_t)i * hist_len + lo + k];
    }
    tp = block_sum_d(tp);
    tq = block_sum_d(tq);
    double acc = 0.0;
    for (int k = threadIdx.x; k < w; k += blockDim.x) {
      double g = 0.0;
      for (int r = 0; r < world; ++r) g += hist[(size_t)r * hist_len + lo + k];
      const double q = g / tq, p = (double)hist[(size_t)i * hist_len + lo + k] / tp;
      const double m = 0.5 * (p + q);
      if (p > 0.0) acc += p * log(p / m);
      if (q > 0.0) acc += q * log(q / m);
    }
    acc = block_sum_d(acc);
    if (threadIdx.x == 0) jsd[(size_t)i * n_cat + c] = (float)sqrt(fmax(0.5 * acc, 0.0));
  } else {
    const int cc = c - n_cat;
    const float* pool = pool_sorted + (size_t)cc * n_total;
    const int s0 = smp_off[cc * world + i], ni = smp_off[cc * world + i + 1] - s0;
    const float* mine = client_sorted + s0;
    float acc = 0.f;
    if (ni > 0) {
      for (int k = threadIdx.x; k + 1 < n_total; k += blockDim.x) {
        float x = pool[k], dx = pool[k + 1] - x;
        if (dx <= 0.f) continue;
        // #mine <= x  (upper bound)
        int a = 0, b = ni;
        while (a < b) {
          int mid = (a + b) >> 1;
          if (mine[mid] <= x) a = mid + 1; else b = mid;
        }
        float fp = (float)(k + 1) / (float)n_total, fi = (float)a / (float)ni;
        acc += fabsf(fp - fi) * dx;
      }
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) wd[(size_t)i * n_cont + cc] = acc;
  }
}

__global__ void k_finish(int world, int n_cat, int n_cont, const float* __restrict__ rows, float* jsd, float* wd,
                         float* weights) {
  __shared__ float sd[64];
  // normalise columns (thread per column)
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
    float tot = 0.f, rt = 0.f, ex[64], es = 0.f;
    for (int i = 0; i < world; ++i) { tot += sd[i]; rt += rows[i]; }
    for (int i = 0; i < world; ++i) {
      float w = (1.f - (tot != 0.f ? sd[i] / tot : 0.f)) * rows[i] / rt;
      ex[i] = expf(w);
      es += ex[i];
    }
    for (int i = 0; i < world; ++i) weights[i] = ex[i] / es;
  }
}

}  // namespace

extern "C" int ftg_stats_to_weights(const float* hist, int world, int n_cat, const int* cat_off,
                                    const float* pool_sorted, int n_total, const float* client_sorted,
                                    const int* smp_off, int n_cont, const float* rows, float* jsd_out, float* wd_out,
                                    float* weights_out, ftg_stream_t stream) {
  if (world < 1 || world > 64) return FTG_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  if (n_cat + n_cont > 0) {
    k_distances<<<dim3(world, n_cat + n_cont), 256, 0, s>>>(hist, world, n_cat, cat_off, pool_sorted, n_total,
                                                            client_sorted, smp_off, n_cont, jsd_out, wd_out);
    ++g_ftg_launch_count;
  }
  k_finish<<<1, 128, 0, s>>>(world, n_cat, n_cont, rows, jsd_out, wd_out, weights_out);
  FTG_LAUNCH_CHECK();
  return 0;
}
