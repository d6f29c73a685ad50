// Variational-Bayes Dirichlet-process GMM fit of many 1-D columns at once.
//
// Same model, priors and stopping rule as sklearn's
//   BayesianGaussianMixture(n_components=K, weight_concentration_prior_type='dirichlet_process',
//                           weight_concentration_prior=dp_prior, n_init=1)
// used by the reference per continuous column on every client and again on the federator
// (Server/dtds/features/transformers.py:334-340, Server/dtds/distributed.py:725-746): one CTA
// fits one column start to finish -- Lloyd iterations from quantile centres (k-means init),
// then E-step / M-step / lower-bound until |delta lower bound| < tol -- with the column streamed
// from L2/HBM each iteration, per-point maths in fp32 and every reduction and the M-step in fp64.
// The oracle is fed_tgan_b200/features/gmm.py::fit_vbgmm_1d (torch restatement of sklearn).
#include "common.cuh"

namespace {

constexpr int KMAX = 12;          // compile-time bound of the component count (registers)
constexpr int VB_THREADS = 512;

__device__ double digamma_d(double x) {
  double r = 0.0;
  while (x < 6.0) { r -= 1.0 / x; x += 1.0; }
  const double f = 1.0 / (x * x);
  return r + log(x) - 0.5 / x - f * (1.0 / 12.0 - f * (1.0 / 120.0 - f * (1.0 / 252.0 - f * (1.0 / 240.0))));
}

// block-wide sum of `cnt` doubles per thread (cnt <= 3*KMAX+1), result in sh_out[0..cnt)
__device__ void block_reduce_vec(const double* v, int cnt, double* sh_part, double* sh_out) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int i = 0; i < cnt; ++i) {
    double t = v[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) sh_part[wid * (3 * KMAX + 2) + i] = t;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    double t = 0.0;
    for (int w = 0; w < nw; ++w) t += sh_part[w * (3 * KMAX + 2) + i];
    sh_out[i] = t;
  }
  __syncthreads();
}

struct VbParams {
  double a[KMAX], b[KMAX], kappa[KMAX], mu[KMAX], nu[KMAX], cov[KMAX];
};

__global__ void __launch_bounds__(VB_THREADS)
k_vbgmm_fit(const float* __restrict__ x_all, int n, const float* __restrict__ init_centers, int K, double dp_prior,
            int max_iter, double tol, double reg_covar, int lloyd_iters, double* __restrict__ out, int* n_iter_out) {
  __shared__ double sh_part[(VB_THREADS / 32) * (3 * KMAX + 2)];
  __shared__ double sh_red[3 * KMAX + 2];
  __shared__ VbParams P;
  __shared__ float s_const[KMAX], s_mu[KMAX], s_h[KMAX];   // E-step constants (fp32)
  __shared__ float s_centers[KMAX];
  __shared__ double s_prior[2];                            // mean prior, covariance prior
  __shared__ int s_stop;
  const int c = blockIdx.x;
  const float* x = x_all + (size_t)c * n;
  const double eps10 = 10.0 * 2.220446049250313e-16;

  // ---- priors: mean and unbiased variance of the column
  {
    double v[2] = {0.0, 0.0};
    for (int i = threadIdx.x; i < n; i += blockDim.x) v[0] += x[i];
    block_reduce_vec(v, 1, sh_part, sh_red);
    const double mean = sh_red[0] / n;
    __syncthreads();
    v[0] = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { const double d = x[i] - mean; v[0] += d * d; }
    block_reduce_vec(v, 1, sh_part, sh_red);
    if (threadIdx.x == 0) { s_prior[0] = mean; s_prior[1] = fmax(sh_red[0] / fmax((double)(n - 1), 1.0), 1e-12); }
    if (threadIdx.x < K) s_centers[threadIdx.x] = init_centers[(size_t)c * K + threadIdx.x];
    __syncthreads();
  }
  // ---- k-means (Lloyd) from the given centres; the last pass yields the hard responsibilities
  double acc[3 * KMAX + 1];
  for (int it = 0; it <= lloyd_iters; ++it) {
#pragma unroll
    for (int k = 0; k < 3 * KMAX + 1; ++k) acc[k] = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float xi = x[i];
      int best = 0;
      float bd = fabsf(xi - s_centers[0]);
#pragma unroll
      for (int k = 1; k < KMAX; ++k)
        if (k < K) {
          const float d = fabsf(xi - s_centers[k]);
          if (d < bd) { bd = d; best = k; }
        }
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {          // static indexing keeps the accumulators in registers
        const double hit = (k == best) ? 1.0 : 0.0;
        acc[k] += hit;
        acc[KMAX + k] += hit * xi;
        acc[2 * KMAX + k] += hit * (double)xi * xi;
      }
    }
    block_reduce_vec(acc, 3 * KMAX, sh_part, sh_red);
    if (it < lloyd_iters) {
      if (threadIdx.x < K && sh_red[threadIdx.x] > 0.0)
        s_centers[threadIdx.x] = (float)(sh_red[KMAX + threadIdx.x] / sh_red[threadIdx.x]);
      __syncthreads();
    }
  }
  double prev_lb = -1e300;
  int iters = 0;
  for (int it = 0; it <= max_iter; ++it) {
    // ---------------- M-step from the sufficient statistics in sh_red (nk, sum r x, sum r x^2, sum r log r)
    if (threadIdx.x < K) {
      const int k = threadIdx.x;
      const double nk = sh_red[k] + eps10;
      const double xk = sh_red[KMAX + k] / nk;
      const double sk = fmax(sh_red[2 * KMAX + k] / nk - xk * xk, 0.0) + reg_covar;
      double tail = 0.0;
      for (int j = k + 1; j < K; ++j) tail += sh_red[j] + eps10;
      P.a[k] = 1.0 + nk;
      P.b[k] = dp_prior + tail;
      P.kappa[k] = 1.0 + nk;
      P.mu[k] = (s_prior[0] + nk * xk) / P.kappa[k];
      P.nu[k] = 1.0 + nk;
      const double diff = xk - s_prior[0];
      P.cov[k] = (s_prior[1] + nk * sk + nk / P.kappa[k] * diff * diff) / P.nu[k];
    }
    __syncthreads();
    // ---------------- lower bound with the E-step's responsibilities and the new parameters
    if (threadIdx.x == 0) {
      s_stop = 0;
      if (it > 0) {
        double log_wishart = 0.0, log_norm_w = 0.0, log_kappa = 0.0;
        for (int k = 0; k < K; ++k) {
          const double ldc = -0.5 * log(P.cov[k]) - 0.5 * log(P.nu[k]);
          log_wishart += -(P.nu[k] * ldc + P.nu[k] * 0.5 * log(2.0) + lgamma(0.5 * P.nu[k]));
          log_norm_w += -(lgamma(P.a[k]) + lgamma(P.b[k]) - lgamma(P.a[k] + P.b[k]));
          log_kappa += log(P.kappa[k]);
        }
        const double lb = -sh_red[3 * KMAX] - log_wishart - log_norm_w - 0.5 * log_kappa;
        if (fabs(lb - prev_lb) < tol) s_stop = 1;
        prev_lb = lb;
        iters = it;
      }
      // E-step constants: log r_k(x) = const_k - h_k (x - mu_k)^2
      double cum = 0.0;
      for (int k = 0; k < K; ++k) {
        const double dsum = digamma_d(P.a[k] + P.b[k]);
        const double e_log_pi = digamma_d(P.a[k]) - dsum + cum;
        cum += digamma_d(P.b[k]) - dsum;
        s_const[k] = (float)(e_log_pi - 0.5 * log(P.cov[k]) - 0.5 * log(P.nu[k]) +
                             0.5 * (digamma_d(0.5 * P.nu[k]) - 1.0 / P.kappa[k]));
        s_mu[k] = (float)P.mu[k];
        s_h[k] = (float)(0.5 / P.cov[k]);
      }
    }
    __syncthreads();
    if (s_stop || it == max_iter) break;
    // ---------------- E-step + sufficient statistics
#pragma unroll
    for (int k = 0; k < 3 * KMAX + 1; ++k) acc[k] = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float xi = x[i];
      float lg[KMAX];
      float mx = -3.0e38f;
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        const float d = xi - s_mu[k];
        lg[k] = (k < K) ? s_const[k] - s_h[k] * d * d : -3.0e38f;
        mx = fmaxf(mx, lg[k]);
      }
      float sum = 0.f;
#pragma unroll
      for (int k = 0; k < KMAX; ++k) sum += (k < K) ? expf(lg[k] - mx) : 0.f;
      const float lse = mx + logf(sum);
      float ent = 0.f;
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (k < K) {
          const float lr = lg[k] - lse;
          const float r = expf(lr);
          acc[k] += r;
          acc[KMAX + k] += r * xi;
          acc[2 * KMAX + k] += (double)r * xi * xi;
          ent += r * lr;
        }
      acc[3 * KMAX] += ent;
    }
    block_reduce_vec(acc, 3 * KMAX + 1, sh_part, sh_red);
  }
  // ---- write [7, K]: means, covs, weights, dof, mean_prec, wc_a, wc_b
  if (threadIdx.x == 0) {
    double* o = out + (size_t)c * 7 * K;
    double w[KMAX], wsum = 0.0, cump = 1.0;
    for (int k = 0; k < K; ++k) {
      const double s = P.a[k] + P.b[k];
      w[k] = P.a[k] / s * cump;
      cump *= P.b[k] / s;
      wsum += w[k];
    }
    for (int k = 0; k < K; ++k) {
      o[0 * K + k] = P.mu[k];
      o[1 * K + k] = P.cov[k];
      o[2 * K + k] = w[k] / wsum;
      o[3 * K + k] = P.nu[k];
      o[4 * K + k] = P.kappa[k];
      o[5 * K + k] = P.a[k];
      o[6 * K + k] = P.b[k];
    }
    if (n_iter_out) n_iter_out[c] = iters;
  }
}

}  // namespace

extern "C" int ftg_vbgmm_fit(const float* x, int C, int n, const float* init_centers, int K, double dp_prior,
                             int max_iter, double tol, double reg_covar, int lloyd_iters, double* out,
                             int* n_iter_out, ftg_stream_t stream) {
  if (K < 1 || K > KMAX || n < 2) return FTG_ERR_UNSUPPORTED;
  k_vbgmm_fit<<<C, VB_THREADS, 0, (cudaStream_t)stream>>>(x, n, init_centers, K, dp_prior, max_iter, tol, reg_covar,
                                                         lloyd_iters, out, n_iter_out);
  FTG_LAUNCH_CHECK();
  return 0;
}
