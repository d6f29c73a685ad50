/* fedtgan_b200 -- C ABI of the sm_100a kernel library (libfedtgan_b200.so).
 *
 * Every entry point launches on the caller's stream, allocates nothing unless stated,
 * and returns a cudaError_t value as int (0 == success).  Matrices are row-major fp32
 * with an explicit leading dimension (floats).  Where a reference op exists, the comment
 * cites the reference file it replaces (paths relative to the Fed-TGAN tree,
 * S/ = Server/).  The PyTorch oracle of each function is the function of the same name
 * in fed_tgan_b200/ops/reference.py.
 */
#ifndef FEDTGAN_B200_H
#define FEDTGAN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* ftg_stream_t; /* cudaStream_t */

/* ---- GEMM with fused epilogue ------------------------------------------------------
 * D[M,N] (+)= mask_in * act(alpha * A[M,K] . B[N,K]^T + bias[n])
 *   a_t: A stored [K,M] (M contiguous)      b_t: B stored [K,N] (N contiguous)
 *   act: 0 none | 1 leaky(slope) | 2 leaky(slope)+dropout(p); act!=0 writes the applied
 *        multiplier to mask_out (if non-null).  Dropout keep = (u >= p) with u from
 *        drop_u (tests) or Philox(rng_state, rng_stream, element index).
 * Replaces: torch Linear / LeakyReLU / Dropout and their autograd backward
 * (S/dtds/synthesizers/ctgan.py:15-64), SURVEY K2/K3/K7.                               */
typedef struct {
  const float* a; const float* b; float* d;
  int M, N, K;
  int lda, ldb, ldd;
  int a_t, b_t;
  const float* bias;
  int act;
  const float* mask_in; int ld_mask_in;
  float* mask_out; int ld_mask_out;
  const float* drop_u; int ld_drop_u;
  const int64_t* rng_state; int rng_stream;
  float alpha, slope, drop_p;
  int accumulate;
  int x3;   /* ftg_gemm_tc only: 0 = single-pass tf32; 1 = fp32x3 (3xTF32 hi/lo split, fp32-equivalent products);
             * 2 = fp32x3 with hi = the tensor core's own truncation of the fp32 word (only lo is written) */
} ftg_gemm_args;

/* fp32 CUDA-core GEMM (exact fp32; fallback + oracle for the tensor-core path) */
int ftg_gemm_simt(const ftg_gemm_args* args, ftg_stream_t stream);

/* tcgen05 / TMEM / TMA GEMM, kind::tf32 with fp32 accumulation in tensor memory.
 * args->x3 = 1 selects the fp32-equivalent mode: every operand stage is split in shared memory into
 * hi = rna_tf32(x) and lo = rna_tf32(x - hi) and A_lo.B_hi + A_hi.B_lo + A_hi.B_hi is accumulated into
 * the same TMEM tile (3xTF32), which reproduces the reference's fp32 matmuls to fp32 round-off.
 * Split-K partial tiles go through `workspace` (>= ftg_gemm_tc_workspace_bytes(...)) and
 * `tile_counters` (int32, zero-initialised once, >= ftg_gemm_tc_num_tiles(...)); both are
 * caller-owned scratch and may be NULL when the call does not split K.
 * Requirements: base pointers 16 B aligned, lda/ldb multiples of 4 floats.
 * Returns 1000 (FTG_ERR_UNSUPPORTED) when the shape/alignment cannot be handled.        */
int ftg_gemm_tc(const ftg_gemm_args* args, int split_k, float* workspace, int* tile_counters,
                ftg_stream_t stream);
/* Residual block forward in ONE launch: D = A.B^T + bias (pre-activation, kept for the backward pass) and
 * out = relu(batchnorm_train(D)) stored into a column window of a wider buffer (the torch.cat of
 * S/dtds/synthesizers/ctgan.py:40-44).  Batch statistics are reduced over all rows inside the kernel
 * (thread-block cluster over the M-tiles, distributed shared memory, Chan's merge); running_mean /
 * running_var (momentum, unbiased variance), save_mean / save_invstd and num_batches_tracked are
 * updated like torch.nn.BatchNorm1d.  Needs M <= 512; otherwise returns 1000 and the caller uses
 * ftg_gemm_tc + ftg_bn_relu_fwd.  SURVEY K2. */
typedef struct {
  const float* gamma; const float* beta;
  float* running_mean; float* running_var;
  float* save_mean; float* save_invstd;      /* may be NULL */
  float* out; int ldo;                       /* relu(bn(.)) destination, row pitch in floats */
  float eps, momentum;
  float* nbt;                                /* num_batches_tracked (1 float), may be NULL */
} ftg_bn_args;
int ftg_gemm_tc_bn(const ftg_gemm_args* args, const ftg_bn_args* bn, ftg_stream_t stream);
int ftg_gemm_tc_bn_plan(int M, int K);   /* split-K factor the fused kernel will use; 0 = shape not supported */
/* (No counterpart in the reference.)  ftg_gemm_tc for callers that cannot allocate device
 * memory -- e.g. a plugin inside a host that owns the whole device heap, an out-of-memory
 * fail-safe path, a loop that cannot take cudaMalloc's device synchronisation.  The split-K
 * workspace and tile counters come from a fixed device array built into the library
 * (8 MiB + 4 KiB per GPU, held from library load to process exit); operands and output are
 * still the caller's.  Returns 1001 (FTG_ERR_SCRATCH_TOO_SMALL, text in
 * ftg_last_error_message()) when split_k * tiles * 32 KiB exceeds
 * ftg_gemm_tc_noalloc_max_workspace_bytes() or the output has more than 1024 tiles.  The
 * first call on a device looks the array up (make it before any stream capture on that
 * device).  All calls on a device share the array: do not let two of them overlap on the
 * GPU.  Calls no cudaMalloc / cudaFree / allocating routine.                              */
int ftg_gemm_tc_noalloc(const ftg_gemm_args* args, int split_k, ftg_stream_t stream);
long long ftg_gemm_tc_noalloc_max_workspace_bytes(void);
const char* ftg_last_error_message(void);
int ftg_gemm_tc_choose_split(int M, int N, int K);
/* With workspace == NULL and split_k in {2,4,8} the split_k CTAs of a tile form a thread-block
 * cluster and reduce their partial tiles through distributed shared memory (no scratch). */
/* Force the TMA ring depth of subsequent tcgen05 GEMM launches (2..6; 0 = automatic policy). */
int ftg_gemm_tc_set_stages(int stages);
/* Experimental: route multi-wave, unsplit tcgen05 GEMMs through the persistent kernel (one CTA per SM,
 * double-buffered TMEM accumulator).  Off by default. */
int ftg_gemm_tc_set_persistent(int on);
int ftg_gemm_tc_choose_cluster_split(int M, int N, int K);
/* In-cluster split-K reduction protocol: 0 (default) = slice owners pull through remote shared-memory loads
 * between two cluster barriers; 1 = every CTA pushes the other ranks' slices of its partial tile into the owners'
 * shared memory with st.async + mbarrier complete_tx.  Both sum in rank order (bitwise identical results). */
int ftg_gemm_tc_set_cluster_push(int on);
long long ftg_gemm_tc_workspace_bytes(int M, int N, int split_k);
int ftg_gemm_tc_num_tiles(int M, int N);

/* ---- small dense helpers ------------------------------------------------ */
int ftg_rowdot(const float* a, int lda, int M, int K, const float* w, const float* bias, float* out,
               ftg_stream_t stream);
/* rowdot + fused scalar  loss_out = sum_m rowscale[m] * out[m]  (WGAN loss value) */
int ftg_rowdot_loss(const float* a, int lda, int M, int K, const float* w, const float* bias, float* out,
                    const float* rowscale, float* loss_out, ftg_stream_t stream);
/* discriminator head: rowdot + loss value + first backward signal Z = rs_sig (x) w (.) mask, one launch */
int ftg_d_head(const float* a, int lda, int M, int K, const float* w, const float* bias, float* y,
               const float* rs_loss, float* loss_out, const float* rs_sig, const float* mask, int ldm, float* Z,
               int ldz, ftg_stream_t stream);
int ftg_colsum(const float* a, int lda, int M, int N, const float* rowscale, float* out, int accumulate,
               ftg_stream_t stream);
/* Fused middle of the two-hidden-layer PacGAN discriminator for the packed batch, one launch, exact fp32:
 *   a1 = dropout(leaky(a0 . W1^T + b1)) (+ mask1);  y = a1 . w + b;  loss = sum rs_loss*y;
 *   Z1 = rs_sig (x) w (.) mask1;  Z0 = (Z1 . W1) (.) mask0
 * (= Linear+LeakyReLU+Dropout, the head Linear, and their backward signals: S/dtds/synthesizers/ctgan.py:22-30).
 * Row local (no BatchNorm in D): a CTA of 16 warps owns 2 rows and streams W1 once through a cp.async ring.  Needs H0, H1 <= 512, H0 % 4 == 0 (ftg_d_mid_supported), `counter`
 * = one zero-initialised device int when loss_out != NULL.  Dropout uniforms as in ftg_gemm_tc's epilogue. */
int ftg_d_mid_supported(int H0, int H1);
int ftg_d_mid(const float* a0, int lda0, int M, int H0, int H1, const float* w1, int ldw1, const float* b1, float* a1,
              int lda1, float* mask1, int ldm1, const float* mask0, int ldm0, const float* w_head, const float* b_head,
              float* y, const float* rs_loss, float* loss_out, const float* rs_sig, float* z1, int ldz1, float* z0,
              int ldz0, const float* drop_u, int ldu, const int64_t* rng_state, int rng_stream, float slope,
              float drop_p, int* counter, ftg_stream_t stream);
int ftg_scale_outer_mask(const float* rowscale, const float* v, const float* mask, int ldm, float* out,
                         int ldo, int M, int N, ftg_stream_t stream);

/* ---- BatchNorm1d(train/eval)+ReLU with concat-store, and its backward -------------------
 * Replaces Residual's BatchNorm1d+ReLU+cat (S/dtds/synthesizers/ctgan.py:40-44).         */
int ftg_bn_relu_fwd(const float* a, int lda, int B, int H, const float* gamma, const float* beta,
                    float* running_mean, float* running_var, float* out, int ldo, float* save_mean,
                    float* save_invstd, int training, float eps, float momentum, float* num_batches_tracked,
                    ftg_stream_t stream);
int ftg_bn_relu_bwd(const float* dout, int lddo, const float* a, int lda, int B, int H, const float* save_mean,
                    const float* save_invstd, const float* gamma, const float* beta, float* da, int ldda,
                    float* dgamma, float* dbeta, ftg_stream_t stream);
/* same, additionally writes d(fc bias)[h] = sum_m da[m,h] (bias of the Linear feeding the BN) */
int ftg_bn_relu_bwd_bias(const float* dout, int lddo, const float* a, int lda, int B, int H,
                         const float* save_mean, const float* save_invstd, const float* gamma, const float* beta,
                         float* da, int ldda, float* dgamma, float* dbeta, float* dbias_fc, ftg_stream_t stream);

/* ---- apply_activate (tanh / gumbel-softmax) fwd+bwd, cond_loss fwd+bwd --------------------
 * Replaces apply_activate (ctgan.py:67-82) and cond_loss (ctgan.py:174-194), SURVEY K4/K5. */
int ftg_activate_fwd(const float* fake, int ldf, int B, int dd, int n_spans, const int* span_start,
                     const int* span_width, const int* span_kind, const int* col_span, float* out, int ldo, float tau,
                     const float* u, int ldu, const int64_t* rng_state, int rng_stream, const float* cond,
                     int ldc, int n_opt, int use_noise, ftg_stream_t stream);
/* activate + cat(c1) + slerp(alpha, real, fake_cat) in one launch (D step) */
int ftg_activate_fwd_slerp(const float* fake, int ldf, int B, int dd, int n_spans, const int* span_start,
                           const int* span_width, const int* span_kind, const int* col_span, float* out, int ldo,
                           float tau, const float* u, int ldu, const int64_t* rng_state, int rng_stream,
                           const float* cond, int ldc, int n_opt, int use_noise, const float* real, int ldr,
                           float* interp, int ldi, const float* alpha, int alpha_stream, ftg_stream_t stream);
/* backward of apply_activate fused with cond_loss value + gradient (G step) */
int ftg_gen_output_bwd(const float* dout, int lddo, const float* act, int lda, int B, int n_spans,
                       const int* span_start, const int* span_width, const int* span_kind, float* dfake, int ldf,
                       float tau, const float* fake, int ldfk, const float* cond, int ldc, const float* mask, int ldm,
                       const int* col_soft, const int* soft_cstart, float* loss_out, ftg_stream_t stream);
int ftg_activate_bwd(const float* dout, int lddo, const float* act, int lda, int B, int n_spans,
                     const int* span_start, const int* span_width, const int* span_kind, float* dfake, int ldf,
                     float tau, int accumulate, ftg_stream_t stream);
int ftg_cond_loss(const float* fake, int ldf, const float* cond, int ldc, const float* mask, int ldm, int B,
                  int dd, int n_col, const int* soft_start, const int* soft_width, const int* soft_cstart,
                  float* dfake, int lddf, float* loss_out, int accumulate, ftg_stream_t stream);

/* ---- slerp interpolation + per-pack gradient-penalty norm (ctgan.py:231-258), K6 --------- */
int ftg_slerp(const float* low, int ldl, const float* high, int ldh, float* out, int ldo, int B, int din,
              const float* alpha, const int64_t* rng_state, int rng_stream, ftg_stream_t stream);
int ftg_gp_penalty(float* g, int ldg, int Bp, int Din, float* pen_out, float lambda_, ftg_stream_t stream);

/* ---- fused Adam over a flat arena (torch.optim.Adam semantics, L2-in-gradient), K8 ------- */
int ftg_adam(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n,
             const int64_t* rng_state /* step = state[1] */, float lr, float beta1, float beta2, float eps,
             float weight_decay, ftg_stream_t stream);

/* ---- device-side sampling (Cond.sample / Sampler.sample, ctgan.py:147-228), K9-K11 ------- */
int ftg_tick(int64_t* rng_state, ftg_stream_t stream);
int ftg_fill_normal(float* out, int rows, int cols, int ld, const int64_t* rng_state, int rng_stream,
                    ftg_stream_t stream);
int ftg_sample_condvec(const float* cum_p, int n_col, int max_w, const int* soft_width, int* col_out,
                       int* opt_out, int B, const float* u_col, const float* u_opt, const int64_t* rng_state,
                       int rng_stream, ftg_stream_t stream);
int ftg_write_condvec(const int* col, const int* opt, const int* soft_cstart, float* c_out, int ldc, int n_opt,
                      float* m_out, int ldm, int n_col, const int* perm, int B, ftg_stream_t stream);
int ftg_argsort_perm(int* perm, int B, const float* u, const int64_t* rng_state, int rng_stream,
                     ftg_stream_t stream);
int ftg_sample_rows(const int* row_ptr, const int* row_ids, const int* soft_cstart, const int* col,
                    const int* opt, const int* perm, int* rid_out, int B, const float* u,
                    const int64_t* rng_state, int rng_stream, ftg_stream_t stream);
int ftg_expand_rows(const float* vals, int n_tanh, const int* codes, int n_col, const int* rid,
                    const int* tanh_start, const int* soft_start, float* out, int ld, int dd, int B,
                    ftg_stream_t stream);

/* fused batch preparation: (col,opt) draw + batch permutation + matching-row draw in one CTA,
 * then one CTA per row writes z, c1, m1 (mode 0, G step) or z, c1 and the expanded real row
 * with its permuted cond vector (mode 1, D step). */
int ftg_prep_select(const float* cum_p, int n_col, int max_w, const int* soft_width, const int* soft_cstart,
                    const int* row_ptr, const int* row_ids, int* col, int* opt, int* perm, int* rid, int B,
                    int64_t* rng_state, int s_col, int s_perm, int s_row, int do_tick, ftg_stream_t stream);
int ftg_prep_rows(int mode, int* col, int* opt, const int* perm, const int* rid, const int* soft_cstart,
                  float* z_out, int ldz, int E, float* c_out, int ldc, int n_opt, float* m_out, int ldm, int n_col,
                  const float* vals, int n_tanh, const int* codes, const int* tanh_start, const int* soft_start,
                  float* real_out, int ldr, int dd, const float* cum_p, int max_w, const int* soft_width, int B,
                  const int64_t* rng_state, int s_z, int s_col, ftg_stream_t stream);

/* ---- VGM codec (BGM_CTGAN_Transformer.transform / inverse_transform,
 *      S/dtds/features/transformers.py:385-464), K12/K13 ---------------------------------- */
int ftg_vgm_decode(const float* act, int lda, int rows, int n_columns, const int* col_is_cont,
                   const int* col_start, const int* col_width, const int* col_aux, const float* mu,
                   const float* sd, int n_clusters, const float* cat_ids, float* out, int ldo,
                   ftg_stream_t stream);
int ftg_vgm_encode(const float* x, int rows, const float* logit_const, const float* mu, const float* var,
                   int k_active, const float* u, const int64_t* rng_state, int rng_stream, float* val_out,
                   int val_stride, int* code_out, int code_stride, ftg_stream_t stream);

/* Variational-Bayes DP-GMM fit (sklearn BayesianGaussianMixture semantics, n_features = 1) of C
 * columns at once, one CTA per column: x [C][n] fp32, init_centers [C][K] (k-means seeds, e.g.
 * quantiles), out [C][7][K] fp64 = means, covs, weights, dof, mean_prec, wc_a, wc_b.  K <= 12.
 * Replaces BayesianGaussianMixture.fit (S/dtds/features/transformers.py:334-340,
 * S/dtds/distributed.py:725-746), SURVEY K14.                                              */
int ftg_vbgmm_fit(const float* x, int C, int n, const float* init_centers, int K, double dp_prior, int max_iter,
                  double tol, double reg_covar, int lloyd_iters, double* out, int* n_iter_out, ftg_stream_t stream);

/* ---- federated weighted average (average_model, S/dtds/distributed.py:86-106 + the RPC
 *      gather/broadcast R9/R10), SURVEY K1 ------------------------------------------------
 * theta <- sum_i w_i theta_i written into every replica; `int_offsets` slots are truncated
 * (num_batches_tracked).
 *  ftg_fedavg_local : all replicas on this GPU, one launch over all ranks' data
 *                     (single-GPU emulation / LocalGroup).
 *  ftg_fedavg_twoshot / _oneshot : one launch per rank on its own GPU; `peers[r]` are
 *     peer-mapped device pointers to every rank's arena, `flags[r]` to every rank's flag
 *     page (>= 4*world ints, zero-initialised).  Ranks synchronise inside the kernel with
 *     system-scope flags; `round_id` must increase by one per call.  `status` (device
 *     int) is set non-zero if a peer did not arrive within `timeout_ms`.                   */
/* `arenas`/`peers`/`flags`/`int_offsets` are small HOST arrays (passed to the kernel by value); the pointers
 * stored in them are device / peer-mapped pointers.  `weights` is either a HOST array (copied by value) or a
 * DEVICE pointer to `world` floats that the kernel reads when it runs -- e.g. `weights_out` of
 * ftg_fed_stats_to_weights, so the similarity weights never leave the GPU. */
int ftg_fedavg_local(float* const* arenas, int world, const float* weights, long long n,
                     const int* int_offsets, int n_int, ftg_stream_t stream);
int ftg_fedavg_twoshot(float* const* peers, int* const* flags, int rank, int world, const float* weights,
                       long long n, const int* int_offsets, int n_int, int round_id, int* status,
                       int timeout_ms, ftg_stream_t stream);
int ftg_fedavg_oneshot(float* const* peers, int* const* flags, float* staging, int rank, int world,
                       const float* weights, long long n, const int* int_offsets, int n_int, int round_id,
                       int* status, int timeout_ms, ftg_stream_t stream);

/* NVLS variant: replicas live in a multicast-bound symmetric allocation; each rank pre-scales its
 * replica by w_rank, then shard r is summed inside the NVSwitch (multimem.ld_reduce.add.f32) and
 * multicast back to all replicas (multimem.st).  Not bitwise reproducible (switch sum order). */
int ftg_fedavg_nvls(float* const* peers, int* const* flags, float* multicast_base, int rank, int world,
                    const float* weights, long long n, const int* int_offsets, int n_int, int round_id,
                    int* status, int timeout_ms, ftg_stream_t stream);

/* peer-memory plumbing (cudaMalloc + CUDA IPC); the only allocating entry points */
int ftg_symm_alloc(long long bytes, void** ptr);
int ftg_symm_free(void* ptr);
int ftg_symm_export(void* ptr, unsigned char handle_out[64]);
int ftg_symm_open(const unsigned char handle[64], void** ptr);
int ftg_symm_close(void* ptr);
int ftg_enable_peer_access(int device, int peer);

/* ---- similarity weights (uniform_meta_category / uniform_continuous_gmm /
 *      calculate_final_weights_for_aggregation, S/dtds/distributed.py:592-783), K15 --------
 * hist: [world][hist_len] category counts (columns concatenated, `cat_off` [n_cat+1]);
 * pool_sorted: [n_cont][n_total] pooled GMM sample of every continuous column, sorted;
 * client_sorted + smp_off [n_cont*world+1]: each client's sorted sample per column;
 * rows [world].  Writes normalised JSD [world][n_cat], WD [world][n_cont] and the final
 * soft-maxed weights [world] (device memory, read directly by the fedavg kernels).       */
int ftg_stats_to_weights(const float* hist, int world, int n_cat, const int* cat_off, const float* pool_sorted,
                         int n_total, const float* client_sorted, const int* smp_off, int n_cont,
                         const float* rows, float* jsd_out, float* wd_out, float* weights_out,
                         ftg_stream_t stream);

/* ---- federated initialisation statistics over peer memory (SURVEY K14 sampler + K15) ----------
 * ftg_gmm_sample_sort: for each of a client's C continuous columns draw n samples from its mixture
 *   (gmm: [C][7][K] fp64 as written by ftg_vbgmm_fit; rows 0..2 = means, variances, weights) with Philox
 *   and sort them inside the kernel (bitonic, one CTA per column).  out: [C][ftg_sort_pad(n)] floats, the
 *   first n of each row ascending, the padding +inf.  Replaces client_gmm.sample(...)
 *   (S/dtds/distributed.py:732-734) and the sort inside scipy's wasserstein_distance.
 * ftg_fed_stats_to_weights: `hist[i]` / `seg[i]` are device pointers -- LOCAL OR PEER-MAPPED -- to client i's
 *   category histogram (global label order, columns concatenated, cat_off [n_cat+1]) and to its sorted samples
 *   from ftg_gmm_sample_sort ([n_cont][ftg_sort_pad(n_i[i])]); `n_i`, `rows` and the pointer arrays are small
 *   HOST arrays (passed by value).  The kernels read every client's statistics in place (over NVLink for peers)
 *   and write the column-normalised JSD [world][n_cat], WD [world][n_cont] and the soft-maxed weights [world]
 *   to device memory; ftg_fedavg_* accept that `weights_out` pointer directly.  Optional outputs: `centers`
 *   [n_cont][K] pooled quantiles ((k+0.5)/K, seeds of the global mixture refit), `pool_out` [n_cont][sum n_i]
 *   the pooled sample.  `partial`: scratch of n_cont*chunks*world floats.
 *   Replaces S/dtds/distributed.py:592-686 (JSD part), 689-765 (WD part), 767-783.              */
int ftg_sort_pad(int n);
/* rows of x [C][n] sorted ascending into out [C][ftg_sort_pad(n)] (padding +inf), same in-kernel bitonic network */
int ftg_sort_rows(const float* x, int C, int n, float* out, ftg_stream_t stream);
int ftg_gmm_sample_sort(const double* gmm, int C, int K, int n, unsigned long long seed, int client_id, float* out,
                        ftg_stream_t stream);
int ftg_fed_stats_to_weights(const float* const* hist, const float* const* seg, const int* n_i, const float* rows,
                             int world, int n_cat, const int* cat_off, int n_cont, float* partial, int chunks,
                             float* centers, int K, float* pool_out, float* jsd_out, float* wd_out,
                             float* weights_out, ftg_stream_t stream);

/* ---- host-side CSV writer (S/dtds/distributed.py:586-590 DataFrame.to_csv) --------------- */
int ftg_csv_write(const char* path, int n_rows, int n_cols, const char* const* names, const int* kinds,
                  const void* const* columns, const char* const* const* str_tables, int n_threads);
/* same, reading float32 columns (kind 1) and float32-coded categories (kind 4, clamped to n_classes[c]) in place from
 * a row-major table through per-column strides: the generator's numeric table goes to text without host-side passes.
 * n_threads < 0: |n_threads| threads and no header line (one part of a table the caller concatenates). */
int ftg_csv_write_strided(const char* path, int n_rows, int n_cols, const char* const* names, const int* kinds,
                          const void* const* columns, const int* strides, const int* n_classes,
                          const char* const* const* str_tables, int n_threads);

const char* ftg_version(void);
long long ftg_last_launch_count(void);   /* kernels launched by this library since load */

#ifdef __cplusplus
}
#endif
#endif
