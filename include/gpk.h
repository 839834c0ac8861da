/* gpk.h -- C ABI of libgpk.so: exact-GP hot path (Gram, Cholesky, LML+gradient, predict) in FP64
 * on sm_100a.  This is the drop-in boundary under the sklearn-style estimator
 * (unmanned_aerial_vehicles_b200.GaussianProcessRegressor).  The reference has no FFI for this path: the
 * interface replaced is the Python estimator used at /root/reference/src/px4/simple_gp.py:167-177,194 and
 * src/px4/gp_trainer.py:163-179,197, whose arithmetic lives in scikit-learn 1.9.0
 * (sklearn/gaussian_process/_gpr.py, kernels.py).  Each entry point cites the lines it replaces.
 *
 * Conventions
 *   - every pointer named *_d is a DEVICE pointer to float64 (int for info/neg counters), row-major;
 *     pointers named *_h are HOST pointers (tiny hyper-parameter vectors, copied by value into launches);
 *   - `stream` is a cudaStream_t passed as void*; all calls are asynchronous on it, no allocation inside;
 *   - factorisation buffers are NP x NP with NP = gpk_pad(N) (multiple of 128), leading dimension NP;
 *   - return value: 0 ok; < 0 bad argument (-k = k-th argument) or -1000-cudaError; the LAPACK-style
 *     "first non-PD leading minor" index is written to *info_d (device int, 0 = ok) so no call blocks.
 */
#ifndef GPK_H
#define GPK_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GPK_ABI_VERSION 1
int gpk_abi_version(void);
/* padded order of an N x N factorisation buffer (next multiple of 128, >= 128) */
int gpk_pad(int N);
/* number of doubles in the LML output vector: [lml, sum log diag L, y.alpha (T), g_c, g_ls (D), g_noise] */
int gpk_lml_out_len(int D, int T);
/* number of doubles of scratch gpk_lml_grad / gpk_lml_eval need */
size_t gpk_grad_workspace_doubles(int N, int D);

/* kernels.py:1530-1587 RBF.__call__(X) + :1374-1419 WhiteKernel + :1244-1296 ConstantKernel, and
 * _gpr.py:589 K[diag] += alpha.  K_d (NP x NP): c*exp(-.5*|xi/l - xj/l|^2), diagonal (c+noise)+jitter,
 * identity in the padding.  lower_only != 0 skips tiles strictly above the diagonal. */
int gpk_gram(const double* X_d, int N, int D, const double* inv_ls_h, double c, double noise, double jitter,
             double* K_d, int NP, int lower_only, void* stream);

/* scipy.linalg.cholesky (dpotrf) at _gpr.py:352,591 fused with the triangular inverse.
 * A_d in: lower triangle of K, out: lower factor L; the 128-blocks strictly above the diagonal of A_d are used as
 * scratch.  Mlow_d = L^-1 (lower), Mup_d = L^-T (upper): only the triangle is written -- the 128-blocks strictly above
 * Mlow_d's block diagonal and strictly below Mup_d's keep whatever the buffers held (inside the diagonal 128-blocks the
 * other triangle is explicit zeros), so a consumer either passes the matching tri hint to gpk_gemm_nt or masks.  One enqueue at a time per device (internal mutex: the
 * recursion's side streams are per device); safe to call from several host threads. */
int gpk_potrf_inv(double* A_d, double* Mlow_d, double* Mup_d, int NP, int* info_d, void* stream);

/* cho_solve((L, True), y) at _gpr.py:363,601 via the explicit inverse: u = L^-1 y, alpha = L^-T u.
 * y_d, u_d, alpha_d are (NP x T) with leading dimension T (rows >= N zero). */
int gpk_potrs(const double* Mlow_d, const double* Mup_d, int NP, const double* y_d, double* u_d, double* alpha_d,
              int T, void* stream);

/* K^-1 = L^-T L^-1 (replaces cho_solve(L, eye(N)) at _gpr.py:631); lower tiles of Kinv_d are written. */
int gpk_lauum(const double* Mup_d, double* Kinv_d, int NP, void* stream);

/* _gpr.py:613-617 (lml) and :629-651 (gradient) without the (N,N,n_theta) tensor.
 * out_d has gpk_lml_out_len(D,T) doubles; gradient entries are d lml / d log(param).
 * Kinv_d may be NULL when with_grad == 0. */
int gpk_lml_grad(const double* X_d, int N, int D, const double* inv_ls_h, double c, double noise,
                 const double* L_d, const double* y_d, const double* alpha_d, int T, const double* Kinv_d, int NP,
                 int with_grad, double* ws_d, double* out_d, void* stream);

/* The gradient contraction of gpk_lml_grad on ONE (rows x cols) block of the matrix (the distributed gradient calls
 * it once per owned tile of K^-1; dist_grad.py).  Row r of the block is training row gi0 + r (inputs Xr_d, nr valid
 * rows, alpha_r_d), column c is training row gj0 + c (Xc_d, nc valid, alpha_c_d); only entries with global column <=
 * global row are counted, strictly lower ones twice.  Kinv_d: the block of K^-1 (rows x cols, leading dimension
 * ldk).  Per-64x64-tile partial sums go to ws_d (gpk_grad_block_workspace_doubles(rows, cols, D) doubles; added to
 * what is there when accumulate != 0).  gpk_grad_block_sum adds the tiles in a fixed order:
 * out_d[0..DP) = sum_ij W_ij K_ij ((x_id - x_jd) / l_d)^2 per feature (DP = gpk_grad_padded_dim(D) >= D),
 * out_d[DP] = sum W_ij K_ij, out_d[DP+1] = tr W; the gradient w.r.t. the log-parameters is 1/2 of these (x noise for
 * the trace), _gpr.py:629-651 + kernels.py:1577-1584.  T <= 8. */
size_t gpk_grad_block_workspace_doubles(int rows, int cols, int D);
int gpk_grad_padded_dim(int D);
int gpk_grad_block(const double* Xr_d, int nr, long long gi0, const double* Xc_d, int nc, long long gj0, int D,
                   const double* inv_ls_h, double c, const double* alpha_r_d, const double* alpha_c_d, int T,
                   const double* Kinv_d, int ldk, int rows, int cols, double* ws_d, int accumulate, void* stream);
int gpk_grad_block_sum(const double* ws_d, int rows, int cols, int D, double* out_d, void* stream);

/* one whole log_marginal_likelihood(theta, eval_gradient) evaluation (_gpr.py:541-656):
 * gram -> potrf_inv -> potrs -> [lauum -> gradient] -> finalize.  A_d, Mlow_d, Mup_d are NP x NP scratch
 * (A_d ends as K^-1 lower tiles when with_grad, else with L's diagonal 128-blocks only: an objective evaluation
 * needs log det alone of L, so the off-diagonal panels stay where their GEMM left them, in Mlow_d, until the inverse
 * overwrites them; gpk_potrf_inv is the call that returns L).  y_d (NP x T, zero padded). */
int gpk_lml_eval(const double* X_d, int N, int D, const double* inv_ls_h, double c, double noise, double jitter,
                 const double* y_d, int T, double* A_d, double* Mlow_d, double* Mup_d, double* u_d, double* alpha_d,
                 int NP, int with_grad, double* ws_d, double* out_d, int* info_d, void* stream);

/* kernels.py:1569-1570 (cdist branch): Ks_d (MP x NP) = c*rbf(Xq, Xtrain), zero in the padding. */
int gpk_cross_gram(const double* Xq_d, int Mq, const double* X_d, int N, int D, const double* inv_ls_h, double c,
                   double* Ks_d, int MP, int NP, void* stream);

/* _gpr.py:446-454: mean_d (Mq x T) = (Ks alpha) * y_std + y_mean */
int gpk_predict_mean(const double* Ks_d, int Mq, int N, int NP, const double* alpha_d, int T, const double* y_mean_d,
                     const double* y_std_d, double* mean_d, void* stream);

/* _gpr.py:446-454 with return_std=False (evaluate_gp_offline.py:198, gp_enhanced_mpc.py:165,175), fused: the
 * cross-Gram row is consumed as it is produced, no Ks in HBM.  Xq_d (Mq x D), MP = Mq rounded up to 64 (or more);
 * partial_d: scratch of gpk_predict_mean_workspace_doubles(MP, N, D, T) doubles (that figure also covers the staged
 * query rows and the result when gpk_predict carves them from one block); mean_d (Mq x T). */
size_t gpk_predict_mean_workspace_doubles(int MP, int N, int D, int T);
int gpk_predict_mean_fused(const double* Xq_d, int Mq, int MP, const double* X_d, int N, int D, const double* inv_ls_h,
                           double c, const double* alpha_d, int T, const double* y_mean_d, const double* y_std_d,
                           double* partial_d, double* mean_d, void* stream);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  gpk_predict_mean_fused with the
 * kernel's scratch (the per-split partial sums: splits(MP, N) * MP * T doubles) taken from a file-scope device array
 * of GPK_STATIC_SCRATCH_DOUBLES doubles (4 MiB, held by every GPU that loads the library until exit) instead of a
 * caller buffer; inputs and outputs are still the caller's.  Allocates nothing: no cudaMalloc / cudaFree / allocating
 * library routine.  A shape whose scratch demand exceeds the array is rejected with GPK_ERR_STATIC_SCRATCH_TOO_SMALL
 * before anything is launched (e.g. MP * T > 524288 at one split; MP = 128 rows against N = 20000, T = 3 needs 56832).
 * The array's device address is looked up once per device, with that device current, on the first call there
 * (cudaGetSymbolAddress, return code checked): make that first call outside graph capture.  Every call on a device
 * shares the array, which keeps the last call's contents -- nothing in it has to start at zero (every slot read is
 * written first by the same call) -- so calls that OVERLAP on the GPU (two streams, two host threads) race on it:
 * serialise them.  Same results as gpk_predict_mean_fused, bit for bit. */
#define GPK_STATIC_SCRATCH_DOUBLES 524288
#define GPK_ERR_STATIC_SCRATCH_TOO_SMALL (-2000)
int gpk_predict_mean_fused_static(const double* Xq_d, int Mq, int MP, const double* X_d, int N, int D,
                                  const double* inv_ls_h, double c, const double* alpha_d, int T,
                                  const double* y_mean_d, const double* y_std_d, double* mean_d, void* stream);

/* _gpr.py:460-500: V = Ks L^-T, var = prior - rowsum(V^2) clamped at 0, std = sqrt(var * y_std^2).
 * rowsq_d: scratch MP x (NP/64).  *neg_d is incremented per clamped row (the caller warns). */
int gpk_predict_std(const double* Ks_d, int Mq, int MP, const double* Mlow_d, int NP, double prior,
                    const double* y_std_d, int T, double* rowsq_d, double* std_d, int* neg_d, void* stream);

/* _gpr.py:370-500 in ONE call from HOST buffers (what simple_gp.py:194 and mpc.py:1490-1509 need per control
 * tick): Xq_h (Mq x D) host -> mean_h (Mq x T) host [, std_h (Mq x T) host, *neg_h = rows whose variance was
 * clamped at 0].  H2D, cross-Gram, mean, L^-1 contraction, std, D2H and one stream sync happen inside.
 * X_d, alpha_d, Mlow_d, y_mean_d, y_std_d are the device-resident model; ws_d is device scratch of at least
 * gpk_predict_workspace_doubles(MP = Mq rounded up to 128, NP, D, T) doubles.  std_h may be NULL: mean only, which
 * takes the fused kernel above and needs only gpk_predict_mean_workspace_doubles(MP, N, D, T) doubles of scratch. */
size_t gpk_predict_workspace_doubles(int MP, int NP, int D, int T);
int gpk_predict(const double* Xq_h, int Mq, int D, const double* X_d, int N, const double* inv_ls_h, double c,
                double noise, const double* alpha_d, int T, const double* Mlow_d, int NP, const double* y_mean_d,
                const double* y_std_d, double* mean_h, double* std_h, int* neg_h, double* ws_d, size_t ws_doubles,
                void* stream);

/* The same for several models that answer the SAME number of query rows each -- the six per-output members (each
 * with its own input scaler, hence its own Xq_h) of the reference's schema B / C pickles: pretrained_gp.py:52-98
 * loops `for name in models: scaler.transform; gp.predict(return_std=True)`, gp_enhanced_mpc.py:150-183 the same.
 * ONE call, ONE stream sync: every model's H2D + launches are enqueued before the first D2H copy.  Fields as the
 * gpk_predict arguments of the same names; each model brings its own scratch (ws_d must differ between models);
 * std_h / neg_h may be NULL per model.  n_models <= GPK_MAX_STACKED. */
#define GPK_MAX_STACKED 16
typedef struct GpkStackedModel {
  const double* Xq_h; int D;
  const double* X_d; int N;
  const double* inv_ls_h; double c; double noise;
  const double* alpha_d; int T;
  const double* Mlow_d; int NP;
  const double* y_mean_d; const double* y_std_d;
  double* mean_h; double* std_h; int* neg_h;
  double* ws_d; size_t ws_doubles;
} GpkStackedModel;
int gpk_predict_stacked(const GpkStackedModel* models_h, int n_models, int Mq, void* stream);

/* evaluate_gp_offline.py:277-312 behind the batched predict: per-component and whole-row error sums of the nominal
 * model (zero residual) and of the GP-corrected one, reduced on the device in a fixed order.
 * Rtrue_d, Rpred_d: (M x C) row-major, C <= 8.  out_d (gpk_residual_metrics_out_len() = 46 doubles):
 *   out[5j + 0..4] = sum t, sum t^2, sum (t-g)^2, #rows t^2 > (t-g)^2, sum (t - mean_j)^2        for component j < C
 *   out[40 + 0..5] = sum se_nom, sum se_gp, #rows se_nom > se_gp, #rows se_nom < se_gp, sum acc_nom, sum acc_gp
 *   (se = squared error summed over the components of a row, acc = components 3..5).
 * ws_d: gpk_residual_metrics_workspace_doubles(M) doubles of scratch. */
int gpk_residual_metrics_out_len(void);
size_t gpk_residual_metrics_workspace_doubles(int M);
int gpk_residual_metrics(const double* Rtrue_d, const double* Rpred_d, int M, int C, double* ws_d, double* out_d,
                         void* stream);

/* utilities */
int gpk_extract_lower(const double* src_d, int NP, double* dst_d, int N, void* stream);
int gpk_pad_copy(const double* src_d, int R, int C, double* dst_d, int RP, int CP, void* stream);
/* C (+)= alpha * A * B^T on the FP64 tensor pipe; M,N,K multiples of 128.  tri: 0 none, 1 lower, 2 upper
 * (k-range trimming; the operand's zero triangle must hold zeros inside its diagonal 128-blocks).
 * lower_only: every element with col <= row is produced, 128-tiles strictly above the diagonal are untouched. */
int gpk_gemm_nt(const double* A_d, int lda, const double* B_d, int ldb, double* C_d, int ldc, int M, int N, int K,
                double alpha, int accumulate, int a_tri, int b_tri, int lower_only, void* stream);
/* register-resident DMMA loop for the FP64 tensor peak; returns flops executed in *flops_h */
int gpk_dmma_probe(int blocks, int iters, double* sink_d, double* flops_h, void* stream);

/* building blocks of the distributed factorisation (dist_chol.py).  There is NO entry point that
 * takes a communicator: the 2-D block-cyclic driver is host-side panel scheduling (dist_chol.py / dist_grad.py) over
 * torch.distributed's NCCL sub-communicators, calling gpk_cross_gram / gpk_potrf_inv / gpk_gemm_nt / gpk_grad_block
 * and the helpers below per tile; a non-Python host would drive the same calls with its own ncclBroadcast. */
/* transpose == 0: out (M x T) = beta*out + alpha * A (M x K) v (K x T);  != 0: out (K x T) += alpha * A^T v (M x T) */
int gpk_gemv(const double* A_d, int lda, int M, int K, const double* v_d, int T, double* out_d, double alpha, double beta,
             int transpose, void* stream);
/* out = tri(M) v with M (NP x NP, ld NP) lower (upper == 0) or upper; v, out (NP x T) */
int gpk_trimv(const double* M_d, int NP, const double* v_d, double* out_d, int T, int upper, void* stream);
/* A[i,i] += add for i < n_valid, A[i,i] = pad_value for n_valid <= i < n */
int gpk_fix_diag(double* A_d, int lda, int n, int n_valid, double add, double pad_value, void* stream);
/* *sum_d += sum_i log A[i,i], i < n */
int gpk_logdiag(const double* A_d, int lda, int n, double* sum_d, void* stream);

/* instrumentation for bench.py (process-global state, NOT thread-safe: one profiling session at a time, and no other
 * thread enqueuing through this library while it is on): kernels launched so far by this library; and per-launch CUDA-event timing of
 * the DMMA GEMM (dominant kernel): begin, run the step, end -> summed launch durations, executed tile flops. */
unsigned long long gpk_launch_count(void);
int gpk_profile_begin(void);
int gpk_profile_end(double* gemm_seconds_h, double* gemm_flops_h, int* n_launches_h);
/* SM-clock cycles thread 0 of the most recent Cholesky leaf spent per phase (8 slots: load, diagonal potrf,
 * panel, trailing update, diagonal inverses, off-diagonal inverse blocks, store, unused); synchronises. */
int gpk_leaf_phase_cycles(long long* out_h);
/* after gpk_profile_end: duration, executed flops and (M,N,K) of the i-th recorded GEMM launch */
int gpk_profile_launch(int i, double* seconds_h, double* flops_h, int* mnk_h);

#ifdef __cplusplus
}
#endif
#endif
