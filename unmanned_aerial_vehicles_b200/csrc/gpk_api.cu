// gpk_api.cu -- extern "C" surface of libgpk.so (see include/gpk.h for the contract and citations).
#include "gpk_common.cuh"
#include "../../include/gpk.h"
#include <string.h>

// launchers defined in the other translation units
int gpk_dp_of(int D);
cudaError_t gpk_launch_gram(const double*, int, int, const GpkKernelParams&, double*, int, int, cudaStream_t);
cudaError_t gpk_launch_cross_gram(const double*, int, const double*, int, int, const GpkKernelParams&, double*, int,
                                  int, cudaStream_t);
cudaError_t gpk_launch_lml_grad(const double*, int, int, const GpkKernelParams&, const double*, size_t, const double*,
                                const double*, int, const double*, int, int, double*, double*, cudaStream_t);
cudaError_t gpk_launch_predict_mean(const double*, int, int, int, const double*, int, const double*, const double*,
                                    double*, cudaStream_t);
cudaError_t gpk_launch_predict_std(const double*, int, int, double, const double*, int, double*, int*, cudaStream_t);
int gpk_mean_fused_splits(int MP, int N);
int gpk_metrics_blocks(int M);
cudaError_t gpk_launch_grad_block(const double*, int, long long, const double*, int, long long, int, const GpkKernelParams&,
                                  const double*, const double*, int, const double*, int, int, int, double*, int,
                                  cudaStream_t);
cudaError_t gpk_launch_grad_sum(const double*, int, int, double*, cudaStream_t);
int gpk_metrics_rec();
cudaError_t gpk_launch_residual_metrics(const double*, const double*, int, int, double*, double*, cudaStream_t);
cudaError_t gpk_launch_predict_mean_fused(const double*, int, int, const double*, int, int, const GpkKernelParams&,
                                          const double*, int, const double*, const double*, double*, double*,
                                          cudaStream_t);
cudaError_t gpk_launch_small_rowsumsq(const double*, int, const double*, int, double*, cudaStream_t);
// Query batches this small take the mat-vec variance path: it re-reads L^-1 once per 8 rows, the DMMA contraction
// pads to 128 rows and runs on NP/128 CTAs only, so the cross-over moves with N (measured at N=1000 and 20000).
static inline bool gpk_small_batch(int Mq, int NP) { return Mq <= 8 || (Mq <= 32 && NP <= 4096); }
cudaError_t gpk_launch_extract_lower(const double*, int, double*, int, cudaStream_t);
cudaError_t gpk_launch_pad_copy(const double*, int, int, double*, int, int, cudaStream_t);
cudaError_t gpk_launch_copy_diag(const double*, int, int, double*, cudaStream_t);
cudaError_t gpk_launch_potrf_inv(double*, int, double*, double*, int, int, int*, cudaStream_t);
cudaError_t gpk_launch_potrf_inv_ex(double*, int, double*, double*, int, int, int*, cudaStream_t, bool keep_L);
cudaError_t gpk_launch_lauum(const double*, int, double*, int, int, cudaStream_t);
cudaError_t gpk_launch_trimv(const double*, int, int, const double*, int, double*, int, int, int, cudaStream_t);
cudaError_t gpk_launch_dmma_probe(int, int, double*, cudaStream_t);
cudaError_t gpk_launch_dmma16816_probe(int, int, double*, cudaStream_t);
cudaError_t gpk_launch_gemv_n(const double*, int, int, int, const double*, int, double*, double, double, cudaStream_t);
cudaError_t gpk_launch_gemv_t(const double*, int, int, int, const double*, int, double*, double, cudaStream_t);
cudaError_t gpk_launch_fix_diag(double*, int, int, int, double, double, cudaStream_t);
cudaError_t gpk_launch_logdiag(const double*, int, int, double*, cudaStream_t);

unsigned long long g_gpk_launches = 0;
GpkProfile g_gpk_prof = {0, 0, 0.0};
double gpk_profile_collect();
int gpk_profile_launch_info(int, double*, double*, int*);
cudaError_t gpk_leaf_cycles(long long* out_h);

static int make_params(GpkKernelParams& p, int D, const double* inv_ls_h, double c, double noise, double jitter) {
  if (D < 1 || D > GPK_MAXD || !inv_ls_h) return -1;
  memset(&p, 0, sizeof(p));
  for (int d = 0; d < D; ++d) p.inv_ls[d] = inv_ls_h[d];
  p.c = c; p.noise = noise; p.jitter = jitter; p.D = D;
  return 0;
}

extern "C" {

int gpk_abi_version(void) { return GPK_ABI_VERSION; }
int gpk_pad(int N) { int n = ((N + GPK_TILE - 1) / GPK_TILE) * GPK_TILE; return n < GPK_TILE ? GPK_TILE : n; }
int gpk_lml_out_len(int D, int T) { return 2 + T + 1 + D + 1; }
size_t gpk_grad_workspace_doubles(int N, int D) {
  size_t nt = (size_t)gpk_pad(N) / 64;
  return nt * nt * (size_t)(gpk_dp_of(D) + 2) + (size_t)gpk_pad(N);
}

int gpk_gram(const double* X_d, int N, int D, const double* inv_ls_h, double c, double noise, double jitter,
             double* K_d, int NP, int lower_only, void* stream) {
  GpkKernelParams p;
  if (!X_d) return -1; if (N < 1) return -2; if (make_params(p, D, inv_ls_h, c, noise, jitter)) return -3;
  if (!K_d) return -8; if (NP != gpk_pad(N)) return -9;
  GPK_CHECK(gpk_launch_gram(X_d, N, D, p, K_d, NP, lower_only, (cudaStream_t)stream));
  return 0;
}

int gpk_potrf_inv(double* A_d, double* Mlow_d, double* Mup_d, int NP, int* info_d, void* stream) {
  if (!A_d) return -1; if (!Mlow_d) return -2; if (!Mup_d) return -3; if (NP < GPK_TILE || NP % GPK_TILE) return -4;
  if (!info_d) return -5;
  GPK_CHECK(gpk_launch_potrf_inv(A_d, NP, Mlow_d, Mup_d, NP, NP, info_d, (cudaStream_t)stream));
  return 0;
}

int gpk_potrs(const double* Mlow_d, const double* Mup_d, int NP, const double* y_d, double* u_d, double* alpha_d,
              int T, void* stream) {
  if (!Mlow_d) return -1; if (!Mup_d) return -2; if (NP % GPK_TILE) return -3; if (T < 1) return -7;
  if (!y_d) return -4; if (!u_d) return -5; if (!alpha_d) return -6;
  GPK_CHECK(gpk_launch_trimv(Mlow_d, NP, NP, y_d, T, u_d, T, 0, T, (cudaStream_t)stream));
  GPK_CHECK(gpk_launch_trimv(Mup_d, NP, NP, u_d, T, alpha_d, T, 1, T, (cudaStream_t)stream));
  return 0;
}

int gpk_lauum(const double* Mup_d, double* Kinv_d, int NP, void* stream) {
  if (!Mup_d) return -1; if (!Kinv_d) return -2; if (NP % GPK_TILE) return -3;
  GPK_CHECK(gpk_launch_lauum(Mup_d, NP, Kinv_d, NP, NP, (cudaStream_t)stream));
  return 0;
}

int gpk_lml_grad(const double* X_d, int N, int D, const double* inv_ls_h, double c, double noise,
                 const double* L_d, const double* y_d, const double* alpha_d, int T, const double* Kinv_d, int NP,
                 int with_grad, double* ws_d, double* out_d, void* stream) {
  GpkKernelParams p;
  if (!X_d) return -1; if (N < 1) return -2;
  if (make_params(p, D, inv_ls_h, c, noise, 0.0)) return -3;
  if (!L_d) return -7; if (!y_d) return -8; if (!alpha_d) return -9;
  if (T < 1) return -10;
  if (with_grad && !Kinv_d) return -11;
  if (!ws_d) return -14; if (!out_d) return -15;
  GPK_CHECK(gpk_launch_lml_grad(X_d, N, D, p, L_d, (size_t)NP + 1, y_d, alpha_d, T, Kinv_d, NP, with_grad, ws_d, out_d,
                                (cudaStream_t)stream));
  return 0;
}

size_t gpk_grad_block_workspace_doubles(int rows, int cols, int D) {
  if (rows < 1 || cols < 1 || rows % 64 || cols % 64) return 0;
  return (size_t)(rows / 64) * (cols / 64) * (size_t)(gpk_dp_of(D) + 2);
}
int gpk_grad_block(const double* Xr_d, int nr, long long gi0, const double* Xc_d, int nc, long long gj0, int D,
                   const double* inv_ls_h, double c, const double* alpha_r_d, const double* alpha_c_d, int T,
                   const double* Kinv_d, int ldk, int rows, int cols, double* ws_d, int accumulate, void* stream) {
  GpkKernelParams p;
  if (!Xr_d) return -1; if (nr < 0 || nr > rows) return -2; if (!Xc_d) return -4; if (nc < 0 || nc > cols) return -5;
  if (make_params(p, D, inv_ls_h, c, 0.0, 0.0)) return -7;
  if (!alpha_r_d) return -10; if (!alpha_c_d) return -11; if (T < 1 || T > 8) return -12;
  if (!Kinv_d) return -13; if (ldk < cols) return -14; if (rows % 64 || cols % 64) return -15; if (!ws_d) return -17;
  GPK_CHECK(gpk_launch_grad_block(Xr_d, nr, gi0, Xc_d, nc, gj0, D, p, alpha_r_d, alpha_c_d, T, Kinv_d, ldk, rows, cols,
                                  ws_d, accumulate, (cudaStream_t)stream));
  return 0;
}
int gpk_grad_block_sum(const double* ws_d, int rows, int cols, int D, double* out_d, void* stream) {
  if (!ws_d) return -1; if (rows % 64 || cols % 64) return -2; if (D < 1 || D > GPK_MAXD) return -4; if (!out_d) return -5;
  GPK_CHECK(gpk_launch_grad_sum(ws_d, (rows / 64) * (cols / 64), D, out_d, (cudaStream_t)stream));
  return 0;
}
int gpk_grad_padded_dim(int D) { return gpk_dp_of(D); }

int gpk_lml_eval(const double* X_d, int N, int D, const double* inv_ls_h, double c, double noise, double jitter,
                 const double* y_d, int T, double* A_d, double* Mlow_d, double* Mup_d, double* u_d, double* alpha_d,
                 int NP, int with_grad, double* ws_d, double* out_d, int* info_d, void* stream) {
  GpkKernelParams p;
  cudaStream_t st = (cudaStream_t)stream;
  if (!X_d) return -1; if (N < 1) return -2;
  if (make_params(p, D, inv_ls_h, c, noise, jitter)) return -3;
  if (!y_d) return -8;
  if (T < 1) return -9;
  if (!A_d) return -10; if (!Mlow_d) return -11; if (!Mup_d) return -12; if (!u_d) return -13; if (!alpha_d) return -14;
  if (NP != gpk_pad(N)) return -15;
  if (!ws_d) return -17; if (!out_d) return -18; if (!info_d) return -19;
  const size_t nt = (size_t)NP / 64;
  double* diag = ws_d + nt * nt * (size_t)(gpk_dp_of(D) + 2);
  GPK_CHECK(gpk_launch_gram(X_d, N, D, p, A_d, NP, 1, st));
  GPK_CHECK(gpk_launch_potrf_inv_ex(A_d, NP, Mlow_d, Mup_d, NP, NP, info_d, st, /*keep_L=*/false));
  GPK_CHECK(gpk_launch_trimv(Mlow_d, NP, NP, y_d, T, u_d, T, 0, T, st));
  GPK_CHECK(gpk_launch_trimv(Mup_d, NP, NP, u_d, T, alpha_d, T, 1, T, st));
  GPK_CHECK(gpk_launch_copy_diag(A_d, NP, NP, diag, st));
  if (with_grad) {
    GPK_CHECK(gpk_launch_lauum(Mup_d, NP, A_d, NP, NP, st));
  }
  p.jitter = 0.0;
  GPK_CHECK(gpk_launch_lml_grad(X_d, N, D, p, diag, 1, y_d, alpha_d, T, A_d, NP, with_grad, ws_d, out_d, st));
  return 0;
}

int gpk_cross_gram(const double* Xq_d, int Mq, const double* X_d, int N, int D, const double* inv_ls_h, double c,
                   double* Ks_d, int MP, int NP, void* stream) {
  GpkKernelParams p;
  if (!Xq_d) return -1; if (Mq < 1) return -2; if (!X_d) return -3; if (N < 1) return -4;
  if (make_params(p, D, inv_ls_h, c, 0.0, 0.0)) return -5;
  if (!Ks_d) return -8;
  if (MP % 64 || MP < Mq) return -9; if (NP < N || NP % 64) return -10;
  GPK_CHECK(gpk_launch_cross_gram(Xq_d, Mq, X_d, N, D, p, Ks_d, MP, NP, (cudaStream_t)stream));
  return 0;
}

int gpk_predict_mean(const double* Ks_d, int Mq, int N, int NP, const double* alpha_d, int T, const double* y_mean_d,
                     const double* y_std_d, double* mean_d, void* stream) {
  if (!Ks_d) return -1; if (Mq < 1) return -2; if (!alpha_d) return -5; if (T < 1) return -6;
  if (!y_mean_d) return -7; if (!y_std_d) return -8; if (!mean_d) return -9;
  GPK_CHECK(gpk_launch_predict_mean(Ks_d, Mq, N, NP, alpha_d, T, y_mean_d, y_std_d, mean_d, (cudaStream_t)stream));
  return 0;
}

int gpk_predict_std(const double* Ks_d, int Mq, int MP, const double* Mlow_d, int NP, double prior,
                    const double* y_std_d, int T, double* rowsq_d, double* std_d, int* neg_d, void* stream) {
  if (!Ks_d) return -1; if (Mq < 1) return -2;
  if (MP % GPK_TILE || MP < Mq) return -3; if (!Mlow_d) return -4; if (NP % GPK_TILE) return -5;
  if (!y_std_d) return -7; if (T < 1) return -8; if (!rowsq_d) return -9; if (!std_d) return -10; if (!neg_d) return -11;
  if (gpk_small_batch(Mq, NP)) {
    GPK_CHECK(gpk_launch_small_rowsumsq(Mlow_d, NP, Ks_d, Mq, rowsq_d, (cudaStream_t)stream));
    GPK_CHECK(gpk_launch_predict_std(rowsq_d, NP / 32, Mq, prior, y_std_d, T, std_d, neg_d, (cudaStream_t)stream));
    return 0;
  }
  GpkGemmArgs g{};
  g.A = Ks_d; g.lda = NP; g.B = Mlow_d; g.ldb = NP; g.C = nullptr; g.ldc = 0; g.M = MP; g.N = NP; g.K = NP;
  g.alpha = 1.0; g.b_tri = GPK_TRI_LOWER; g.epilogue = GPK_EPI_ROWSUMSQ; g.rowsq = rowsq_d;
  GPK_CHECK(gpk_launch_gemm_nt(g, (cudaStream_t)stream));
  GPK_CHECK(gpk_launch_predict_std(rowsq_d, NP / GPK_ROWSQ_TILE, Mq, prior, y_std_d, T, std_d, neg_d, (cudaStream_t)stream));
  return 0;
}

size_t gpk_predict_mean_workspace_doubles(int MP, int N, int D, int T) {
  if (MP < 1 || MP % 64 || N < 1 || D < 1 || T < 1) return 0;
  return (size_t)MP * D + (size_t)MP * T + (size_t)gpk_mean_fused_splits(MP, N) * MP * T + 2;
}
size_t gpk_predict_workspace_doubles(int MP, int NP, int D, int T) {
  if (MP < 1 || NP < 1 || D < 1 || T < 1) return 0;
  const size_t full = (size_t)MP * NP + (size_t)MP * D + 2 * (size_t)MP * T + (size_t)MP * (NP / GPK_ROWSQ_TILE) + 2;
  const size_t mean_only = gpk_predict_mean_workspace_doubles(MP, NP, D, T);   // a scratch sized here serves both
  return full > mean_only ? full : mean_only;
}

int gpk_predict_mean_fused(const double* Xq_d, int Mq, int MP, const double* X_d, int N, int D, const double* inv_ls_h,
                           double c, const double* alpha_d, int T, const double* y_mean_d, const double* y_std_d,
                           double* partial_d, double* mean_d, void* stream) {
  GpkKernelParams p;
  if (!Xq_d) return -1; if (Mq < 1) return -2; if (MP % 64 || MP < Mq) return -3; if (!X_d) return -4; if (N < 1) return -5;
  if (make_params(p, D, inv_ls_h, c, 0.0, 0.0)) return -6;
  if (!alpha_d) return -9; if (T < 1) return -10; if (!y_mean_d) return -11; if (!y_std_d) return -12;
  if (!partial_d) return -13; if (!mean_d) return -14;
  GPK_CHECK(gpk_launch_predict_mean_fused(Xq_d, Mq, MP, X_d, N, D, p, alpha_d, T, y_mean_d, y_std_d, partial_d, mean_d,
                                          (cudaStream_t)stream));
  return 0;
}

// one model's share of a host-to-host predict: everything up to (not including) the device->host copies
static int predict_enqueue(const double* Xq_h, int Mq, int D, const double* X_d, int N, const double* inv_ls_h, double c,
                           double noise, const double* alpha_d, int T, const double* Mlow_d, int NP,
                           const double* y_mean_d, const double* y_std_d, int want_std, double* ws_d, size_t ws_doubles,
                           cudaStream_t st, double** mean_d_out, double** std_d_out, int** neg_d_out) {
  if (!Xq_h || !ws_d) return -1;
  if (Mq < 1) return -2;
  if (NP % GPK_TILE || NP < N) return -12;
  if (T < 1) return -10;
  const int MP = ((Mq + GPK_TILE - 1) / GPK_TILE) * GPK_TILE;
  GpkKernelParams p;
  if (make_params(p, D, inv_ls_h, c, 0.0, 0.0)) return -3;
  if (!want_std) {
    // mean only (evaluate_gp_offline.py:198, gp_enhanced_mpc.py:165,175): the fused kernel, no M x N cross-Gram in HBM
    if (ws_doubles < gpk_predict_mean_workspace_doubles(MP, N, D, T)) return -19;
    double* xq = ws_d;
    double* mean_d = xq + (size_t)MP * D;
    double* partial = mean_d + (size_t)MP * T;
    GPK_CHECK(cudaMemcpyAsync(xq, Xq_h, sizeof(double) * (size_t)Mq * D, cudaMemcpyHostToDevice, st));
    GPK_CHECK(gpk_launch_predict_mean_fused(xq, Mq, MP, X_d, N, D, p, alpha_d, T, y_mean_d, y_std_d, partial, mean_d, st));
    *mean_d_out = mean_d; *std_d_out = nullptr; *neg_d_out = nullptr;
    return 0;
  }
  if (ws_doubles < gpk_predict_workspace_doubles(MP, NP, D, T)) return -19;
  double* Ks = ws_d;
  double* xq = Ks + (size_t)MP * NP;
  double* mean_d = xq + (size_t)MP * D;
  double* std_d = mean_d + (size_t)MP * T;
  double* rowsq = std_d + (size_t)MP * T;
  int* neg_d = reinterpret_cast<int*>(rowsq + (size_t)MP * (NP / GPK_ROWSQ_TILE));
  GPK_CHECK(cudaMemcpyAsync(xq, Xq_h, sizeof(double) * (size_t)Mq * D, cudaMemcpyHostToDevice, st));
  GPK_CHECK(gpk_launch_cross_gram(xq, Mq, X_d, N, D, p, Ks, MP, NP, st));
  GPK_CHECK(gpk_launch_predict_mean(Ks, Mq, N, NP, alpha_d, T, y_mean_d, y_std_d, mean_d, st));
  if (want_std) {
    GPK_CHECK(cudaMemsetAsync(neg_d, 0, sizeof(int), st));
    if (gpk_small_batch(Mq, NP)) {
      GPK_CHECK(gpk_launch_small_rowsumsq(Mlow_d, NP, Ks, Mq, rowsq, st));
      GPK_CHECK(gpk_launch_predict_std(rowsq, NP / 32, Mq, c + noise, y_std_d, T, std_d, neg_d, st));
    } else {
      GpkGemmArgs g{};
      g.A = Ks; g.lda = NP; g.B = Mlow_d; g.ldb = NP; g.C = nullptr; g.ldc = 0; g.M = MP; g.N = NP; g.K = NP;
      g.alpha = 1.0; g.b_tri = GPK_TRI_LOWER; g.epilogue = GPK_EPI_ROWSUMSQ; g.rowsq = rowsq;
      GPK_CHECK(gpk_launch_gemm_nt(g, st));
      GPK_CHECK(gpk_launch_predict_std(rowsq, NP / GPK_ROWSQ_TILE, Mq, c + noise, y_std_d, T, std_d, neg_d, st));
    }
  }
  *mean_d_out = mean_d; *std_d_out = std_d; *neg_d_out = neg_d;
  return 0;
}

static int predict_readback(int Mq, int T, const double* mean_d, const double* std_d, const int* neg_d, double* mean_h,
                            double* std_h, int* neg_h, cudaStream_t st) {
  GPK_CHECK(cudaMemcpyAsync(mean_h, mean_d, sizeof(double) * (size_t)Mq * T, cudaMemcpyDeviceToHost, st));
  if (std_h) {
    GPK_CHECK(cudaMemcpyAsync(std_h, std_d, sizeof(double) * (size_t)Mq * T, cudaMemcpyDeviceToHost, st));
    if (neg_h) GPK_CHECK(cudaMemcpyAsync(neg_h, neg_d, sizeof(int), cudaMemcpyDeviceToHost, st));
  }
  return 0;
}

int gpk_predict(const double* Xq_h, int Mq, int D, const double* X_d, int N, const double* inv_ls_h, double c,
                double noise, const double* alpha_d, int T, const double* Mlow_d, int NP, const double* y_mean_d,
                const double* y_std_d, double* mean_h, double* std_h, int* neg_h, double* ws_d, size_t ws_doubles,
                void* stream) {
  if (!mean_h) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  double *mean_d, *std_d; int* neg_d;
  int rc = predict_enqueue(Xq_h, Mq, D, X_d, N, inv_ls_h, c, noise, alpha_d, T, Mlow_d, NP, y_mean_d, y_std_d,
                           std_h != nullptr, ws_d, ws_doubles, st, &mean_d, &std_d, &neg_d);
  if (rc) return rc;
  rc = predict_readback(Mq, T, mean_d, std_d, neg_d, mean_h, std_h, neg_h, st);
  if (rc) return rc;
  GPK_CHECK(cudaStreamSynchronize(st));
  return 0;
}

// Several models, the same number of query rows each, ONE call and ONE stream sync: the per-output members of the
// reference's schema B / C pickles (pretrained_gp.py:52-98, gp_enhanced_mpc.py:150-183).  Every model's launches are
// enqueued before the first device->host copy, so the host never waits between models.
int gpk_predict_stacked(const GpkStackedModel* models_h, int n_models, int Mq, void* stream) {
  if (!models_h) return -1;
  if (n_models < 1 || n_models > GPK_MAX_STACKED) return -2;
  if (Mq < 1) return -3;
  cudaStream_t st = (cudaStream_t)stream;
  double* mean_d[GPK_MAX_STACKED]; double* std_d[GPK_MAX_STACKED]; int* neg_d[GPK_MAX_STACKED];
  for (int m = 0; m < n_models; ++m) {
    const GpkStackedModel& k = models_h[m];
    if (!k.mean_h) return -1;
    for (int o = 0; o < m; ++o) if (models_h[o].ws_d == k.ws_d) return -19;   // each model needs its own scratch
    const int rc = predict_enqueue(k.Xq_h, Mq, k.D, k.X_d, k.N, k.inv_ls_h, k.c, k.noise, k.alpha_d, k.T, k.Mlow_d, k.NP,
                                   k.y_mean_d, k.y_std_d, k.std_h != nullptr, k.ws_d, k.ws_doubles, st, &mean_d[m],
                                   &std_d[m], &neg_d[m]);
    if (rc) { if (m > 0) cudaStreamSynchronize(st); return rc; }   // earlier models' launches still use their buffers
  }
  for (int m = 0; m < n_models; ++m) {
    const GpkStackedModel& k = models_h[m];
    const int rc = predict_readback(Mq, k.T, mean_d[m], std_d[m], neg_d[m], k.mean_h, k.std_h, k.neg_h, st);
    if (rc) return rc;
  }
  GPK_CHECK(cudaStreamSynchronize(st));
  return 0;
}

int gpk_residual_metrics_out_len(void) { return gpk_metrics_rec(); }
size_t gpk_residual_metrics_workspace_doubles(int M) { return M < 1 ? 0 : (size_t)gpk_metrics_blocks(M) * gpk_metrics_rec(); }
int gpk_residual_metrics(const double* Rtrue_d, const double* Rpred_d, int M, int C, double* ws_d, double* out_d,
                         void* stream) {
  if (!Rtrue_d) return -1; if (!Rpred_d) return -2; if (M < 1) return -3; if (C < 1 || C > 8) return -4;
  if (!ws_d) return -5; if (!out_d) return -6;
  GPK_CHECK(gpk_launch_residual_metrics(Rtrue_d, Rpred_d, M, C, ws_d, out_d, (cudaStream_t)stream));
  return 0;
}

int gpk_extract_lower(const double* src_d, int NP, double* dst_d, int N, void* stream) {
  if (!src_d) return -1; if (!dst_d) return -3; if (N < 1 || N > NP) return -4;
  GPK_CHECK(gpk_launch_extract_lower(src_d, NP, dst_d, N, (cudaStream_t)stream));
  return 0;
}
int gpk_pad_copy(const double* src_d, int R, int C, double* dst_d, int RP, int CP, void* stream) {
  if (!src_d) return -1; if (!dst_d) return -4;
  if (RP < R) return -5; if (CP < C) return -6;
  GPK_CHECK(gpk_launch_pad_copy(src_d, R, C, dst_d, RP, CP, (cudaStream_t)stream));
  return 0;
}
int gpk_gemm_nt(const double* A_d, int lda, const double* B_d, int ldb, double* C_d, int ldc, int M, int N, int K,
                double alpha, int accumulate, int a_tri, int b_tri, int lower_only, void* stream) {
  if (!A_d) return -1; if (!B_d) return -3; if (!C_d) return -5;
  if ((M % GPK_TILE) || (N % GPK_TILE) || (K % GPK_TILE)) return -7;
  if (lda < K) return -2; if (ldb < K) return -4; if (ldc < N) return -6;
  GpkGemmArgs g{};
  g.A = A_d; g.lda = lda; g.B = B_d; g.ldb = ldb; g.C = C_d; g.ldc = ldc; g.M = M; g.N = N; g.K = K; g.alpha = alpha;
  g.accumulate = accumulate; g.a_tri = a_tri; g.b_tri = b_tri; g.lower_only = lower_only; g.epilogue = GPK_EPI_STORE;
  GPK_CHECK(gpk_launch_gemm_nt(g, (cudaStream_t)stream));
  return 0;
}
int gpk_dmma_probe(int blocks, int iters, double* sink_d, double* flops_h, void* stream) {
  if (!sink_d) return -3;
  if (blocks < 0) {  // negative block count selects the m16n8k16 shape: 8 warps x 8 accumulators x 4096 flops
    GPK_CHECK(gpk_launch_dmma16816_probe(-blocks, iters, sink_d, (cudaStream_t)stream));
    if (flops_h) *flops_h = (double)(-blocks) * 8.0 * 8.0 * 4096.0 * (double)iters;
    return 0;
  }
  GPK_CHECK(gpk_launch_dmma_probe(blocks, iters, sink_d, (cudaStream_t)stream));
  // 256 threads = 8 warps, 16 independent m8n8k4 accumulators per warp, 512 flops each
  if (flops_h) *flops_h = (double)blocks * 8.0 * 16.0 * 512.0 * (double)iters;
  return 0;
}

int gpk_gemv(const double* A_d, int lda, int M, int K, const double* v_d, int T, double* out_d, double alpha, double beta,
             int transpose, void* stream) {
  if (!A_d) return -1; if (!v_d) return -5; if (T < 1) return -6; if (!out_d) return -7;
  if (transpose) GPK_CHECK(gpk_launch_gemv_t(A_d, lda, M, K, v_d, T, out_d, alpha, (cudaStream_t)stream));
  else GPK_CHECK(gpk_launch_gemv_n(A_d, lda, M, K, v_d, T, out_d, alpha, beta, (cudaStream_t)stream));
  return 0;
}
int gpk_trimv(const double* M_d, int NP, const double* v_d, double* out_d, int T, int upper, void* stream) {
  if (!M_d) return -1; if (!v_d) return -3; if (!out_d) return -4; if (T < 1) return -5;
  GPK_CHECK(gpk_launch_trimv(M_d, NP, NP, v_d, T, out_d, T, upper, T, (cudaStream_t)stream));
  return 0;
}
int gpk_fix_diag(double* A_d, int lda, int n, int n_valid, double add, double pad_value, void* stream) {
  if (!A_d) return -1; if (n < 0 || lda < n) return -2;
  GPK_CHECK(gpk_launch_fix_diag(A_d, lda, n, n_valid, add, pad_value, (cudaStream_t)stream));
  return 0;
}
int gpk_logdiag(const double* A_d, int lda, int n, double* sum_d, void* stream) {
  if (!A_d) return -1; if (!sum_d) return -4;
  GPK_CHECK(gpk_launch_logdiag(A_d, lda, n, sum_d, (cudaStream_t)stream));
  return 0;
}
unsigned long long gpk_launch_count(void) { return g_gpk_launches; }
int gpk_leaf_phase_cycles(long long* out_h) {
  if (!out_h) return -1;
  if (cudaDeviceSynchronize() != cudaSuccess) return -1000;
  GPK_CHECK(gpk_leaf_cycles(out_h));
  return 0;
}
int gpk_profile_launch(int i, double* seconds_h, double* flops_h, int* mnk_h) { return gpk_profile_launch_info(i, seconds_h, flops_h, mnk_h); }
int gpk_profile_begin(void) { g_gpk_prof.enabled = 1; g_gpk_prof.n = 0; g_gpk_prof.flops = 0.0; return 0; }
int gpk_profile_end(double* gemm_seconds_h, double* gemm_flops_h, int* n_launches_h) {
  if (cudaDeviceSynchronize() != cudaSuccess) return -1000;
  if (gemm_seconds_h) *gemm_seconds_h = gpk_profile_collect();
  if (gemm_flops_h) *gemm_flops_h = g_gpk_prof.flops;
  if (n_launches_h) *n_launches_h = g_gpk_prof.n;
  g_gpk_prof.enabled = 0;
  return 0;
}

}  // extern "C"
