// gpk_kernels.cu -- kernel-function kernels: Gram assembly, cross-Gram, LML-gradient contraction,
// prediction finalisation.  Mirrors sklearn/gaussian_process/kernels.py:
//   RBF.__call__ (:1530-1587), WhiteKernel.__call__/diag (:1374-1440), ConstantKernel (:1244-1296),
//   Sum/Product (:838-990) for the algebra  c * RBF(iso|ARD) + WhiteKernel(noise).
// Squared distances use the direct sum of squared differences of X/length_scale (what pdist/cdist
// 'sqeuclidean' computes), not the |a|^2+|b|^2-2ab expansion, to keep 1e-13 parity near duplicates.
#include "gpk_common.cuh"

#define GRAD_TC 8  // targets per gradient pass
#define KTILE 64   // 64x64 output tile, 256 threads, 4x4 micro-tile per thread

template <int DP>
__device__ __forceinline__ void load_scaled_rows(double (*s)[DP + 1], const double* __restrict__ X, int ldx, int r0,
                                                 int nrows_valid, const GpkKernelParams& p, int tid) {
  for (int idx = tid; idx < KTILE * DP; idx += 256) {
    int r = idx / DP, d = idx % DP;
    double v = 0.0;
    if (r0 + r < nrows_valid && d < p.D) v = X[(size_t)(r0 + r) * ldx + d] * p.inv_ls[d];
    s[r][d] = v;
  }
}

// K[i,j] for the padded NP x NP buffer.  lower_only: only tiles with tj <= ti are produced.
template <int DP>
__global__ void __launch_bounds__(256, 2) gpk_gram_kernel(const double* __restrict__ X, int ldx, int N, int NP,
                                                       GpkKernelParams p, double* __restrict__ K, int ldk,
                                                       int lower_only) {
  const int tj = blockIdx.x, ti = blockIdx.y;
  if (lower_only && (tj * KTILE) > (ti * KTILE + KTILE - 1)) return;
  __shared__ double xi[KTILE][DP + 1], xj[KTILE][DP + 1];
  const int tid = threadIdx.x, i0 = ti * KTILE, j0 = tj * KTILE;
  load_scaled_rows<DP>(xi, X, ldx, i0, N, p, tid);
  load_scaled_rows<DP>(xj, X, ldx, j0, N, p, tid);
  __syncthreads();
  const int ty = tid >> 4, tx = tid & 15;
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int li = ty * 4 + a, i = i0 + li;
    double out[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int lj = tx + 16 * b, j = j0 + lj;    // cyclic columns: lanes read xj with an odd row stride (no bank
      double d2 = 0.0;                            // conflicts) and store 16 consecutive doubles per row
#pragma unroll
      for (int d = 0; d < DP; ++d) { double df = xi[li][d] - xj[lj][d]; d2 += df * df; }
      double v;
      if (i >= N || j >= N) v = (i == j) ? 1.0 : 0.0;
      else if (i == j) v = (p.c + p.noise) + p.jitter;
      else v = p.c * exp(-0.5 * d2);
      out[b] = v;
    }
    double* dst = K + (size_t)i * ldk + j0 + tx;
#pragma unroll
    for (int b = 0; b < 4; ++b) dst[16 * b] = out[b];
  }
}

// Kstar[q, j] = c * rbf(xq, xtrain_j)  (q < Mq, j < N), zero in the padding; MP x NP output.
template <int DP>
__global__ void __launch_bounds__(256, 2) gpk_cross_gram_kernel(const double* __restrict__ Xq, int ldq, int Mq,
                                                             const double* __restrict__ X, int ldx, int N,
                                                             GpkKernelParams p, double* __restrict__ Ks, int ldk) {
  __shared__ double xi[KTILE][DP + 1], xj[KTILE][DP + 1];
  const int tid = threadIdx.x, i0 = blockIdx.y * KTILE, j0 = blockIdx.x * KTILE;
  load_scaled_rows<DP>(xi, Xq, ldq, i0, Mq, p, tid);
  load_scaled_rows<DP>(xj, X, ldx, j0, N, p, tid);
  __syncthreads();
  const int ty = tid >> 4, tx = tid & 15;
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int li = ty * 4 + a, i = i0 + li;
    double out[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int lj = tx + 16 * b, j = j0 + lj;    // cyclic columns (see gpk_gram_kernel)
      double d2 = 0.0;
#pragma unroll
      for (int d = 0; d < DP; ++d) { double df = xi[li][d] - xj[lj][d]; d2 += df * df; }
      out[b] = (i < Mq && j < N) ? p.c * exp(-0.5 * d2) : 0.0;
    }
    double* dst = Ks + (size_t)i * ldk + j0 + tx;
#pragma unroll
    for (int b = 0; b < 4; ++b) dst[16 * b] = out[b];
  }
}

// Gradient contraction (sklearn _gpr.py:629-651 without the (N,N,n_theta) tensor):
//   W = sum_t alpha_t alpha_t^T - T * Kinv ;   acc[d]   = sum_ij W_ij Kc_ij (dx_d/l_d)^2   (d < D)
//   acc[DP] = sum_ij W_ij Kc_ij  (d/dlog c),   acc[DP+1] = tr(W)  (d/dlog noise = 0.5*noise*tr W)
// over the lower tiles (off-diagonal pairs counted twice).  partial[(block), DP+2].
// General form: a (rows x cols) block of the matrix whose row / column entries carry the GLOBAL indices gi0 + r,
// gj0 + c (the distributed factorisation calls it once per owned tile; the single-GPU path once for the whole matrix
// with Xr == Xc, gi0 == gj0 == 0).  Only entries with global column <= global row are counted.
template <int DP>
__global__ void __launch_bounds__(256) gpk_grad_kernel(const double* __restrict__ Xr, const double* __restrict__ Xc, int ldx,
                                                       int nr, int nc, long long gi0, long long gj0, GpkKernelParams p,
                                                       const double* __restrict__ ar, const double* __restrict__ ac,
                                                       int lda, int T, const double* __restrict__ Kinv, int ldk,
                                                       double* __restrict__ partial, int accum) {
  const int tj = blockIdx.x, ti = blockIdx.y;
  const int bid = blockIdx.y * gridDim.x + blockIdx.x;
  __shared__ double xi[KTILE][DP + 1], xj[KTILE][DP + 1];
  __shared__ double ai[KTILE][GRAD_TC + 1], aj[KTILE][GRAD_TC + 1];
  __shared__ double red[8][DP + 2];
  const int tid = threadIdx.x, i0 = ti * KTILE, j0 = tj * KTILE;
  double acc[DP + 2];
#pragma unroll
  for (int d = 0; d < DP + 2; ++d) acc[d] = 0.0;
  if (gj0 + j0 <= gi0 + i0 + KTILE - 1) {
    load_scaled_rows<DP>(xi, Xr, ldx, i0, nr, p, tid);
    load_scaled_rows<DP>(xj, Xc, ldx, j0, nc, p, tid);
    for (int idx = tid; idx < KTILE * T; idx += 256) {
      int r = idx / T, t = idx % T;
      ai[r][t] = (i0 + r < nr) ? ar[(size_t)(i0 + r) * lda + t] : 0.0;
      aj[r][t] = (j0 + r < nc) ? ac[(size_t)(j0 + r) * lda + t] : 0.0;
    }
    __syncthreads();
    const int ty = tid >> 4, tx = tid & 15;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int li = ty * 4 + a, i = i0 + li;
      const long long gi = gi0 + i;
      const double* kr = Kinv + (size_t)i * ldk + j0 + tx;      // cyclic columns: 16 lanes read 128 contiguous bytes
      double kin[4] = {kr[0], kr[16], kr[32], kr[48]};
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int lj = tx + 16 * b, j = j0 + lj;
        const long long gj = gj0 + j;
        if (i >= nr || j >= nc || gj > gi) continue;
        double w = 0.0;
        for (int t = 0; t < T; ++t) w += ai[li][t] * aj[lj][t];
        w -= (double)T * kin[b];
        if (gi == gj) { acc[DP] += w * p.c; acc[DP + 1] += w; continue; }
        double sq[DP], d2 = 0.0;
#pragma unroll
        for (int d = 0; d < DP; ++d) { double df = xi[li][d] - xj[lj][d]; sq[d] = df * df; d2 += sq[d]; }
        double we = 2.0 * w * p.c * exp(-0.5 * d2);
#pragma unroll
        for (int d = 0; d < DP; ++d) acc[d] += we * sq[d];
        acc[DP] += we;
      }
    }
  }
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int d = 0; d < DP + 2; ++d) {
    double s = acc[d];
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) red[warp][d] = s;
  }
  __syncthreads();
  if (tid < DP + 2) {
    double s = 0.0;
    for (int w = 0; w < 8; ++w) s += red[w][tid];
    if (accum) s += partial[(size_t)bid * (DP + 2) + tid];
    partial[(size_t)bid * (DP + 2) + tid] = s;
  }
}

// Deterministic final reduction. out layout (doubles):
//   [0] lml  [1] sum log diag L  [2..2+T) y_t.alpha_t   [2+T ..) g_c, g_ls[0..D), g_noise  (already * 0.5)
__global__ void gpk_lml_finalize_kernel(const double* __restrict__ partial, int nblocks, int DPp2, int D, int N, int T,
                                        const double* __restrict__ L, size_t lstep, const double* __restrict__ y, int ldy,
                                        const double* __restrict__ alpha, int lda, GpkKernelParams p, int with_grad,
                                        double* __restrict__ out) {
  // one CTA per reduced quantity q (grid = 1 + T [+ DP + 2]); fixed summation order -> deterministic
  __shared__ double sh[256];
  const int tid = threadIdx.x;
  const int q = blockIdx.x;
  double s = 0.0;
  if (q == 0) { for (int i = tid; i < N; i += 256) s += log(L[(size_t)i * lstep]); }
  else if (q <= T) { int t = q - 1; for (int i = tid; i < N; i += 256) s += y[(size_t)i * ldy + t] * alpha[(size_t)i * lda + t]; }
  else { int d = q - 1 - T; for (int b = tid; b < nblocks; b += 256) s += partial[(size_t)b * DPp2 + d]; }
  sh[tid] = s;
  __syncthreads();
  for (int o = 128; o; o >>= 1) { if (tid < o) sh[tid] += sh[tid + o]; __syncthreads(); }
  if (tid == 0) {
    double v = sh[0];
    if (q == 0) out[1] = v;
    else if (q <= T) out[2 + q - 1] = v;
    else {
      int d = q - 1 - T, DP = DPp2 - 2;
      if (d < D) out[2 + T + 1 + d] = 0.5 * v;
      else if (d == DP) out[2 + T] = 0.5 * v;
      else if (d == DP + 1) out[2 + T + 1 + D] = 0.5 * p.noise * v;
    }
  }
}
// lml = sum_t (-1/2 y_t.alpha_t - sum log L_ii - N/2 log 2 pi)   (_gpr.py:613-617)
__global__ void gpk_lml_combine_kernel(int N, int T, double* __restrict__ out) {
  double lml = 0.0;
  for (int t = 0; t < T; ++t) lml += -0.5 * out[2 + t] - out[1] - 0.5 * (double)N * 1.8378770664093453;
  out[0] = lml;
}

// mean[q,t] = (sum_j Ks[q,j] alpha[j,t]) * y_std[t] + y_mean[t]; one warp per query row.
__global__ void gpk_predict_mean_kernel(const double* __restrict__ Ks, int ldk, int Mq, int N,
                                        const double* __restrict__ alpha, int lda, int T,
                                        const double* __restrict__ y_mean, const double* __restrict__ y_std,
                                        double* __restrict__ mean, int ldm) {
  int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= Mq) return;
  const double* kr = Ks + (size_t)row * ldk;
  for (int t0 = 0; t0 < T; t0 += 4) {
    double acc[4] = {0, 0, 0, 0};
    for (int j = lane; j < N; j += 32) {
      double k = kr[j];
#pragma unroll
      for (int t = 0; t < 4; ++t) if (t0 + t < T) acc[t] += k * alpha[(size_t)j * lda + t0 + t];
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      double s = acc[t];
      for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0 && t0 + t < T) mean[(size_t)row * ldm + t0 + t] = s * y_std[t0 + t] + y_mean[t0 + t];
    }
  }
}

// Mean-only predict, fused (sklearn _gpr.py:446-454 with return_std=False; what evaluate_gp_offline.py:198 and
// gp_enhanced_mpc.py:165,175 call): mean[q,t] = y_mean[t] + y_std[t] * sum_j c*exp(-1/2 |xq - x_j|^2_l) alpha[j,t].
// The cross-Gram row is consumed as it is produced and never stored: no M x N pass through HBM at all (the two-pass
// route writes and re-reads 8*M*N bytes).  Bound by FP64 exp + FMA issue: per (query, training row) pair 2*DP
// operations for the distance, ~25 for exp, TT for the accumulation; algorithmic HBM bytes per query 8*(D + T).
// CTA = 64 query rows x one slice of the training rows (blockIdx.y of `nsplit`), 64 training rows per step staged in
// shared memory with the alpha rows that go with them.  Thread (ty, tx) of a 32 x 8 grid keeps its two query rows
// (ty, ty+32) in registers and walks training rows tx+8b of the step (cyclic: lanes read shared memory with an odd row
// stride, conflict-free; the 8 tx lanes that share a query row are reduced with shuffles, fixed order).  partial[(split*MP + q)*T + t]; a second tiny kernel adds
// the splits in order and applies the target scale: no atomics, so a given (MP, N) always sums in the same order.
template <int DP, int TT>
__global__ void __launch_bounds__(256, 2) gpk_predict_mean_fused_kernel(
    const double* __restrict__ Xq, int ldq, int Mq, const double* __restrict__ X, int ldx, int N, GpkKernelParams p,
    const double* __restrict__ alpha, int lda, int t0, int T, double* __restrict__ partial, int MP, int tiles_per_split) {
  // training rows and their alpha rows are double-buffered: the next step's global loads are issued before this
  // step's arithmetic and parked in registers, so their latency hides behind ~700 FP64 instructions per thread
  __shared__ double xj[2][KTILE][DP + 1], aj[2][KTILE][TT + 1];
  const int tid = threadIdx.x, tx = tid & 7, ty = tid >> 3;      // 32 x 8 threads: query rows ty, ty+32; columns tx+8b
  const int i0 = blockIdx.x * KTILE, split = blockIdx.y;
  const int ntiles = (N + KTILE - 1) / KTILE;
  const int jt0 = split * tiles_per_split, jt1 = min(ntiles, jt0 + tiles_per_split);
  constexpr int NX = (KTILE * DP + 255) / 256, NA = (KTILE * TT + 255) / 256;   // staged elements per thread
  double px[NX], pa[NA];
  auto fetch = [&](int jt) {          // global -> registers (scaled inputs, alpha; zero beyond N / T)
    const int j0 = jt * KTILE;
#pragma unroll
    for (int u = 0; u < NX; ++u) {
      const int idx = tid + 256 * u, r = idx / DP, d = idx % DP;
      px[u] = (idx < KTILE * DP && j0 + r < N && d < p.D) ? X[(size_t)(j0 + r) * ldx + d] * p.inv_ls[d] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < NA; ++u) {
      const int idx = tid + 256 * u, r = idx / TT, t = idx % TT;
      pa[u] = (idx < KTILE * TT && j0 + r < N && t0 + t < T) ? alpha[(size_t)(j0 + r) * lda + t0 + t] : 0.0;
    }
  };
  auto park = [&](int buf) {          // registers -> shared
#pragma unroll
    for (int u = 0; u < NX; ++u) { const int idx = tid + 256 * u; if (idx < KTILE * DP) xj[buf][idx / DP][idx % DP] = px[u]; }
#pragma unroll
    for (int u = 0; u < NA; ++u) { const int idx = tid + 256 * u; if (idx < KTILE * TT) aj[buf][idx / TT][idx % TT] = pa[u]; }
  };
  if (jt0 < jt1) { fetch(jt0); park(0); }
  __syncthreads();
  double xq0[DP], xq1[DP], acc0[TT], acc1[TT];   // this thread's two query rows, scaled, straight from global (once)
#pragma unroll
  for (int d = 0; d < DP; ++d) {
    const bool dv = d < p.D;
    xq0[d] = (dv && i0 + ty < Mq) ? Xq[(size_t)(i0 + ty) * ldq + d] * p.inv_ls[d] : 0.0;
    xq1[d] = (dv && i0 + ty + 32 < Mq) ? Xq[(size_t)(i0 + ty + 32) * ldq + d] * p.inv_ls[d] : 0.0;
  }
#pragma unroll
  for (int t = 0; t < TT; ++t) acc0[t] = acc1[t] = 0.0;
  for (int jt = jt0; jt < jt1; ++jt) {
    const int cur = (jt - jt0) & 1;
    if (jt + 1 < jt1) fetch(jt + 1);
#pragma unroll 2
    for (int b = 0; b < 8; ++b) {
      const int lj = tx + 8 * b;
      double d20 = 0.0, d21 = 0.0;
#pragma unroll
      for (int d = 0; d < DP; ++d) {
        const double x = xj[cur][lj][d], f0 = xq0[d] - x, f1 = xq1[d] - x;
        d20 += f0 * f0; d21 += f1 * f1;
      }
      const double k0 = exp(-0.5 * d20), k1 = exp(-0.5 * d21);
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        const double av = aj[cur][lj][t];
        acc0[t] = fma(k0, av, acc0[t]); acc1[t] = fma(k1, av, acc1[t]);
      }
    }
    if (jt + 1 < jt1) park(cur ^ 1);   // the other buffer was last read in step jt - 1, behind the barrier below
    __syncthreads();
  }
#pragma unroll
  for (int t = 0; t < TT; ++t) {
    double s0 = acc0[t], s1 = acc1[t];
#pragma unroll
    for (int o = 4; o; o >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o, 8); s1 += __shfl_xor_sync(0xffffffffu, s1, o, 8); }
    if (tx == 0 && t0 + t < T) {
      partial[((size_t)split * MP + i0 + ty) * T + t0 + t] = p.c * s0;
      partial[((size_t)split * MP + i0 + ty + 32) * T + t0 + t] = p.c * s1;
    }
  }
}
// mean[q,t] = (sum_split partial) * y_std[t] + y_mean[t]
__global__ void gpk_predict_mean_finish_kernel(const double* __restrict__ partial, int nsplit, int MP, int Mq, int T,
                                               const double* __restrict__ y_mean, const double* __restrict__ y_std,
                                               double* __restrict__ mean) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Mq * T) return;
  const int t = idx % T;
  double s = 0.0;
  for (int k = 0; k < nsplit; ++k) s += partial[(size_t)k * MP * T + idx];
  mean[idx] = s * y_std[t] + y_mean[t];
}

// Few query rows (control-tick calls): one CTA per row, so a single row at N=20k is spread over 8 warps
// (the warp-per-row kernel above takes 0.31 ms for it).  Fixed reduction order -> deterministic.
__global__ void __launch_bounds__(256) gpk_predict_mean_cta_kernel(const double* __restrict__ Ks, int ldk, int N,
                                                                   const double* __restrict__ alpha, int lda, int T,
                                                                   const double* __restrict__ y_mean,
                                                                   const double* __restrict__ y_std,
                                                                   double* __restrict__ mean, int ldm) {
  __shared__ double red[4][8];
  const int row = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double* kr = Ks + (size_t)row * ldk;
  for (int t0 = 0; t0 < T; t0 += 4) {
    double acc[4] = {0, 0, 0, 0};
    for (int j = threadIdx.x; j < N; j += 256) {
      const double k = kr[j];
#pragma unroll
      for (int t = 0; t < 4; ++t) if (t0 + t < T) acc[t] += k * alpha[(size_t)j * lda + t0 + t];
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      double s = acc[t];
      for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) red[t][warp] = s;
    }
    __syncthreads();
    if (threadIdx.x < 4 && t0 + (int)threadIdx.x < T) {
      const double* r = red[threadIdx.x];
      const double s = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
      mean[(size_t)row * ldm + t0 + threadIdx.x] = s * y_std[t0 + threadIdx.x] + y_mean[t0 + threadIdx.x];
    }
    __syncthreads();
  }
}

// var = prior - sum_tiles rowsq ; negatives clamped to 0 and counted (sklearn _gpr.py:480-500)
__global__ void gpk_predict_std_kernel(const double* __restrict__ rowsq, int ntiles, int Mq, double prior,
                                       const double* __restrict__ y_std, int T, double* __restrict__ std_out,
                                       int lds, int* __restrict__ neg_count) {
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Mq) return;
  double s = 0.0;
  for (int t = 0; t < ntiles; ++t) s += rowsq[(size_t)q * ntiles + t];
  double var = prior - s;
  if (var < 0.0) { atomicAdd(neg_count, 1); var = 0.0; }
  for (int t = 0; t < T; ++t) std_out[(size_t)q * lds + t] = sqrt(var * (y_std[t] * y_std[t]));
}

// Small query batches (the 1-row control-tick calls): ||L^-1 k*||^2 as a triangular mat-vec against up to MQ
// cross-Gram rows at once, HBM/L2-bound on the N^2/2 doubles of L^-1 (the padded 128-row DMMA contraction would
// run on NP/128 CTAs only).  CTA b owns rows [32b, 32b+32) (4 per warp) and writes
// part[q, b] = sum_rows (Mlow[i, :i+1] . Ks[q, :i+1])^2  -- the layout gpk_predict_std_kernel sums.
template <int MQ>
__global__ void __launch_bounds__(256) gpk_small_rowsumsq_kernel(const double* __restrict__ Mlow, int NP,
                                                                 const double* __restrict__ Ks, int mq,
                                                                 double* __restrict__ part, int ldp) {
  __shared__ double red[MQ][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double sq[MQ];
#pragma unroll
  for (int q = 0; q < MQ; ++q) sq[q] = 0.0;
#pragma unroll 1
  for (int rr = 0; rr < 4; ++rr) {
    const int i = blockIdx.x * 32 + warp + 8 * rr;
    const double* mrow = Mlow + (size_t)i * NP;
    double acc[MQ];
#pragma unroll
    for (int q = 0; q < MQ; ++q) acc[q] = 0.0;
    for (int k = 2 * lane; k <= i; k += 64) {
      double2 m = *reinterpret_cast<const double2*>(mrow + k);
      if (k + 1 > i) m.y = 0.0;
#pragma unroll
      for (int q = 0; q < MQ; ++q)
        if (q < mq) {
          const double2 v = *reinterpret_cast<const double2*>(Ks + (size_t)q * NP + k);
          acc[q] = fma(m.x, v.x, fma(m.y, v.y, acc[q]));
        }
    }
#pragma unroll
    for (int q = 0; q < MQ; ++q) {
      double a = acc[q];
#pragma unroll
      for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
      sq[q] = fma(a, a, sq[q]);
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int q = 0; q < MQ; ++q) red[q][warp] = sq[q];
  }
  __syncthreads();
  if (threadIdx.x < MQ && (int)threadIdx.x < mq) {
    const double* r = red[threadIdx.x];
    part[(size_t)threadIdx.x * ldp + blockIdx.x] = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
  }
}
// part: (Mq x NP/32); Mq <= 32 is handled 8 rows per launch
cudaError_t gpk_launch_small_rowsumsq(const double* Mlow, int NP, const double* Ks, int Mq, double* part,
                                      cudaStream_t st) {
  const int nb = NP / 32;
  for (int q0 = 0; q0 < Mq; q0 += 8) {
    const int mq = Mq - q0 < 8 ? Mq - q0 : 8;
    const double* ks = Ks + (size_t)q0 * NP;
    double* pp = part + (size_t)q0 * nb;
    if (mq == 1) gpk_small_rowsumsq_kernel<1><<<nb, 256, 0, st>>>(Mlow, NP, ks, mq, pp, nb);
    else gpk_small_rowsumsq_kernel<8><<<nb, 256, 0, st>>>(Mlow, NP, ks, mq, pp, nb);
    GPK_LAUNCHED();
  }
  return cudaGetLastError();
}

// dst (N x N, ld N) = lower triangle of src's leading N x N block, zeros above (scipy's L_ layout)
__global__ void gpk_extract_lower_kernel(const double* __restrict__ src, int lds, double* __restrict__ dst, int N) {
  int c = blockIdx.y * blockDim.x + threadIdx.x, r = blockIdx.x;   // rows in grid.x: no 65535 limit (N > 65535)
  if (c < N) dst[(size_t)r * N + c] = (c <= r) ? src[(size_t)r * lds + c] : 0.0;
}

// zero-pad copy: dst (RP x ldd) <- src (R x C), rest zero
__global__ void gpk_pad_copy_kernel(const double* __restrict__ src, int lds, int R, int C, double* __restrict__ dst,
                                    int ldd, int RP, int CP) {
  int c = blockIdx.y * blockDim.x + threadIdx.x, r = blockIdx.x;   // rows in grid.x: no 65535 limit
  if (c < CP && r < RP) dst[(size_t)r * ldd + c] = (r < R && c < C) ? src[(size_t)r * lds + c] : 0.0;
}

// ---- launchers -----------------------------------------------------------------------------
#define DISPATCH_DP(D, CALL)                       \
  do {                                             \
    if ((D) <= 4) { CALL(4); }                     \
    else if ((D) <= 8) { CALL(8); }                \
    else if ((D) <= 10) { CALL(10); }              \
    else if ((D) <= 12) { CALL(12); }              \
    else if ((D) <= 16) { CALL(16); }              \
    else { CALL(32); }                             \
  } while (0)

int gpk_dp_of(int D) { return D <= 4 ? 4 : D <= 8 ? 8 : D <= 10 ? 10 : D <= 12 ? 12 : D <= 16 ? 16 : 32; }

cudaError_t gpk_launch_gram(const double* X, int N, int D, const GpkKernelParams& p, double* K, int NP,
                            int lower_only, cudaStream_t st) {
  dim3 grid(NP / KTILE, NP / KTILE);
#define CALL(DPV) gpk_gram_kernel<DPV><<<grid, 256, 0, st>>>(X, D, N, NP, p, K, NP, lower_only)
  DISPATCH_DP(D, CALL); GPK_LAUNCHED();
#undef CALL
  return cudaGetLastError();
}

cudaError_t gpk_launch_cross_gram(const double* Xq, int Mq, const double* X, int N, int D, const GpkKernelParams& p,
                                  double* Ks, int MP, int NP, cudaStream_t st) {
  dim3 grid(NP / KTILE, MP / KTILE);
#define CALL(DPV) gpk_cross_gram_kernel<DPV><<<grid, 256, 0, st>>>(Xq, D, Mq, X, D, N, p, Ks, NP)
  DISPATCH_DP(D, CALL); GPK_LAUNCHED();
#undef CALL
  return cudaGetLastError();
}

// ws: (NP/64)^2 * (DP+2) doubles
cudaError_t gpk_launch_lml_grad(const double* X, int N, int D, const GpkKernelParams& p, const double* L,
                                size_t lstep, const double* y, const double* alpha, int T, const double* Kinv, int NP,
                                int with_grad, double* ws, double* out, cudaStream_t st) {
  const int DP = gpk_dp_of(D);
  const int nt = NP / KTILE, nblocks = nt * nt;
  if (with_grad) {
    dim3 grid(nt, nt);
    for (int t0 = 0; t0 < T; t0 += GRAD_TC) {
      const int tc = (T - t0 < GRAD_TC) ? (T - t0) : GRAD_TC;
#define CALL(DPV) gpk_grad_kernel<DPV><<<grid, 256, 0, st>>>(X, X, D, N, N, 0, 0, p, alpha + t0, alpha + t0, T, tc, Kinv, NP, ws, t0 > 0)
      DISPATCH_DP(D, CALL); GPK_LAUNCHED();
#undef CALL
    }
  }
  gpk_lml_finalize_kernel<<<1 + T + (with_grad ? DP + 2 : 0), 256, 0, st>>>(ws, nblocks, DP + 2, D, N, T, L, lstep, y, T, alpha, T, p, with_grad, out); GPK_LAUNCHED();
  gpk_lml_combine_kernel<<<1, 1, 0, st>>>(N, T, out); GPK_LAUNCHED();
  return cudaGetLastError();
}

// One (rows x cols) block with global offsets (distributed gradient): partial sums ACCUMULATE into
// ws ((rows/64) * (cols/64) * (DP+2) doubles, zeroed by the caller before the first block) when accum != 0.
cudaError_t gpk_launch_grad_block(const double* Xr, int nr, long long gi0, const double* Xc, int nc, long long gj0, int D,
                                  const GpkKernelParams& p, const double* ar, const double* ac, int T, const double* Kinv,
                                  int ldk, int rows, int cols, double* ws, int accum, cudaStream_t st) {
  if (T > GRAD_TC || rows % KTILE || cols % KTILE) return cudaErrorInvalidValue;
  dim3 grid(cols / KTILE, rows / KTILE);
#define CALL(DPV) gpk_grad_kernel<DPV><<<grid, 256, 0, st>>>(Xr, Xc, D, nr, nc, gi0, gj0, p, ar, ac, T, T, Kinv, ldk, ws, accum)
  DISPATCH_DP(D, CALL); GPK_LAUNCHED();
#undef CALL
  return cudaGetLastError();
}
// out[d] = sum over blocks of ws[b][d], d < DP+2, fixed order (one thread per quantity)
__global__ void gpk_grad_sum_kernel(const double* __restrict__ ws, int nblocks, int DPp2, double* __restrict__ out) {
  const int d = threadIdx.x;
  if (d >= DPp2) return;
  double s = 0.0;
  for (int b = 0; b < nblocks; ++b) s += ws[(size_t)b * DPp2 + d];
  out[d] = s;
}
cudaError_t gpk_launch_grad_sum(const double* ws, int nblocks, int D, double* out, cudaStream_t st) {
  gpk_grad_sum_kernel<<<1, 64, 0, st>>>(ws, nblocks, gpk_dp_of(D) + 2, out); GPK_LAUNCHED();
  return cudaGetLastError();
}

cudaError_t gpk_launch_predict_mean(const double* Ks, int Mq, int N, int NP, const double* alpha, int T,
                                    const double* y_mean, const double* y_std, double* mean, cudaStream_t st) {
  if (Mq <= 1024) {   // fewer than ~7 warps per SM with a warp per row: give each row a CTA instead
    gpk_predict_mean_cta_kernel<<<Mq, 256, 0, st>>>(Ks, NP, N, alpha, T, T, y_mean, y_std, mean, T); GPK_LAUNCHED();
    return cudaGetLastError();
  }
  int blocks = (int)(((size_t)Mq * 32 + 255) / 256);
  gpk_predict_mean_kernel<<<blocks, 256, 0, st>>>(Ks, NP, Mq, N, alpha, T, T, y_mean, y_std, mean, T); GPK_LAUNCHED();
  return cudaGetLastError();
}

// number of training-row slices the fused mean kernel is split into: enough CTAs for two per SM when there are few
// query rows (the 25-row control tick), one slice when the query tiles alone fill the GPU
int gpk_mean_fused_splits(int MP, int N) {
  const int rowtiles = MP / KTILE, ntiles = (N + KTILE - 1) / KTILE;
  int ns = (2 * 148 + rowtiles - 1) / rowtiles;
  if (ns > ntiles) ns = ntiles;
  return ns < 1 ? 1 : ns;
}
// partial: nsplit * MP * T doubles
cudaError_t gpk_launch_predict_mean_fused(const double* Xq, int Mq, int MP, const double* X, int N, int D,
                                          const GpkKernelParams& p, const double* alpha, int T, const double* y_mean,
                                          const double* y_std, double* partial, double* mean, cudaStream_t st) {
  const int ntiles = (N + KTILE - 1) / KTILE, nsplit = gpk_mean_fused_splits(MP, N);
  const int tps = (ntiles + nsplit - 1) / nsplit;
  dim3 grid(MP / KTILE, nsplit);
  for (int t0 = 0; t0 < T; t0 += 8) {
    const int tc = T - t0;
#define CALLF(DPV, TTV) gpk_predict_mean_fused_kernel<DPV, TTV><<<grid, 256, 0, st>>>(Xq, D, Mq, X, D, N, p, alpha, T, t0, T, partial, MP, tps)
#define CALLD(DPV) do { if (tc <= 4) CALLF(DPV, 4); else CALLF(DPV, 8); } while (0)
    if (D <= 4) CALLD(4); else if (D <= 8) CALLD(8); else if (D <= 10) CALLD(10); else if (D <= 12) CALLD(12);
    else if (D <= 16) CALLD(16); else CALLD(32);
#undef CALLD
#undef CALLF
    GPK_LAUNCHED();
  }
  gpk_predict_mean_finish_kernel<<<(Mq * T + 255) / 256, 256, 0, st>>>(partial, nsplit, MP, Mq, T, y_mean, y_std, mean); GPK_LAUNCHED();
  return cudaGetLastError();
}

cudaError_t gpk_launch_predict_std(const double* rowsq, int ntiles, int Mq, double prior, const double* y_std, int T,
                                   double* std_out, int* neg, cudaStream_t st) {
  gpk_predict_std_kernel<<<(Mq + 255) / 256, 256, 0, st>>>(rowsq, ntiles, Mq, prior, y_std, T, std_out, T, neg); GPK_LAUNCHED();
  return cudaGetLastError();
}

cudaError_t gpk_launch_extract_lower(const double* src, int NP, double* dst, int N, cudaStream_t st) {
  dim3 grid(N, (N + 255) / 256);
  gpk_extract_lower_kernel<<<grid, 256, 0, st>>>(src, NP, dst, N); GPK_LAUNCHED();
  return cudaGetLastError();
}

cudaError_t gpk_launch_pad_copy(const double* src, int R, int C, double* dst, int RP, int CP, cudaStream_t st) {
  dim3 grid(RP, (CP + 127) / 128);
  gpk_pad_copy_kernel<<<grid, 128, 0, st>>>(src, C, R, C, dst, CP, RP, CP); GPK_LAUNCHED();
  return cudaGetLastError();
}

__global__ void gpk_copy_diag_kernel(const double* __restrict__ A, int lda, int n, double* __restrict__ d) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) d[i] = A[(size_t)i * lda + i];
}
cudaError_t gpk_launch_copy_diag(const double* A, int lda, int n, double* d, cudaStream_t st) {
  gpk_copy_diag_kernel<<<(n + 255) / 256, 256, 0, st>>>(A, lda, n, d); GPK_LAUNCHED();
  return cudaGetLastError();
}

// ---- small dense helpers used by the distributed factorisation (dist_chol.py) -----------------
// out[m,t] = beta*out[m,t] + alpha * sum_k A[m,k] v[k,t]; one warp per row
__global__ void gpk_gemv_n_kernel(const double* __restrict__ A, int lda, int M, int K, const double* __restrict__ v,
                                  int T, double* __restrict__ out, double alpha, double beta) {
  int row = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const double* ar = A + (size_t)row * lda;
  for (int t0 = 0; t0 < T; t0 += 4) {
    double acc[4] = {0, 0, 0, 0};
    for (int k = lane; k < K; k += 32) {
      double a = ar[k];
#pragma unroll
      for (int t = 0; t < 4; ++t) if (t0 + t < T) acc[t] += a * v[(size_t)k * T + t0 + t];
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      double s = acc[t];
      for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0 && t0 + t < T) {
        double* o = out + (size_t)row * T + t0 + t;
        *o = (beta == 0.0 ? 0.0 : beta * *o) + alpha * s;
      }
    }
  }
}
// out[c,t] += alpha * sum_r A[r,c] v[r,t]  (out pre-zeroed by the caller when needed); atomics over row chunks
__global__ void gpk_gemv_t_kernel(const double* __restrict__ A, int lda, int M, int K, const double* __restrict__ v,
                                  int T, double* __restrict__ out, double alpha) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  int r0 = blockIdx.y * 512, r1 = min(M, r0 + 512);
  if (c >= K) return;
  for (int t0 = 0; t0 < T; t0 += 4) {
    double acc[4] = {0, 0, 0, 0};
    for (int r = r0; r < r1; ++r) {
      double a = A[(size_t)r * lda + c];
#pragma unroll
      for (int t = 0; t < 4; ++t) if (t0 + t < T) acc[t] += a * v[(size_t)r * T + t0 + t];
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) if (t0 + t < T) atomicAdd(out + (size_t)c * T + t0 + t, alpha * acc[t]);
  }
}
// A[i,i] = (i < n_valid) ? A[i,i] + add : pad_value
__global__ void gpk_fix_diag_kernel(double* __restrict__ A, int lda, int n, int n_valid, double add, double pad_value) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { double* p = A + (size_t)i * lda + i; *p = (i < n_valid) ? (*p + add) : pad_value; }
}
// s[0] += sum_i log A[i,i], i < n  (single block)
__global__ void gpk_logdiag_kernel(const double* __restrict__ A, int lda, int n, double* __restrict__ s) {
  __shared__ double sh[256];
  double a = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) a += log(A[(size_t)i * lda + i]);
  sh[threadIdx.x] = a; __syncthreads();
  for (int o = 128; o; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) s[0] += sh[0];
}
cudaError_t gpk_launch_gemv_n(const double* A, int lda, int M, int K, const double* v, int T, double* out, double alpha,
                              double beta, cudaStream_t st) {
  if (M <= 0) return cudaSuccess;
  gpk_gemv_n_kernel<<<(unsigned)(((size_t)M * 32 + 255) / 256), 256, 0, st>>>(A, lda, M, K, v, T, out, alpha, beta); GPK_LAUNCHED();
  return cudaGetLastError();
}
cudaError_t gpk_launch_gemv_t(const double* A, int lda, int M, int K, const double* v, int T, double* out, double alpha,
                              cudaStream_t st) {
  if (M <= 0) return cudaSuccess;
  dim3 grid((K + 127) / 128, (M + 511) / 512);
  gpk_gemv_t_kernel<<<grid, 128, 0, st>>>(A, lda, M, K, v, T, out, alpha); GPK_LAUNCHED();
  return cudaGetLastError();
}
cudaError_t gpk_launch_fix_diag(double* A, int lda, int n, int n_valid, double add, double pad, cudaStream_t st) {
  gpk_fix_diag_kernel<<<(n + 255) / 256, 256, 0, st>>>(A, lda, n, n_valid, add, pad); GPK_LAUNCHED();
  return cudaGetLastError();
}
cudaError_t gpk_launch_logdiag(const double* A, int lda, int n, double* s, cudaStream_t st) {
  gpk_logdiag_kernel<<<1, 256, 0, st>>>(A, lda, n, s); GPK_LAUNCHED();
  return cudaGetLastError();
}

// ---- residual metrics behind the batched predict (evaluate_gp_offline.py:277-312) -------------
// For true residuals R (M x C) and GP-predicted residuals G (M x C), C <= 8, per component j:
//   S[j][0] = sum t          S[j][1] = sum t^2 (nominal model = zero residual: its squared error)
//   S[j][2] = sum (t - g)^2  S[j][3] = #rows with t^2 > (t - g)^2      S[j][4] = sum (t - mean_j)^2   (second pass)
// and over whole rows (se = squared error summed over components; acc = components 3..5 when C >= 6):
//   Gl[0] = sum se_nom, Gl[1] = sum se_gp, Gl[2] = #rows se_nom > se_gp, Gl[3] = #rows se_nom < se_gp,
//   Gl[4] = sum acc_nom, Gl[5] = sum acc_gp.
// HBM-bound: 16*C bytes per row, each read once per pass (the second pass re-reads R only).  Two stages, fixed order:
// every CTA owns a contiguous row range and leaves one partial record; one CTA adds the records in order.
#define MET_C 8
#define MET_REC (MET_C * 5 + 6)
__device__ __forceinline__ double met_block_sum(double v, double* sh) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0) for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += sh[w];
  return s;   // valid in thread 0
}
__global__ void __launch_bounds__(256) gpk_metrics_pass1_kernel(const double* __restrict__ R, const double* __restrict__ G,
                                                                int M, int Cc, int rows_per_block,
                                                                double* __restrict__ partial) {
  __shared__ double sh[8];
  const int r0 = blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
  double st[MET_C], st2[MET_C], se2[MET_C], nb[MET_C], gl[6];
#pragma unroll
  for (int j = 0; j < MET_C; ++j) st[j] = st2[j] = se2[j] = nb[j] = 0.0;
#pragma unroll
  for (int q = 0; q < 6; ++q) gl[q] = 0.0;
  for (int r = r0 + threadIdx.x; r < r1; r += 256) {
    double sn = 0.0, sg = 0.0, an = 0.0, ag = 0.0;
#pragma unroll
    for (int j = 0; j < MET_C; ++j)
      if (j < Cc) {
        const double t = R[(size_t)r * Cc + j], e = t - G[(size_t)r * Cc + j], t2 = t * t, e2 = e * e;
        st[j] += t; st2[j] += t2; se2[j] += e2; nb[j] += (t2 > e2) ? 1.0 : 0.0;
        sn += t2; sg += e2;
        if (j >= 3 && j < 6) { an += t2; ag += e2; }
      }
    gl[0] += sn; gl[1] += sg; gl[2] += (sn > sg) ? 1.0 : 0.0; gl[3] += (sn < sg) ? 1.0 : 0.0; gl[4] += an; gl[5] += ag;
  }
  double* rec = partial + (size_t)blockIdx.x * MET_REC;
#pragma unroll
  for (int j = 0; j < MET_C; ++j) {
    double a = met_block_sum(st[j], sh), b = met_block_sum(st2[j], sh), c = met_block_sum(se2[j], sh), d = met_block_sum(nb[j], sh);
    if (threadIdx.x == 0) { rec[j * 5 + 0] = a; rec[j * 5 + 1] = b; rec[j * 5 + 2] = c; rec[j * 5 + 3] = d; rec[j * 5 + 4] = 0.0; }
  }
#pragma unroll
  for (int q = 0; q < 6; ++q) { double a = met_block_sum(gl[q], sh); if (threadIdx.x == 0) rec[MET_C * 5 + q] = a; }
}
__global__ void __launch_bounds__(256) gpk_metrics_pass2_kernel(const double* __restrict__ R, int M, int Cc, int rows_per_block,
                                                                const double* __restrict__ out, double* __restrict__ partial) {
  __shared__ double sh[8];
  const int r0 = blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
  double dv[MET_C], mu[MET_C];
#pragma unroll
  for (int j = 0; j < MET_C; ++j) { dv[j] = 0.0; mu[j] = (j < Cc) ? out[j * 5 + 0] / (double)M : 0.0; }
  for (int r = r0 + threadIdx.x; r < r1; r += 256)
#pragma unroll
    for (int j = 0; j < MET_C; ++j)
      if (j < Cc) { const double d = R[(size_t)r * Cc + j] - mu[j]; dv[j] += d * d; }
#pragma unroll
  for (int j = 0; j < MET_C; ++j) { double a = met_block_sum(dv[j], sh); if (threadIdx.x == 0) partial[(size_t)blockIdx.x * MET_REC + j * 5 + 4] = a; }
}
// out[q] = sum over blocks of partial[b][q] for q in [q0, q1) with (q % 5 == 4) == second_pass among the component slots
__global__ void gpk_metrics_reduce_kernel(const double* __restrict__ partial, int nblocks, int second_pass, double* __restrict__ out) {
  const int q = threadIdx.x;
  if (q >= MET_REC) return;
  const bool comp = q < MET_C * 5, dev = comp && (q % 5 == 4);
  if (second_pass ? !dev : dev) return;
  double s = 0.0;
  for (int b = 0; b < nblocks; ++b) s += partial[(size_t)b * MET_REC + q];
  out[q] = s;
}
int gpk_metrics_blocks(int M) { int b = (M + 1023) / 1024; return b < 1 ? 1 : (b > 592 ? 592 : b); }
int gpk_metrics_rec() { return MET_REC; }
cudaError_t gpk_launch_residual_metrics(const double* R, const double* G, int M, int Cc, double* partial, double* out,
                                        cudaStream_t st) {
  const int nb = gpk_metrics_blocks(M), rpb = (M + nb - 1) / nb;
  gpk_metrics_pass1_kernel<<<nb, 256, 0, st>>>(R, G, M, Cc, rpb, partial); GPK_LAUNCHED();
  gpk_metrics_reduce_kernel<<<1, 64, 0, st>>>(partial, nb, 0, out); GPK_LAUNCHED();
  gpk_metrics_pass2_kernel<<<nb, 256, 0, st>>>(R, M, Cc, rpb, out, partial); GPK_LAUNCHED();
  gpk_metrics_reduce_kernel<<<1, 64, 0, st>>>(partial, nb, 1, out); GPK_LAUNCHED();
  return cudaGetLastError();
}
