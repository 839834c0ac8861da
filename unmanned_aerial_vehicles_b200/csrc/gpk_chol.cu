// gpk_chol.cu -- recursive blocked FP64 Cholesky fused with the triangular inverse.
//
// Replaces scipy.linalg.cholesky (dpotrf; sklearn _gpr.py:352,591) and the N-right-hand-side
// cho_solve(L, I) sklearn uses to form K^-1 (_gpr.py:631, 2N^3 flops) by
//   potrf + trtri (2N^3/3) followed by lauum (N^3/3), all as large-K NT GEMMs on the FP64 tensor pipe.
//
// Buffers (NP x NP, NP % 128 == 0):
//   A    : in  = lower triangle of K (+ identity padding), out = lower Cholesky factor L
//   Mlow : out = L^-1 (lower; diagonal 128-blocks carry explicit zeros above the diagonal)
//   Mup  : out = L^-T (upper; = Mlow transposed)
// Recursion on [off, off+n):  n1 | n2 split at a multiple of 128
//   rec(11);  L21 = A21 * M11^T;  A22 -= L21 L21^T;  rec(22);  M21 = -M22 * L21 * M11.
#include "gpk_common.cuh"

#define LEAF_SMEM_BYTES 0

// One CTA factors a 128x128 diagonal block and inverts the factor in the same sweep.
// The block lives in REGISTERS: 16x16 threads, thread (ty,tx) owns S[ty+16a][tx+16b], a,b < 8 (2-D cyclic).
// Step j (one __syncthreads each, buffers double-buffered):
//   column j (raw) and row j (raw) are published to shared memory; d = sqrt(pivot);
//   for every row i > j and column c <= i:   c > j : A-part   S[i][c] -= l_ij * l_cj
//                                            c < j : M-part   S[i][c] -= l_ij * M[j][c]
//                                            c = j :          S[i][j]  = -l_ij / d
//   i.e. forward elimination applied to [A | I] at once, so S ends as L^-1 while column j of L is streamed out.
// info: first failing global pivot index + 1 (LAPACK convention), 0 if none.
#define LEAF_THREADS 512
#define LEAF_RA 4   // rows per thread: 128 / 32
// publish helpers: the owner index is uniform across the CTA, so a switch replaces 8-way register selects
#define LEAF_COL_CASE(B)                                                            \
  case B:                                                                           \
    _Pragma("unroll") for (int a = 0; a < LEAF_RA; ++a) {                           \
      const int i = ty + 32 * a;                                                    \
      if (i > j) buf[p][i] = S[a][B];                                               \
      else if (i == j) { piv[p] = S[a][B]; buf[p][j] = 1.0; }                       \
      S[a][B] = 0.0;  /* column j restarts as the M-part: S[i][j] = 0 - l_ij/d */   \
    }                                                                               \
    break;
#define LEAF_ROW_CASE(AA)                                                           \
  case AA:                                                                          \
    _Pragma("unroll") for (int b = 0; b < 8; ++b) {                                 \
      const int c = tx + 16 * b;                                                    \
      if (c < j) buf[p][c] = S[AA][b];                                              \
    }                                                                               \
    break;

__global__ void __launch_bounds__(LEAF_THREADS, 1)
gpk_leaf_potrf_inv_kernel(double* __restrict__ A, int lda, double* __restrict__ Mlow, double* __restrict__ Mup,
                          int ldm, int off, int* __restrict__ info) {
  // buf[p][x]: x < j  -> raw M[j][x] (row j);  x == j -> 1;  x > j -> raw A[x][j] (column j).  piv[p] = A[j][j].
  __shared__ double buf[2][GPK_TILE], piv[2];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;   // 32 x 16 threads; rows ty+32a, cols tx+16b
  double* Ab = A + (size_t)off * lda + off;
  double S[LEAF_RA][8];
#pragma unroll
  for (int a = 0; a < LEAF_RA; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int i = ty + 32 * a, c = tx + 16 * b;
      S[a][b] = (c <= i) ? Ab[(size_t)i * lda + c] : 0.0;
    }
  for (int j = 0; j < GPK_TILE; ++j) {
    const int p = j & 1;
    if (ty == (j & 31)) {          // row j first (reads S[j][c], c < j), then column j (which zeroes S[.][j])
      switch (j >> 5) { LEAF_ROW_CASE(0) LEAF_ROW_CASE(1) LEAF_ROW_CASE(2) LEAF_ROW_CASE(3) }
    }
    if (tx == (j & 15)) {
      switch (j >> 4) {
        LEAF_COL_CASE(0) LEAF_COL_CASE(1) LEAF_COL_CASE(2) LEAF_COL_CASE(3)
        LEAF_COL_CASE(4) LEAF_COL_CASE(5) LEAF_COL_CASE(6) LEAF_COL_CASE(7)
      }
    }
    __syncthreads();
    double d = piv[p];
    if (!(d > 0.0)) {  // non-positive or NaN pivot: record and poison
      if (tid == 0) { int want = off + j + 1; int old = atomicCAS(info, 0, want); if (old != 0 && old > want) atomicMin(info, want); }
      d = __longlong_as_double(0x7ff8000000000000LL);
    }
    // 1/sqrt(d) with one Newton step each on sqrt and reciprocal (no FP64 division on the critical path)
    double dinv = rsqrt(d);
    double ljj = d * dinv;
    ljj = fma(0.5 * dinv, fma(-ljj, ljj, d), ljj);
    dinv = fma(dinv, fma(-ljj, dinv, 1.0), dinv);
    double v[8];
#pragma unroll
    for (int b = 0; b < 8; ++b) v[b] = buf[p][tx + 16 * b] * dinv;
#pragma unroll
    for (int a = 0; a < LEAF_RA; ++a) {
      const int i = ty + 32 * a;
      if (i > j) {
        const double li = buf[p][i] * dinv;
        // whole row at once: c > j is the Cholesky update, c < j the inverse update, c == j gives -l_ij/d;
        // entries above the diagonal (c > i) hold finite garbage that nothing reads
#pragma unroll
        for (int b = 0; b < 8; ++b) S[a][b] = fma(-li, v[b], S[a][b]);
        if (tx == (j & 15)) Ab[(size_t)i * lda + j] = li;          // column j of L
      } else if (i == j) {                                         // row j of M becomes final
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          const int c = tx + 16 * b;
          if (c < j) S[a][b] *= dinv;
          else if (c == j) { S[a][b] = dinv; Ab[(size_t)j * lda + j] = ljj; }
        }
      }
    }
  }
  double* Ml = Mlow + (size_t)off * ldm + off;
  double* Mu = Mup + (size_t)off * ldm + off;
#pragma unroll
  for (int a = 0; a < LEAF_RA; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int i = ty + 32 * a, c = tx + 16 * b;
      const double m = (c <= i) ? S[a][b] : 0.0;
      Ml[(size_t)i * ldm + c] = m;
      Mu[(size_t)c * ldm + i] = m;
    }
}

struct CholCtx { double* A; int lda; double* Ml; double* Mu; int ldm; int* info; cudaStream_t st; cudaError_t err; };

static void chol_rec(CholCtx& c, int off, int n) {
  if (c.err != cudaSuccess) return;
  if (n <= GPK_TILE) {
    gpk_leaf_potrf_inv_kernel<<<1, LEAF_THREADS, LEAF_SMEM_BYTES, c.st>>>(c.A, c.lda, c.Ml, c.Mu, c.ldm, off, c.info); GPK_LAUNCHED();
    c.err = cudaGetLastError();
    return;
  }
  const int n1 = GPK_TILE * ((n / GPK_TILE + 1) / 2), n2 = n - n1, o2 = off + n1;
  chol_rec(c, off, n1);
  if (c.err != cudaSuccess) return;
  double* A21 = c.A + (size_t)o2 * c.lda + off;
  double* Ml21 = c.Ml + (size_t)o2 * c.ldm + off;
  double* Mu12 = c.Mu + (size_t)off * c.ldm + o2;
  GpkGemmArgs g{};
  // L21 = A21 * M11^T  (tmp in Mlow21, then copied over A21)
  g.A = A21; g.lda = c.lda; g.B = c.Ml + (size_t)off * c.ldm + off; g.ldb = c.ldm; g.C = Ml21; g.ldc = c.ldm;
  g.M = n2; g.N = n1; g.K = n1; g.alpha = 1.0; g.accumulate = 0; g.a_tri = GPK_TRI_NONE; g.b_tri = GPK_TRI_LOWER;
  g.lower_only = 0; g.epilogue = GPK_EPI_STORE; g.rowsq = nullptr;
  if ((c.err = gpk_launch_gemm_nt(g, c.st)) != cudaSuccess) return;
  if ((c.err = gpk_launch_copy_block(Ml21, c.ldm, A21, c.lda, n2, n1, c.st)) != cudaSuccess) return;
  // A22 -= L21 L21^T (lower tiles)
  g = GpkGemmArgs{};
  g.A = A21; g.lda = c.lda; g.B = A21; g.ldb = c.lda; g.C = c.A + (size_t)o2 * c.lda + o2; g.ldc = c.lda;
  g.M = n2; g.N = n2; g.K = n1; g.alpha = -1.0; g.accumulate = 1; g.lower_only = 1; g.epilogue = GPK_EPI_STORE;
  if ((c.err = gpk_launch_gemm_nt(g, c.st)) != cudaSuccess) return;
  chol_rec(c, o2, n2);
  if (c.err != cudaSuccess) return;
  // Wt = Mt11 * L21^T  -> Mup12 (n1 x n2)
  g = GpkGemmArgs{};
  g.A = c.Mu + (size_t)off * c.ldm + off; g.lda = c.ldm; g.B = A21; g.ldb = c.lda; g.C = Mu12; g.ldc = c.ldm;
  g.M = n1; g.N = n2; g.K = n1; g.alpha = 1.0; g.a_tri = GPK_TRI_UPPER; g.epilogue = GPK_EPI_STORE;
  if ((c.err = gpk_launch_gemm_nt(g, c.st)) != cudaSuccess) return;
  // M21 = -M22 * W  : C[m,j] = -sum_k M22[m,k] Wt[j,k]
  g = GpkGemmArgs{};
  g.A = c.Ml + (size_t)o2 * c.ldm + o2; g.lda = c.ldm; g.B = Mu12; g.ldb = c.ldm; g.C = Ml21; g.ldc = c.ldm;
  g.M = n2; g.N = n1; g.K = n2; g.alpha = -1.0; g.a_tri = GPK_TRI_LOWER; g.epilogue = GPK_EPI_STORE;
  if ((c.err = gpk_launch_gemm_nt(g, c.st)) != cudaSuccess) return;
  c.err = gpk_launch_transpose_block(Ml21, c.ldm, Mu12, c.ldm, n2, n1, c.st);
}

cudaError_t gpk_launch_potrf_inv(double* A, int lda, double* Mlow, double* Mup, int ldm, int NP, int* info,
                                 cudaStream_t st) {
  if (NP % GPK_TILE) return cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync(info, 0, sizeof(int), st);
  if (e != cudaSuccess) return e;
  CholCtx c{A, lda, Mlow, Mup, ldm, info, st, cudaSuccess};
  chol_rec(c, 0, NP);
  return c.err;
}

// K^-1 = L^-T L^-1 = Mup * Mup^T  (lower tiles of Kinv are written; diagonal tiles in full)
cudaError_t gpk_launch_lauum(const double* Mup, int ldm, double* Kinv, int ldk, int NP, cudaStream_t st) {
  GpkGemmArgs g{};
  g.A = Mup; g.lda = ldm; g.B = Mup; g.ldb = ldm; g.C = Kinv; g.ldc = ldk; g.M = NP; g.N = NP; g.K = NP;
  g.alpha = 1.0; g.accumulate = 0; g.a_tri = GPK_TRI_UPPER; g.b_tri = GPK_TRI_UPPER; g.lower_only = 1;
  g.epilogue = GPK_EPI_STORE;
  return gpk_launch_gemm_nt(g, st);
}

// out[i, t] = sum_k Mt[i, k] * v[k, t]  over k <= i (lower) or k >= i (upper); one warp per row.
template <int TT>
__global__ void gpk_trimv_kernel(const double* __restrict__ Mt, int ldm, int NP, const double* __restrict__ v,
                                 int ldv, double* __restrict__ out, int ldo, int upper, int t0, int T) {
  int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= NP) return;
  int kb = upper ? row : 0, ke = upper ? NP : row + 1;
  double acc[TT];
#pragma unroll
  for (int t = 0; t < TT; ++t) acc[t] = 0.0;
  const double* mr = Mt + (size_t)row * ldm;
  for (int k = kb + lane; k < ke; k += 32) {
    double m = mr[k];
#pragma unroll
    for (int t = 0; t < TT; ++t)
      if (t0 + t < T) acc[t] += m * v[(size_t)k * ldv + t0 + t];
  }
#pragma unroll
  for (int t = 0; t < TT; ++t) {
    double s = acc[t];
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0 && t0 + t < T) out[(size_t)row * ldo + t0 + t] = s;
  }
}
cudaError_t gpk_launch_trimv(const double* Mt, int ldm, int NP, const double* v, int ldv, double* out, int ldo,
                             int upper, int T, cudaStream_t st) {
  int blocks = (NP * 32 + 255) / 256;
  for (int t0 = 0; t0 < T; t0 += 4) {
    gpk_trimv_kernel<4><<<blocks, 256, 0, st>>>(Mt, ldm, NP, v, ldv, out, ldo, upper, t0, T); GPK_LAUNCHED();
  }
  return cudaGetLastError();
}
