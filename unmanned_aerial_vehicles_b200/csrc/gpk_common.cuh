// gpk_common.cuh -- shared declarations for the GP hot-path kernels (sm_100a, FP64).
//
// Layout convention used by every kernel in this library:
//   * all matrices are row-major float64 with an explicit leading dimension;
//   * every square factorisation buffer is padded to NP = gpk_pad(N), a multiple of
//     GPK_TILE (128).  The padding block of the Gram matrix is the identity, so its
//     Cholesky factor / inverse are the identity as well and log-det, alpha and the
//     gradient are unaffected.  Kernels therefore never bounds-check tiles.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GPK_TILE 128          // CTA tile edge of the DMMA GEMM and the Cholesky leaf
#define GPK_ROWSQ_TILE 64     // column-tile width of the GPK_EPI_ROWSUMSQ partial sums (rowsq is M x N/64)
#define GPK_MAXD 32           // max feature dimension supported by the kernel-function kernels
#define GPK_MAXT 16           // max number of targets handled per launch of the small-T kernels

// Hyper-parameters of  c * RBF(length_scale) + WhiteKernel(noise)  passed by value.
struct GpkKernelParams {
  double inv_ls[GPK_MAXD];    // 1/length_scale per feature (iso: all equal); 0 beyond D
  double c;                   // ConstantKernel value (1.0 when absent)
  double noise;               // WhiteKernel noise_level (0 when absent)
  double jitter;              // GaussianProcessRegressor.alpha added to the diagonal
  int D;
};

// triangular-structure hints for the NT GEMM (k-range trimming; the zero triangle of the
// operand's diagonal tiles must really hold zeros)
enum { GPK_TRI_NONE = 0, GPK_TRI_LOWER = 1, GPK_TRI_UPPER = 2 };
// epilogues
enum { GPK_EPI_STORE = 0, GPK_EPI_ROWSUMSQ = 1 };

struct GpkGemmArgs {
  const double* A; int lda;   // (M x K) row-major
  const double* B; int ldb;   // (N x K) row-major   -> C = alpha * A * B^T (+ C)
  double* C; int ldc;         // (M x N) row-major
  int M, N, K;                // all multiples of GPK_TILE
  double alpha;
  int accumulate;             // 0: C = alpha*AB^T ; 1: C += alpha*AB^T
  int a_tri, b_tri;           // GPK_TRI_*
  int lower_only;             // skip tiles strictly above the diagonal (needs M == N)
  int epilogue;               // GPK_EPI_*
  double* rowsq;              // GPK_EPI_ROWSUMSQ: (M x N/GPK_TILE) partial row sums of squares
  int raster_rows;            // tile-rows per raster group; 0 = the launcher's default (8, or GPK_GEMM_RASTER)
  double* Ct; int ldct;       // GPK_EPI_STORE, optional: the result is ALSO stored transposed, Ct (N x M) row-major
};

// internal launchers (C++ linkage), all asynchronous on `st`
cudaError_t gpk_launch_gemm_nt(const GpkGemmArgs& g, cudaStream_t st);
cudaError_t gpk_launch_copy_block(const double* src, int lds, double* dst, int ldd, int rows, int cols, cudaStream_t st);
cudaError_t gpk_launch_transpose_block(const double* src, int lds, double* dst, int ldd, int rows, int cols, cudaStream_t st);

// per-device one-time state (function attributes, side streams): indexed by the current device ordinal
#define GPK_MAX_DEV 64
static inline int gpk_current_device() { int d = 0; cudaGetDevice(&d); return (d >= 0 && d < GPK_MAX_DEV) ? d : 0; }

extern unsigned long long g_gpk_launches;   // kernels launched by this library (bench.py's gpu_launches)
#define GPK_LAUNCHED() ((void)__atomic_fetch_add(&g_gpk_launches, 1ULL, __ATOMIC_RELAXED))   // host threads may launch concurrently
// optional per-launch timing of the DMMA GEMM (bench.py roofline): event pairs around each launch
struct GpkProfile { int enabled; int n; double flops; };
extern GpkProfile g_gpk_prof;
#define GPK_CHECK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) return (int)(-1000 - (int)e__); } while (0)
