// gpk_noalloc.cu -- ONE entry point that allocates nothing: gpk_predict_mean_fused_static.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops such as the 50 Hz control tick, harnesses that forbid
// it).  It is gpk_predict_mean_fused (mean-only predict, sklearn _gpr.py:446-454 with return_std=False) with the
// kernel's scratch -- the per-split partial sums, gpk_mean_fused_splits(MP, N) * MP * T doubles -- taken from a
// file-scope device array instead of a caller-supplied buffer.  Inputs and outputs are still the caller's.
//
//   * The array is GPK_STATIC_SCRATCH_DOUBLES doubles (4 MiB); every GPU that loads the library holds it until exit.
//     A call that needs more is rejected with GPK_ERR_STATIC_SCRATCH_TOO_SMALL before anything is launched.
//   * In HOST code the array's name is a host-side stand-in, not a device address (a kernel handed it would fault, or
//     on a GPU that can reach host memory silently use the wrong memory).  The device address is looked up here, in the
//     file that defines the array, once per device with that device current, by cudaGetSymbolAddress; the return code
//     is checked and the pointer cached.  That first call per device is the set-up: make it outside graph capture.
//   * No cudaMalloc / cudaFree / allocating library routine is called; the launches are the existing kernels'.
//   * Every call on a device shares the array, which keeps the last call's contents.  Nothing has to start at zero:
//     the fused kernel writes every (split, row < MP, target) slot the finishing kernel reads before it is read, so
//     stale contents are never observed.  Calls that overlap on the GPU (two streams, two host threads) RACE on it:
//     serialise them (one stream, or an event between them).
#include "gpk_common.cuh"
#include "../../include/gpk.h"
#include <mutex>
#include <string.h>

// launchers / helpers defined in gpk_kernels.cu
int gpk_mean_fused_splits(int MP, int N);
cudaError_t gpk_launch_predict_mean_fused(const double*, int, int, const double*, int, int, const GpkKernelParams&,
                                          const double*, int, const double*, const double*, double*, double*,
                                          cudaStream_t);

static __device__ __align__(256) double g_static_scratch[GPK_STATIC_SCRATCH_DOUBLES];

static double* g_static_scratch_ptr[GPK_MAX_DEV];   // device address of the array on each device (nullptr: not looked up)
static std::mutex g_static_scratch_mu;

// device address of g_static_scratch on the CURRENT device (looked up once per device)
static cudaError_t static_scratch_address(double** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= GPK_MAX_DEV) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (!g_static_scratch_ptr[dev]) {
    double* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    g_static_scratch_ptr[dev] = p;
  }
  *out = g_static_scratch_ptr[dev];
  return cudaSuccess;
}

extern "C" int gpk_predict_mean_fused_static(const double* Xq_d, int Mq, int MP, const double* X_d, int N, int D,
                                             const double* inv_ls_h, double c, const double* alpha_d, int T,
                                             const double* y_mean_d, const double* y_std_d, double* mean_d,
                                             void* stream) {
  if (!Xq_d) return -1; if (Mq < 1) return -2; if (MP % 64 || MP < Mq) return -3; if (!X_d) return -4; if (N < 1) return -5;
  if (D < 1 || D > GPK_MAXD || !inv_ls_h) return -6;
  if (!alpha_d) return -9; if (T < 1) return -10; if (!y_mean_d) return -11; if (!y_std_d) return -12;
  if (!mean_d) return -13;
  const size_t need = (size_t)gpk_mean_fused_splits(MP, N) * (size_t)MP * (size_t)T;
  if (need > (size_t)GPK_STATIC_SCRATCH_DOUBLES) return GPK_ERR_STATIC_SCRATCH_TOO_SMALL;   // nothing launched
  GpkKernelParams p;
  memset(&p, 0, sizeof(p));
  for (int d = 0; d < D; ++d) p.inv_ls[d] = inv_ls_h[d];
  p.c = c; p.noise = 0.0; p.jitter = 0.0; p.D = D;
  double* scratch = nullptr;
  GPK_CHECK(static_scratch_address(&scratch));
  GPK_CHECK(gpk_launch_predict_mean_fused(Xq_d, Mq, MP, X_d, N, D, p, alpha_d, T, y_mean_d, y_std_d, scratch, mean_d,
                                          (cudaStream_t)stream));
  return 0;
}
