// Times the phases of the Cholesky leaf in isolation (one CTA), to find what a column step really costs.
#include "../../unmanned_aerial_vehicles_b200/csrc/gpk_chol.cu"
#include <cstdio>
#include <vector>
unsigned long long g_gpk_launches = 0;
GpkProfile g_gpk_prof = {0, 0, 0.0};
cudaError_t gpk_launch_gemm_nt(const GpkGemmArgs&, cudaStream_t) { return cudaSuccess; }
cudaError_t gpk_launch_copy_block(const double*, int, double*, int, int, int, cudaStream_t) { return cudaSuccess; }
cudaError_t gpk_launch_transpose_block(const double*, int, double*, int, int, int, cudaStream_t) { return cudaSuccess; }

__global__ void __launch_bounds__(LEAF_THREADS, 1) probe_p1(const double* A, long long* cyc, int* info, int mode) {
  extern __shared__ __align__(16) unsigned char raw[];
  LeafSmem& sm = *reinterpret_cast<LeafSmem*>(raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int idx = tid; idx < 128 * 128; idx += LEAF_THREADS) {
    int i = idx >> 7, c = idx & 127;
    sm.S[i * LF_LD + c] = (c <= i) ? A[i * 128 + c] : 0.0;
  }
  if (tid < 128) sm.col[0][tid] = 0.0;
  __syncthreads();
  long long t0 = clock64();
  if (warp < 4) leaf_diag_potrf4(sm, 0, lane, warp, 0, info);
  if (mode == 0) __syncthreads();
  long long t1 = clock64();
  if (tid == 0) cyc[mode] = t1 - t0;
}
int main() {
  std::vector<double> h(128 * 128);
  for (int i = 0; i < 128; ++i) for (int j = 0; j < 128; ++j) h[i * 128 + j] = (i == j ? 2.0 : 0.0) + 0.01 + 0.001 * ((i * 7 + j * 3) % 5 == 0 && i >= j);
  double* A; long long* cyc; int* info;
  cudaMalloc(&A, 128 * 128 * 8); cudaMalloc(&cyc, 64); cudaMalloc(&info, 4); cudaMemset(info, 0, 4);
  cudaMemcpy(A, h.data(), 128 * 128 * 8, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe_p1, cudaFuncAttributeMaxDynamicSharedMemorySize, LEAF_SMEM_BYTES);
  for (int rep = 0; rep < 2; ++rep)
    for (int mode = 0; mode < 2; ++mode) {
      probe_p1<<<1, LEAF_THREADS, LEAF_SMEM_BYTES>>>(A, cyc, info, mode);
      long long c[4]; cudaMemcpy(c, cyc, 32, cudaMemcpyDeviceToHost);
      printf("mode %d: P1 one 32-block = %lld cycles (%s)\n", mode, c[mode], cudaGetErrorString(cudaGetLastError()));
    }
  {
    double *Ml, *Mu; cudaMalloc(&Ml, 128 * 128 * 8); cudaMalloc(&Mu, 128 * 128 * 8);
    leaf_configure();
    for (int rep = 0; rep < 3; ++rep) {
      cudaMemcpy(A, h.data(), 128 * 128 * 8, cudaMemcpyHostToDevice);
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0);
      gpk_leaf_potrf_inv_kernel<<<1, LEAF_THREADS, LEAF_SMEM_BYTES>>>(A, 128, Ml, Mu, 128, 0, info);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      long long c[8]; gpk_leaf_cycles(c);
      printf("full leaf on fresh A: %.1f us; load %lld P1 %lld P2 %lld P3 %lld I1 %lld I2 %lld store %lld\n", ms * 1e3, c[0], c[1], c[2], c[3], c[4], c[5], c[6]);
    }
  }
  int hi; cudaMemcpy(&hi, info, 4, cudaMemcpyDeviceToHost); printf("info %d\n", hi);
  return 0;
}
