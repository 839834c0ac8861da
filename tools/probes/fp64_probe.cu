// FP64 pipe probe: dependent-chain latency and multi-warp throughput of DFMA / DMUL / rsqrt on one SM.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void lat_dfma(double* out, long long* cyc, int n) {
  double a = out[0], b = out[1], c = out[2];
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) { a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); }
  long long t1 = clock64();
  out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void lat_rsqrt(double* out, long long* cyc, int n) {
  double a = out[0] + 2.0;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) { a = rsqrt(a) + 1.5; a = rsqrt(a) + 1.5; }
  long long t1 = clock64();
  out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int ILP>
__global__ void thr_dfma(double* out, long long* cyc, int n) {
  double a[ILP]; double b = out[1], c = out[2];
  for (int k = 0; k < ILP; ++k) a[k] = out[0] + k + threadIdx.x;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) a[k] = fma(a[k], b, c);
  }
  __syncthreads();
  long long t1 = clock64();
  double s = 0; for (int k = 0; k < ILP; ++k) s += a[k];
  out[4 + threadIdx.x] = s; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void lat_shfl(double* out, long long* cyc, int n) {
  double a = out[0] + threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) { a = __shfl_sync(0xffffffffu, a, (threadIdx.x + 1) & 31); a = __shfl_sync(0xffffffffu, a, (threadIdx.x + 3) & 31); }
  long long t1 = clock64();
  out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 8 * 2048); cudaMalloc(&cyc, 64);
  double h[3] = {1.0000001, 0.9999999, 1e-9}; cudaMemcpy(out, h, 24, cudaMemcpyHostToDevice);
  long long c; int n = 4096;
  for (int rep = 0; rep < 2; ++rep) {
    lat_dfma<<<1, 32>>>(out, cyc, n); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("DFMA dependent latency: %.1f cycles\n", (double)c / (4.0 * n));
    lat_rsqrt<<<1, 32>>>(out, cyc, n); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("rsqrt(double)+DADD dependent latency: %.1f cycles\n", (double)c / (2.0 * n));
    lat_shfl<<<1, 32>>>(out, cyc, n); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("shfl(double) dependent latency: %.1f cycles\n", (double)c / (2.0 * n));
    int warps[] = {1, 4, 8, 16, 32};
    for (int w : warps) {
      thr_dfma<8><<<1, 32 * w>>>(out, cyc, n); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
      printf("DFMA throughput %2d warps x ILP8: %.2f DFMA lanes/clk/SM\n", w, 8.0 * n * 32 * w / (double)c);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
