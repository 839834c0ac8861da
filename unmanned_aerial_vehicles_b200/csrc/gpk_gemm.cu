// gpk_gemm.cu -- FP64 tensor-core (DMMA mma.sync.m8n8k4.f64) NT GEMM for sm_100a.
//   C (+)= alpha * A * B^T,  A (M x K), B (N x K), C (M x N), all row-major, dims % 128 == 0.
// Replaces the OpenBLAS dgemm/dsyrk/dtrsm/dtrmm calls hidden under scipy.linalg.cholesky /
// cho_solve / solve_triangular (sklearn/gaussian_process/_gpr.py:352,363,460,591,601,631).
// One kernel template, three CTA tiles: 64x64 (4 warps, warp tile 32x32; the default), 32x32 (small or triangular
// products that would leave SMs idle) and 128x128 (8 warps; kept for sweeps, not selected); k-tile 32, 2-stage
// cp.async ring, padded shared-memory rows (stride k-tile + 4 doubles: conflict-free 8x4 DMMA fragment reads).
#include "gpk_common.cuh"
#include <stdlib.h>

#ifndef KT
#define KT 32                 // k-tile (sweep on B200, 64-tile kernel at 8192^3: KT 32 x 2 stages 35.9 TF/s, 16 x 3 35.5, 16 x 4 35.0, 32 x 3 34.1, 8 x 4 35.7)
#endif
#define LDSM (KT + 4)         // padded smem row stride in doubles (== 4 mod 16: conflict-free 8x4 fragment reads)
#ifndef GEMM_WR
#define GEMM_WR 2
#define GEMM_WC 4
#endif
#ifndef STAGES
#define STAGES 2
#endif
#ifndef GEMM64_WR
#define GEMM64_WR 2        // warp grid of the 64-tile kernel (warp tile 32 x 32)
#define GEMM64_WC 2
#endif
#define GEMM_THREADS 256
#define GEMM_SMEM_BYTES_TS(TS) (STAGES * 2 * (TS) * LDSM * 8)
#define GPK_PROF_CAP 16384
static cudaEvent_t g_ev0[GPK_PROF_CAP], g_ev1[GPK_PROF_CAP];
static int g_shape[GPK_PROF_CAP][3];
static double g_lflops[GPK_PROF_CAP];

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// TS: CTA tile edge (128 / 64 / 32); WR x WC warps, warp tile (TS/WR) x (TS/WC)
template <int TS, int WR, int WC>
__global__ void __launch_bounds__(32 * WR * WC, (TS == 128) ? 1 : 2) gpk_gemm_nt_kernel(GpkGemmArgs g) {
  constexpr int NTHR = 32 * WR * WC, MT = TS / WR / 8, NT = TS / WC / 8, STAGE_DOUBLES = 2 * TS * LDSM;
  extern __shared__ __align__(16) double smem[];
  // grouped raster: G tile-rows per group (8 unless GPK_GEMM_RASTER says otherwise), column-major inside a group,
  // so co-resident CTAs share A and B panels in L2
  int tn, tm;
  {
    const int G = g.raster_rows;
    const int ntn = gridDim.x, ntm = gridDim.y, id = blockIdx.y * ntn + blockIdx.x;
    const int grp = id / (G * ntn), within = id - grp * G * ntn;
    const int gsz = min(G, ntm - grp * G);
    tm = grp * G + within % gsz;
    tn = within / gsz;
  }
  if (g.lower_only && tn > tm) return;
  const int m0 = tm * TS, n0 = tn * TS;
  int kb = 0, ke = g.K;
  if (g.a_tri == GPK_TRI_LOWER) ke = min(ke, m0 + TS);
  if (g.a_tri == GPK_TRI_UPPER) kb = max(kb, m0);
  if (g.b_tri == GPK_TRI_LOWER) ke = min(ke, n0 + TS);
  if (g.b_tri == GPK_TRI_UPPER) kb = max(kb, n0);
  const int nk = (ke > kb) ? (ke - kb) / KT : 0;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = (warp / WC) * (TS / WR), wn = (warp % WC) * (TS / WC);
  const int fr = lane >> 2, fk = lane & 3;

  double acc[MT][NT][2];
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  // per-thread incremental global pointers: chunk i of this thread is row (tid/CPR + i*NTHR/CPR), 16-byte column tid%CPR
  constexpr int CPR = KT / 2, NCH = TS * CPR / NTHR, RSTEP = NTHR / CPR;
  const int lr = tid / CPR, lc = (tid % CPR) * 2;
  const double* pA = g.A + (size_t)(m0 + lr) * g.lda + kb + lc;
  const double* pB = g.B + (size_t)(n0 + lr) * g.ldb + kb + lc;
  const size_t stepA = (size_t)RSTEP * g.lda, stepB = (size_t)RSTEP * g.ldb;
  const int soff = lr * LDSM + lc;
  auto prefetch = [&](int slot) {
    double* sA = smem + slot * STAGE_DOUBLES + soff;
    double* sB = sA + TS * LDSM;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      cp_async16(sA + i * RSTEP * LDSM, pA + i * stepA);
      cp_async16(sB + i * RSTEP * LDSM, pB + i * stepB);
    }
    pA += KT; pB += KT;
  };
  // prologue
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < nk) prefetch(s);
    cp_async_commit();
  }
  // The two warps that share an SM sub-partition (warp, warp + 4) issue their prefetch at different points of the
  // k-tile, so one of them always keeps the DMMA pipe fed while the other does address math / LDGSTS issue.
  const int pf_kk = ((warp >> 2) & 1) ? (KT / 2) : 0;
  for (int kt = 0; kt < nk; ++kt) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    const double* sA = smem + (kt % STAGES) * STAGE_DOUBLES;
    const double* sB = sA + TS * LDSM;
#pragma unroll
    for (int kk = 0; kk < KT; kk += 4) {
      double a[MT], b[NT];
#pragma unroll
      for (int i = 0; i < MT; ++i) a[i] = sA[(wm + i * 8 + fr) * LDSM + kk + fk];
#pragma unroll
      for (int j = 0; j < NT; ++j) b[j] = sB[(wn + j * 8 + fr) * LDSM + kk + fk];
#pragma unroll
      for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
      if (kk == pf_kk) {  // prefetch tile kt + STAGES - 1 into the slot consumed at iteration kt - 1
        const int nxt = kt + STAGES - 1;
        if (nxt < nk) prefetch(nxt % STAGES);
        cp_async_commit();
      }
    }
  }
  cp_async_wait<0>();

  if (g.epilogue == GPK_EPI_STORE) {
#pragma unroll
    for (int i = 0; i < MT; ++i) {
      int row = m0 + wm + i * 8 + fr;
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        int col = n0 + wn + j * 8 + fk * 2;
        double2* p = reinterpret_cast<double2*>(g.C + (size_t)row * g.ldc + col);
        double2 v = make_double2(g.alpha * acc[i][j][0], g.alpha * acc[i][j][1]);
        if (g.accumulate) { double2 o = *p; v.x += o.x; v.y += o.y; }
        *p = v;
        if (g.Ct) {   // transposed copy: the 8 lanes that share fk write 8 consecutive doubles (one 64-byte segment)
          g.Ct[(size_t)col * g.ldct + row] = v.x;
          g.Ct[(size_t)(col + 1) * g.ldct + row] = v.y;
        }
      }
    }
  } else if (TS == GPK_ROWSQ_TILE) {  // GPK_EPI_ROWSUMSQ: rowsq[row, tn] = sum_n (alpha*acc)^2 over this tile's columns
    __syncthreads();
    double* red = smem;  // [TS][WC]
#pragma unroll
    for (int i = 0; i < MT; ++i) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < NT; ++j) s += acc[i][j][0] * acc[i][j][0] + acc[i][j][1] * acc[i][j][1];
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      if (fk == 0) red[(wm + i * 8 + fr) * WC + (warp % WC)] = s;
    }
    __syncthreads();
    if (tid < TS) {
      double s = 0.0;
#pragma unroll
      for (int w = 0; w < WC; ++w) s += red[tid * WC + w];
      g.rowsq[(size_t)(m0 + tid) * (g.N / TS) + tn] = g.alpha * g.alpha * s;
    }
  }
}

template <int TS, int WR, int WC>
static cudaError_t launch_ts(const GpkGemmArgs& g_in, cudaStream_t st) {
  // tuning hook (tools/gemm_raster_sweep.py): GPK_GEMM_RASTER=G tile-rows per raster group.  The A panel of a group
  // (G*TS rows x K) is what has to stay in the 126 MB L2 while B streams past it once per group, so a larger G
  // cuts B's DRAM re-reads (ntm/G passes) until that panel no longer fits.  Default 8: the measured configuration.
  static const int raster_env = [] { const char* e = getenv("GPK_GEMM_RASTER"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 0; }();
  GpkGemmArgs g = g_in;
  if (g.raster_rows <= 0) g.raster_rows = raster_env ? raster_env : 8;
  static bool attr_set[GPK_MAX_DEV] = {false};
  const int dev = gpk_current_device();
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(gpk_gemm_nt_kernel<TS, WR, WC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         GEMM_SMEM_BYTES_TS(TS));
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  dim3 grid(g.N / TS, g.M / TS);
  gpk_gemm_nt_kernel<TS, WR, WC><<<grid, 32 * WR * WC, GEMM_SMEM_BYTES_TS(TS), st>>>(g);
  return cudaGetLastError();
}

cudaError_t gpk_launch_gemm_nt(const GpkGemmArgs& g, cudaStream_t st) {
  if (g.M <= 0 || g.N <= 0) return cudaSuccess;
  if ((g.M % GPK_TILE) || (g.N % GPK_TILE) || (g.K % GPK_TILE)) return cudaErrorInvalidValue;
  // tile size by available parallelism: small problems are split finer so they cover more of the 148 SMs
  const long t128 = (long)(g.M / 128) * (g.N / 128);
  // Measured on B200 (tools/gemm_tile_sweep*.py): the 64-tile kernel matches the 128-tile one on plain shapes
  // (35.0 TF/s at 5120^3) and beats it by 2-9 % on triangular / lower-only shapes at every size (finer k-range
  // trimming, less wasted work in diagonal tiles); the 32-tile kernel wins up to n = 2560 (t128 <= 400), where
  // 128-tiles leave most of the 148 SMs idle (1280^3 tri: 188 -> 95 us; 640^3: 100 -> 29 us).
  int ts = 128;
  if (g.epilogue == GPK_EPI_STORE) {
    const bool tri = g.a_tri != GPK_TRI_NONE || g.b_tri != GPK_TRI_NONE || g.lower_only;
    if (t128 <= 100 || (tri && t128 <= 400)) ts = 32;
    else ts = 64;
  } else {
    ts = GPK_ROWSQ_TILE;   // the row-sum-of-squares epilogue exists for this tile size only
  }
  {  // tuning hook (tools/gemm_tile_sweep.py): GPK_FORCE_TS=32|64|128 overrides the choice for STORE epilogues
    static const int forced = [] { const char* e = getenv("GPK_FORCE_TS"); return e ? atoi(e) : 0; }();
    if (forced && g.epilogue == GPK_EPI_STORE) ts = forced;
  }
  const bool prof = g_gpk_prof.enabled && g_gpk_prof.n < GPK_PROF_CAP;
  if (prof) {
    if (!g_ev0[g_gpk_prof.n]) { cudaEventCreate(&g_ev0[g_gpk_prof.n]); cudaEventCreate(&g_ev1[g_gpk_prof.n]); }
    cudaEventRecord(g_ev0[g_gpk_prof.n], st);
  }
  cudaError_t le = ts == 128 ? launch_ts<128, GEMM_WR, GEMM_WC>(g, st) : ts == 64 ? launch_ts<64, GEMM64_WR, GEMM64_WC>(g, st) : launch_ts<32, 2, 2>(g, st);
  if (le != cudaSuccess) return le;
  GPK_LAUNCHED();
  if (prof) {
    cudaEventRecord(g_ev1[g_gpk_prof.n], st);
    // executed tile flops: 2*ts*ts*(trimmed k range) summed over launched tiles
    double fl = 0.0;
    for (int tm = 0; tm < g.M / ts; ++tm)
      for (int tn = 0; tn < g.N / ts; ++tn) {
        if (g.lower_only && tn > tm) continue;
        int kb = 0, ke = g.K, m0 = tm * ts, n0 = tn * ts;
        if (g.a_tri == GPK_TRI_LOWER) ke = ke < m0 + ts ? ke : m0 + ts;
        if (g.a_tri == GPK_TRI_UPPER) kb = kb > m0 ? kb : m0;
        if (g.b_tri == GPK_TRI_LOWER) ke = ke < n0 + ts ? ke : n0 + ts;
        if (g.b_tri == GPK_TRI_UPPER) kb = kb > n0 ? kb : n0;
        if (ke > kb) fl += 2.0 * ts * ts * (double)(ke - kb);
      }
    g_gpk_prof.flops += fl;
    g_lflops[g_gpk_prof.n] = fl; g_shape[g_gpk_prof.n][0] = g.M; g_shape[g_gpk_prof.n][1] = g.N; g_shape[g_gpk_prof.n][2] = g.K;
    ++g_gpk_prof.n;
  }
  return cudaGetLastError();
}

// sum of the recorded GEMM launch durations (seconds); stream must be idle (caller synchronises)
double gpk_profile_collect() {
  double s = 0.0;
  for (int i = 0; i < g_gpk_prof.n; ++i) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, g_ev0[i], g_ev1[i]) == cudaSuccess) s += ms * 1e-3;
  }
  return s;
}

// ---- small data-movement helpers ---------------------------------------------------------
__global__ void gpk_copy_block_kernel(const double* __restrict__ src, int lds, double* __restrict__ dst, int ldd,
                                      int rows, int cols) {
  int c = (blockIdx.y * blockDim.x + threadIdx.x) * 2;
  int r = blockIdx.x;   // rows in grid.x: no 65535 limit
  if (c < cols)
    *reinterpret_cast<double2*>(dst + (size_t)r * ldd + c) =
        *reinterpret_cast<const double2*>(src + (size_t)r * lds + c);
}
cudaError_t gpk_launch_copy_block(const double* src, int lds, double* dst, int ldd, int rows, int cols,
                                  cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  dim3 grid(rows, (cols / 2 + 255) / 256);
  gpk_copy_block_kernel<<<grid, 256, 0, st>>>(src, lds, dst, ldd, rows, cols); GPK_LAUNCHED();
  return cudaGetLastError();
}

__global__ void gpk_transpose_block_kernel(const double* __restrict__ src, int lds, double* __restrict__ dst,
                                           int ldd, int rows, int cols) {
  __shared__ double t[32][33];
  int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) t[i][threadIdx.x] = src[(size_t)(r0 + i) * lds + c0 + threadIdx.x];
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) dst[(size_t)(c0 + i) * ldd + r0 + threadIdx.x] = t[threadIdx.x][i];
}
cudaError_t gpk_launch_transpose_block(const double* src, int lds, double* dst, int ldd, int rows, int cols,
                                       cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  dim3 grid(cols / 32, rows / 32), block(32, 8);
  gpk_transpose_block_kernel<<<grid, block, 0, st>>>(src, lds, dst, ldd, rows, cols); GPK_LAUNCHED();
  return cudaGetLastError();
}

// ---- FP64 tensor-pipe peak probe: register-resident DMMA chain, no memory traffic ----------
__global__ void __launch_bounds__(256) gpk_dmma_probe_kernel(int iters, double* sink) {
  double acc[16][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
  double a = 1.0 + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) dmma884(acc[i][0], acc[i][1], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i][0] + acc[i][1];
  if (s == 12345.678) sink[0] = s;  // keep the chain alive
}
cudaError_t gpk_launch_dmma_probe(int blocks, int iters, double* sink, cudaStream_t st) {
  gpk_dmma_probe_kernel<<<blocks, 256, 0, st>>>(iters, sink); GPK_LAUNCHED();
  return cudaGetLastError();
}

int gpk_profile_launch_info(int i, double* seconds, double* flops, int* mnk) {
  if (i < 0 || i >= g_gpk_prof.n) return -1;
  float ms = 0.f;
  cudaEventElapsedTime(&ms, g_ev0[i], g_ev1[i]);
  *seconds = ms * 1e-3; *flops = g_lflops[i]; mnk[0] = g_shape[i][0]; mnk[1] = g_shape[i][1]; mnk[2] = g_shape[i][2];
  return 0;
}

// ---- probe of the sm_90+ m16n8k16 f64 shape (does sm_100a run it at the m8n8k4 rate or better?) ------------
__device__ __forceinline__ void dmma16816(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, "
      "{%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
      : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
      : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]),
        "d"(b[2]), "d"(b[3]));
}
__global__ void __launch_bounds__(256) gpk_dmma16816_probe_kernel(int iters, double* sink) {
  double acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = 1.0 + 1e-9 * (threadIdx.x + i);
#pragma unroll
  for (int i = 0; i < 4; ++i) b[i] = 1.0 - 1e-9 * (threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma16816(acc[i], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
  if (s == 12345.678) sink[0] = s;
}
cudaError_t gpk_launch_dmma16816_probe(int blocks, int iters, double* sink, cudaStream_t st) {
  gpk_dmma16816_probe_kernel<<<blocks, 256, 0, st>>>(iters, sink);
  return cudaGetLastError();
}
