// gpk_i8gemm.cu -- EXPERIMENTAL, not on the default path (nothing in engine.py calls it yet).
//
// FP64-accurate NT GEMM on the tcgen05 INT8 pipe (Ozaki splitting): C = alpha * A * B^T with A (M x K), B (N x K)
// given as S int8 "slice planes" per operand plus one power-of-two scale per row (gpk_i8_slice produces them).
// DESIGN.md section 8 has the accuracy study (tools/ozaki_study.py) that says where this may replace the DMMA
// kernel: the right-most-path inverse blocks, LAUUM and predict's Ks * L^-T, at 7 slices (28 slice products).
//
//   a_ik = mu_i * sum_s q^A_s[i,k] * 2^-(6+7s),  |q| <= 64      (slice 0 keeps 6 bits + sign, later ones 7 + sign)
//   C_ij = alpha * mu_i * nu_j * sum_{g=0}^{S-1} 2^-(12+7g) * P_g[i,j],   P_g = sum_{s+t=g} q^A_s (q^B_t)^T
//
// P_g is an exact integer: one int32 TMEM accumulator per g (64 columns each, S <= 8 -> <= 512 columns), fed by
// tcgen05.mma.kind::i8 (M=128, N=64, K=32 per instruction) from TMA-staged, 64-byte-swizzled shared-memory tiles.
// No intermediate drain: |P_g| <= (g+1) * K * 4096 < 2^31 for K <= 65536.
//
// CTA = one 128 x 64 tile of C, 256 threads: warp 0 lane 0 = TMA producer, warp 1 lane 0 = MMA issuer,
// warp 2 = TMEM allocator, warps 4-7 = epilogue (tcgen05.ld, int32 -> FP64, scale, store or row-sum-of-squares).
// State of validation: compiles for sm_100a; NOT yet run on a GPU (no GPU minutes were left in round 1).
#include "gpk_common.cuh"
#include "../../include/gpk.h"
#include <cuda.h>

#define I8_BM 128
#define I8_BN 64
#define I8_KB 64                 // int8 elements (= bytes) of K per pipeline stage: one 64-byte swizzle row
#define I8_STAGES 2
#define I8_MAXS 8
#define I8_TMEM_COLS 512
#define I8_A_SLICE_BYTES (I8_BM * I8_KB)     // 8192
#define I8_B_SLICE_BYTES (I8_BN * I8_KB)     // 4096

struct I8GemmParams {
  int S;                       // slices per operand (1..8)
  int n_tiles;                 // tiles along N (grid.x = m_tiles * n_tiles)
  int K;                       // padded K, multiple of I8_KB
  const double* mu_a;          // (M) power-of-two row scales of A
  const double* mu_b;          // (N) power-of-two row scales of B
  double* C; int ldc;          // GPK_EPI_STORE
  double alpha; int accumulate;
  int a_tri, b_tri, lower_only, epilogue;
  double* rowsq; int rowsq_ld; // GPK_EPI_ROWSUMSQ: (M x N/64) partial sums of squares
};

// ---------------------------------------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"((uint64_t)map), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, int8 x int8 -> int32, M=128, N=64, K=32
__device__ __forceinline__ void tc_mma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// 32 lanes x 16 consecutive 32-bit columns: thread l of the warp gets row (lane base + l)
__device__ __forceinline__ void tc_ld16(uint32_t taddr, int (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major operand tile, rows of 64 bytes, 64-byte swizzle: 8-row atoms of 512 bytes stacked along M/N.
// Bit layout of the tcgen05 shared-memory matrix descriptor: start address >> 4 in [0,14), leading byte offset >> 4
// in [16,30) (one atom along K: ignored, set to 1), stride byte offset >> 4 in [32,46) (512 B between 8-row groups),
// version 1 in [46,48), base offset 0 (tiles are 512-byte aligned), layout type in [61,64): 4 = SWIZZLE_64B.
__device__ __forceinline__ uint64_t smem_desc_k64(uint32_t addr) {
  uint64_t d = (uint64_t)((addr >> 4) & 0x3FFFu);
  d |= (uint64_t)1 << 16;                                  // leading byte offset: not used by swizzled K-major layouts
  d |= (uint64_t)((512u >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;
  return d;
}
// instruction descriptor, kind::i8: D format S32 (2) at [4,6), A/B format signed 8-bit (1) at [7,10) / [10,13),
// both operands K-major (bits 15, 16 = 0), N >> 3 at [17,23), M >> 4 at [24,29)
__device__ __forceinline__ uint32_t idesc_i8_128x64() {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(I8_BN >> 3) << 17) | ((uint32_t)(I8_BM >> 4) << 24);
}

// k-range of one tile after triangular trimming, in units of I8_KB blocks: [kb_lo, kb_hi)
__device__ __forceinline__ void tile_k_range(const I8GemmParams& p, int i0, int j0, int& kb_lo, int& kb_hi) {
  int lo = 0, hi = p.K;
  if (p.a_tri == GPK_TRI_LOWER) hi = min(hi, i0 + I8_BM);
  if (p.a_tri == GPK_TRI_UPPER) lo = max(lo, i0);
  if (p.b_tri == GPK_TRI_LOWER) hi = min(hi, j0 + I8_BN);
  if (p.b_tri == GPK_TRI_UPPER) lo = max(lo, j0);
  kb_lo = lo / I8_KB;
  kb_hi = (hi + I8_KB - 1) / I8_KB;
  if (kb_hi < kb_lo) kb_hi = kb_lo;
}

// ---------------------------------------------------------------------------------------------- GEMM kernel
__global__ void __launch_bounds__(256, 1)
gpk_i8_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const I8GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[I8_STAGES], empty_bar[I8_STAGES], tmem_full_bar;
  __shared__ uint32_t tmem_slot;

  const int tile_m = blockIdx.x / p.n_tiles, tile_n = blockIdx.x % p.n_tiles;
  const int i0 = tile_m * I8_BM, j0 = tile_n * I8_BN;
  if (p.lower_only && j0 >= i0 + I8_BM) return;            // whole CTA leaves before any barrier / TMEM state exists
  int kb_lo, kb_hi;
  tile_k_range(p, i0, j0, kb_lo, kb_hi);
  const int num_kb = kb_hi - kb_lo;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S;
  const uint32_t stage_bytes = (uint32_t)S * (I8_A_SLICE_BYTES + I8_B_SLICE_BYTES);
  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;   // swizzle atoms need 512-byte alignment

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < I8_STAGES; ++s) { mbar_init(smem_u32(&full_bar[s]), 1); mbar_init(smem_u32(&empty_bar[s]), 1); }
    mbar_init(smem_u32(&tmem_full_bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {                                          // one warp allocates and later frees the tensor memory
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(I8_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == 0 && lane == 0) {
    // ---- TMA producer: one 3-D box per operand per stage = all S slices of a (rows x 64 B) tile
    for (int kb = 0; kb < num_kb; ++kb) {
      const int st = kb % I8_STAGES;
      const uint32_t ph = (uint32_t)(kb / I8_STAGES) & 1u;
      mbar_wait(smem_u32(&empty_bar[st]), ph ^ 1u);
      const uint32_t full = smem_u32(&full_bar[st]);
      mbar_expect_tx(full, stage_bytes);
      const uint32_t a_dst = tiles + (uint32_t)st * stage_bytes;
      const uint32_t b_dst = a_dst + (uint32_t)S * I8_A_SLICE_BYTES;
      tma_load_3d(a_dst, &tma_a, full, (kb_lo + kb) * I8_KB, i0, 0);
      tma_load_3d(b_dst, &tma_b, full, (kb_lo + kb) * I8_KB, j0, 0);
    }
  } else if (warp == 1 && lane == 0) {
    // ---- MMA issuer: per 32-byte k-step, S(S+1)/2 slice products, product (s, t) into accumulator g = s + t
    const uint32_t idesc = idesc_i8_128x64();
    for (int kb = 0; kb < num_kb; ++kb) {
      const int st = kb % I8_STAGES;
      const uint32_t ph = (uint32_t)(kb / I8_STAGES) & 1u;
      mbar_wait(smem_u32(&full_bar[st]), ph);
      tc_fence_after();
      const uint32_t a_base = tiles + (uint32_t)st * stage_bytes;
      const uint32_t b_base = a_base + (uint32_t)S * I8_A_SLICE_BYTES;
#pragma unroll
      for (int kk = 0; kk < I8_KB / 32; ++kk) {
        for (int g = 0; g < S; ++g) {
          for (int s = 0; s <= g; ++s) {
            const int t = g - s;
            const uint64_t ad = smem_desc_k64(a_base + (uint32_t)s * I8_A_SLICE_BYTES + (uint32_t)kk * 32u);
            const uint64_t bd = smem_desc_k64(b_base + (uint32_t)t * I8_B_SLICE_BYTES + (uint32_t)kk * 32u);
            tc_mma_i8(tmem + (uint32_t)g * I8_BN, ad, bd, idesc, (kb | kk | s) != 0 ? 1u : 0u);
          }
        }
      }
      tc_commit(smem_u32(&empty_bar[st]));                  // frees the stage once these MMAs have read it
    }
    if (num_kb > 0) tc_commit(smem_u32(&tmem_full_bar));
  } else if (warp >= 4) {
    // ---- epilogue: thread = one row of the tile (TMEM lane), 64 columns in 4 chunks of 16
    const int q = warp - 4;
    const int row = q * 32 + lane;
    const size_t grow = (size_t)(i0 + row);
    if (num_kb > 0) {
      mbar_wait(smem_u32(&tmem_full_bar), 0);
      tc_fence_after();
    }
    const double ra = p.alpha * p.mu_a[grow];
    double ssq = 0.0;
    for (int c0 = 0; c0 < I8_BN; c0 += 16) {
      double acc[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[j] = 0.0;
      if (num_kb > 0) {
        for (int g = S - 1; g >= 0; --g) {                  // smallest scale first
          int v[16];
          tc_ld16(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(g * I8_BN + c0), v);
          const double sc = __longlong_as_double((long long)(1023 - (12 + 7 * g)) << 52);   // 2^-(12+7g)
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = fma((double)v[j], sc, acc[j]);
        }
      }
      if (p.epilogue == GPK_EPI_ROWSUMSQ) {
#pragma unroll
        for (int j = 0; j < 16; ++j) { const double x = acc[j] * ra * p.mu_b[j0 + c0 + j]; ssq = fma(x, x, ssq); }
      } else {
        double* crow = p.C + grow * (size_t)p.ldc + (size_t)(j0 + c0);
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          double2 o;
          o.x = acc[j] * ra * p.mu_b[j0 + c0 + j];
          o.y = acc[j + 1] * ra * p.mu_b[j0 + c0 + j + 1];
          if (p.accumulate) { const double2 old = *reinterpret_cast<const double2*>(crow + j); o.x += old.x; o.y += old.y; }
          *reinterpret_cast<double2*>(crow + j) = o;
        }
      }
    }
    if (p.epilogue == GPK_EPI_ROWSUMSQ) p.rowsq[grow * (size_t)p.rowsq_ld + (size_t)tile_n] = ssq;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(I8_TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------- slicing
// One CTA per row: row maximum -> power-of-two scale mu (|a| / mu < 1), then S rounds of
// r *= 2^b; q = rint(r); r -= q  (b = 6 for the first slice, 7 afterwards; every step exact in FP64).
// planes: [S][rowsP][KP] int8, K contiguous; rows >= rows and columns >= K are zero; mu = 1 for empty rows.
// tri: only the stated triangle of src is read (row r: k <= r lower, k >= r upper), the rest is taken as zero.
__global__ void __launch_bounds__(256)
gpk_i8_slice_kernel(const double* __restrict__ src, int ld, int rows, int K, int tri, int S,
                    int8_t* __restrict__ planes, int rowsP, int KP, double* __restrict__ mu_out) {
  __shared__ double red[8];
  __shared__ double inv_mu_s;
  const int r = blockIdx.x;
  int k_lo = 0, k_hi = (r < rows) ? K : 0;
  if (tri == GPK_TRI_LOWER) k_hi = min(k_hi, r + 1);
  if (tri == GPK_TRI_UPPER) k_lo = min(r, k_hi);
  const double* row = src + (size_t)r * ld;

  double amax = 0.0;
  for (int k = k_lo + threadIdx.x; k < k_hi; k += 256) amax = fmax(amax, fabs(row[k]));
  for (int o = 16; o; o >>= 1) amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
  __syncthreads();
  if (threadIdx.x == 0) {
    double m = red[0];
    for (int w = 1; w < 8; ++w) m = fmax(m, red[w]);
    int e = 0;
    if (m > 0.0) frexp(m, &e);                       // m = f * 2^e, f in [0.5, 1)  ->  m < 2^e
    const double mu = (m > 0.0) ? ldexp(1.0, e) : 1.0;
    mu_out[r] = mu;
    inv_mu_s = (m > 0.0) ? ldexp(1.0, -e) : 1.0;
  }
  __syncthreads();
  const double inv_mu = inv_mu_s;

  const size_t plane = (size_t)rowsP * KP;
  int8_t* out = planes + (size_t)r * KP;
  for (int k0 = threadIdx.x * 16; k0 < KP; k0 += 256 * 16) {     // 16 consecutive k per thread: one 16-byte store per plane
    double rr[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int k = k0 + j;
      rr[j] = (k >= k_lo && k < k_hi) ? row[k] * inv_mu : 0.0;
    }
    for (int s = 0; s < S; ++s) {
      const double up = (s == 0) ? 64.0 : 128.0;
      uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const double x = rr[j] * up;
        const double qd = rint(x);
        rr[j] = x - qd;
        w[j >> 2] |= ((uint32_t)(uint8_t)(int8_t)(int)qd) << (8 * (j & 3));
      }
      *reinterpret_cast<uint4*>(out + (size_t)s * plane + k0) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

// ---------------------------------------------------------------------------------------------- host side
typedef CUresult (*GpkEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static GpkEncodeTiledFn gpk_encode_tiled() {
  static GpkEncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = (GpkEncodeTiledFn)p;
  }
  return fn;
}

// planes [S][rows][KP] int8 -> 3-D map (k, row, slice), box (64 B, box_rows, S), 64-byte swizzle
static int make_plane_map(CUtensorMap* map, const int8_t* planes, int rows, int KP, int S, int box_rows) {
  GpkEncodeTiledFn enc = gpk_encode_tiled();
  if (!enc) return -2000;
  cuuint64_t dims[3] = {(cuuint64_t)KP, (cuuint64_t)rows, (cuuint64_t)S};
  cuuint64_t strides[2] = {(cuuint64_t)KP, (cuuint64_t)rows * (cuuint64_t)KP};
  cuuint32_t box[3] = {(cuuint32_t)I8_KB, (cuuint32_t)box_rows, (cuuint32_t)S};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)planes, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2000 - (int)r;
}

extern "C" {

int gpk_i8_slice(const double* src_d, int ld, int rows, int K, int tri, int S, signed char* planes_d, int rowsP, int KP,
                 double* mu_d, void* stream) {
  if (!src_d || !planes_d || !mu_d) return -1;
  if (rows < 0 || K < 0 || ld < K || rowsP < rows || KP < K) return -2;
  if (rowsP % I8_BN || KP % I8_KB) return -3;
  if (S < 1 || S > I8_MAXS) return -6;
  if (rowsP == 0) return 0;
  gpk_i8_slice_kernel<<<rowsP, 256, 0, (cudaStream_t)stream>>>(src_d, ld, rows, K, tri, S, (int8_t*)planes_d, rowsP, KP, mu_d);
  GPK_LAUNCHED();
  GPK_CHECK(cudaGetLastError());
  return 0;
}

int gpk_i8_gemm_nt(const signed char* Aq_d, const double* mu_a_d, int M, const signed char* Bq_d, const double* mu_b_d,
                   int N, int KP, int S, double* C_d, int ldc, double alpha, int accumulate, int a_tri, int b_tri,
                   int lower_only, double* rowsq_d, void* stream) {
  if (!Aq_d || !mu_a_d || !Bq_d || !mu_b_d) return -1;
  if (M <= 0 || N <= 0 || KP <= 0 || M % I8_BM || N % I8_BN || KP % I8_KB) return -3;
  if (S < 1 || S > I8_MAXS) return -8;
  if ((long long)S * KP * 4096 >= (1ll << 31)) return -7;          // int32 accumulators would overflow
  if (!C_d && !rowsq_d) return -9;
  if (lower_only && M != N) return -15;
  static bool attr_set[GPK_MAX_DEV] = {};
  const int dev = gpk_current_device();
  const size_t smem = (size_t)I8_STAGES * S * (I8_A_SLICE_BYTES + I8_B_SLICE_BYTES) + 1024;
  if (!attr_set[dev]) {
    GPK_CHECK(cudaFuncSetAttribute(gpk_i8_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   I8_STAGES * I8_MAXS * (I8_A_SLICE_BYTES + I8_B_SLICE_BYTES) + 1024));
    attr_set[dev] = true;
  }
  CUtensorMap ma, mb;
  int rc = make_plane_map(&ma, (const int8_t*)Aq_d, M, KP, S, I8_BM);
  if (rc) return rc;
  rc = make_plane_map(&mb, (const int8_t*)Bq_d, N, KP, S, I8_BN);
  if (rc) return rc;
  I8GemmParams p;
  p.S = S; p.n_tiles = N / I8_BN; p.K = KP; p.mu_a = mu_a_d; p.mu_b = mu_b_d;
  p.C = C_d; p.ldc = ldc; p.alpha = alpha; p.accumulate = accumulate;
  p.a_tri = a_tri; p.b_tri = b_tri; p.lower_only = lower_only;
  p.epilogue = rowsq_d ? GPK_EPI_ROWSUMSQ : GPK_EPI_STORE;
  p.rowsq = rowsq_d; p.rowsq_ld = N / I8_BN;
  gpk_i8_gemm_kernel<<<(M / I8_BM) * (N / I8_BN), 256, smem, (cudaStream_t)stream>>>(ma, mb, p);
  GPK_LAUNCHED();
  GPK_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
