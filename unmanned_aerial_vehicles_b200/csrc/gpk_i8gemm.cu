// gpk_i8gemm.cu -- EXPERIMENTAL, not on the default path (engine.py registers scratch for it only when GPK_I8 is set).
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
// tcgen05.mma.kind::i8 from TMA-staged, swizzled shared-memory tiles.  The B slices of a stage lie back to back
// ([t][64 rows][KB bytes]) and the accumulators lie back to back in tensor memory ([g][64 columns]), so for a fixed
// A slice s the products with B slices t = 0..S-1-s -- which belong to accumulators g = s..S-1 -- are ONE wide
// instruction: M=128, N=64*(S-s) (cut at N=256), K=32.  10 instructions per k-step at S=7 instead of 28, and the
// A slice is read from shared memory once per wide instruction instead of once per product.
// No intermediate drain: |P_g| <= (g+1) * K * 4096 < 2^31 for K <= 65536.
//
// CTA = one 128 x 64 tile of C, 256 threads: warp 0 lane 0 = TMA producer, warp 1 lane 0 = MMA issuer,
// warp 2 = TMEM allocator, warps 4-7 = epilogue (tcgen05.ld, int32 -> FP64, scale, store or row-sum-of-squares).
// State of validation: compiles for sm_100a; NOT yet run on a GPU (no GPU minutes were left in round 1).
#include "gpk_common.cuh"
#include "../../include/gpk.h"
#include <cuda.h>
#include <stdlib.h>

#define I8_BM 128
#define I8_BN 64
#define I8_KPAD 64               // planes are padded to multiples of this along K
#define I8_MAXS 8
#define I8_TMEM_COLS 512
#define I8_SMEM_BUDGET (226 * 1024)   // dynamic shared memory the pipeline may use (227 KB per CTA minus statics)
#define I8_MAX_STAGES 8
// KB = int8 elements (= bytes) of K per pipeline stage = one swizzle row: 64 (SWIZZLE_64B, 2 k-steps per stage)
// or 32 (SWIZZLE_32B, 1 k-step per stage, twice as many stages in the same shared memory)

struct I8GemmParams {
  int S;                       // slices per operand (1..8)
  int stages;                  // pipeline depth (host: as many as fit in I8_SMEM_BUDGET, <= I8_MAX_STAGES)
  int n_tiles;                 // tiles along N (grid.x = m_tiles * n_tiles)
  int K;                       // padded K, multiple of I8_KPAD
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

// K-major operand tile, rows of KB bytes (KB = 64 or 32), swizzle width = KB: 8-row atoms of 8*KB bytes stacked
// along M/N.  Bit layout of the tcgen05 shared-memory matrix descriptor: start address >> 4 in [0,14), leading byte
// offset >> 4 in [16,30) (one atom along K: ignored, set to 1), stride byte offset >> 4 in [32,46) (8*KB bytes
// between 8-row groups), version 1 in [46,48), base offset 0 (tiles are aligned to the swizzle repeat), layout type
// in [61,64): 4 = SWIZZLE_64B, 6 = SWIZZLE_32B.
template <int KB>
__device__ __forceinline__ uint64_t smem_desc_kmajor(uint32_t addr) {
  uint64_t d = (uint64_t)((addr >> 4) & 0x3FFFu);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(((8u * KB) >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(KB == 64 ? 4 : 6) << 61;
  return d;
}
// instruction descriptor, kind::i8: D format S32 (2) at [4,6), A/B format signed 8-bit (1) at [7,10) / [10,13),
// both operands K-major (bits 15, 16 = 0), N >> 3 at [17,23), M >> 4 at [24,29)
__device__ __forceinline__ uint32_t idesc_i8(int n) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(I8_BM >> 4) << 24);
}

// k-range of one tile after triangular trimming, in units of KB-byte blocks: [kb_lo, kb_hi)
template <int KB>
__device__ __forceinline__ void tile_k_range(const I8GemmParams& p, int i0, int j0, int& kb_lo, int& kb_hi) {
  int lo = 0, hi = p.K;
  if (p.a_tri == GPK_TRI_LOWER) hi = min(hi, i0 + I8_BM);
  if (p.a_tri == GPK_TRI_UPPER) lo = max(lo, i0);
  if (p.b_tri == GPK_TRI_LOWER) hi = min(hi, j0 + I8_BN);
  if (p.b_tri == GPK_TRI_UPPER) lo = max(lo, j0);
  kb_lo = lo / KB;
  kb_hi = (hi + KB - 1) / KB;
  if (kb_hi < kb_lo) kb_hi = kb_lo;
}

// ---------------------------------------------------------------------------------------------- GEMM kernel
template <int KB>
__global__ void __launch_bounds__(256, 1)
gpk_i8_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const I8GemmParams p) {
  constexpr uint32_t A_SLICE = I8_BM * KB, B_SLICE = I8_BN * KB;     // bytes of one slice tile in a stage
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[I8_MAX_STAGES], empty_bar[I8_MAX_STAGES], tmem_full_bar;
  __shared__ uint32_t tmem_slot;

  const int tile_m = blockIdx.x / p.n_tiles, tile_n = blockIdx.x % p.n_tiles;
  const int i0 = tile_m * I8_BM, j0 = tile_n * I8_BN;
  if (p.lower_only && j0 >= i0 + I8_BM) return;            // whole CTA leaves before any barrier / TMEM state exists
  int kb_lo, kb_hi;
  tile_k_range<KB>(p, i0, j0, kb_lo, kb_hi);
  const int num_kb = kb_hi - kb_lo;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S, stages = p.stages;
  const uint32_t stage_bytes = (uint32_t)S * (A_SLICE + B_SLICE);
  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;   // swizzle atoms need (8*KB)-byte alignment

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(smem_u32(&full_bar[s]), 1); mbar_init(smem_u32(&empty_bar[s]), 1); }
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
    // ---- TMA producer: one 3-D box per operand per stage = all S slices of a (rows x KB bytes) tile
    int st = 0;
    uint32_t ph = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      mbar_wait(smem_u32(&empty_bar[st]), ph ^ 1u);
      const uint32_t full = smem_u32(&full_bar[st]);
      mbar_expect_tx(full, stage_bytes);
      const uint32_t a_dst = tiles + (uint32_t)st * stage_bytes;
      const uint32_t b_dst = a_dst + (uint32_t)S * A_SLICE;
      tma_load_3d(a_dst, &tma_a, full, (kb_lo + kb) * KB, i0, 0);
      tma_load_3d(b_dst, &tma_b, full, (kb_lo + kb) * KB, j0, 0);
      if (++st == stages) { st = 0; ph ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {
    // ---- MMA issuer.  Per 32-byte k-step and A slice s: B slices 0..S-1-s in one or two wide instructions
    //      (N <= 256) into accumulator columns [64 s, 64 S).  s = 0 touches every column first, so it alone
    //      carries the "overwrite" flag on the tile's first k-step.
    int st = 0;
    uint32_t ph = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      mbar_wait(smem_u32(&full_bar[st]), ph);
      tc_fence_after();
      const uint32_t a_base = tiles + (uint32_t)st * stage_bytes;
      const uint32_t b_base = a_base + (uint32_t)S * A_SLICE;
#pragma unroll
      for (int kk = 0; kk < KB / 32; ++kk) {
        for (int s = 0; s < S; ++s) {
          const uint64_t ad = smem_desc_kmajor<KB>(a_base + (uint32_t)s * A_SLICE + (uint32_t)kk * 32u);
          for (int t0 = 0; t0 < S - s; t0 += 4) {            // 4 slices of 64 rows = N 256
            const int nt = min(4, S - s - t0);
            const uint64_t bd = smem_desc_kmajor<KB>(b_base + (uint32_t)t0 * B_SLICE + (uint32_t)kk * 32u);
            tc_mma_i8(tmem + (uint32_t)(s + t0) * I8_BN, ad, bd, idesc_i8(nt * I8_BN), (kb | kk | s) != 0 ? 1u : 0u);
          }
        }
      }
      tc_commit(smem_u32(&empty_bar[st]));                  // frees the stage once these MMAs have read it
      if (++st == stages) { st = 0; ph ^= 1u; }
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
  // 4 consecutive k per thread: a warp reads 1 KB of the row and writes 128 contiguous bytes of every plane
  for (int k0 = threadIdx.x * 4; k0 < KP; k0 += 256 * 4) {
    double rr[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + j;
      rr[j] = (k >= k_lo && k < k_hi) ? row[k] * inv_mu : 0.0;
    }
    for (int s = 0; s < S; ++s) {
      const double up = (s == 0) ? 64.0 : 128.0;
      uint32_t w = 0u;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const double x = rr[j] * up;
        const double qd = rint(x);
        rr[j] = x - qd;
        w |= ((uint32_t)(uint8_t)(int8_t)(int)qd) << (8 * j);
      }
      *reinterpret_cast<uint32_t*>(out + (size_t)s * plane + k0) = w;
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

// planes [S][rows][KP] int8 -> 3-D map (k, row, slice), box (KB bytes, box_rows, S), swizzle width = KB
static int make_plane_map(CUtensorMap* map, const int8_t* planes, int rows, int KP, int S, int box_rows, int KB) {
  GpkEncodeTiledFn enc = gpk_encode_tiled();
  if (!enc) return -2000;
  cuuint64_t dims[3] = {(cuuint64_t)KP, (cuuint64_t)rows, (cuuint64_t)S};
  cuuint64_t strides[2] = {(cuuint64_t)KP, (cuuint64_t)rows * (cuuint64_t)KP};
  cuuint32_t box[3] = {(cuuint32_t)KB, (cuuint32_t)box_rows, (cuuint32_t)S};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)planes, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, KB == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2000 - (int)r;
}

// stage width along K: 64 bytes (default) or 32 (GPK_I8_KB=32: half the stage, twice the pipeline depth)
static int i8_stage_kb() {
  static int kb = 0;
  if (!kb) {
    const char* e = getenv("GPK_I8_KB");
    kb = (e && atoi(e) == 32) ? 32 : 64;
  }
  return kb;
}

extern "C" {

int gpk_i8_slice(const double* src_d, int ld, int rows, int K, int tri, int S, signed char* planes_d, int rowsP, int KP,
                 double* mu_d, void* stream) {
  if (!src_d || !planes_d || !mu_d) return -1;
  if (rows < 0 || K < 0 || ld < K || rowsP < rows || KP < K) return -2;
  if (rowsP % I8_BN || KP % I8_KPAD) return -3;
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
  if (M <= 0 || N <= 0 || KP <= 0 || M % I8_BM || N % I8_BN || KP % I8_KPAD) return -3;
  if (S < 1 || S > I8_MAXS) return -8;
  if ((long long)S * KP * 4096 >= (1ll << 31)) return -7;          // int32 accumulators would overflow
  if (!C_d && !rowsq_d) return -9;
  if (lower_only && M != N) return -15;
  static bool attr_set[GPK_MAX_DEV] = {};
  const int dev = gpk_current_device();
  const int KB = i8_stage_kb();
  const size_t stage_bytes = (size_t)S * (I8_BM + I8_BN) * KB;
  int stages = (int)((I8_SMEM_BUDGET - 1024) / stage_bytes);
  if (stages > I8_MAX_STAGES) stages = I8_MAX_STAGES;
  if (stages < 2) return -8;
  const size_t smem = (size_t)stages * stage_bytes + 1024;
  if (!attr_set[dev]) {
    GPK_CHECK(cudaFuncSetAttribute(gpk_i8_gemm_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, I8_SMEM_BUDGET));
    GPK_CHECK(cudaFuncSetAttribute(gpk_i8_gemm_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, I8_SMEM_BUDGET));
    attr_set[dev] = true;
  }
  CUtensorMap ma, mb;
  int rc = make_plane_map(&ma, (const int8_t*)Aq_d, M, KP, S, I8_BM, KB);
  if (rc) return rc;
  rc = make_plane_map(&mb, (const int8_t*)Bq_d, N, KP, S, I8_BN, KB);
  if (rc) return rc;
  I8GemmParams p;
  p.S = S; p.stages = stages; p.n_tiles = N / I8_BN; p.K = KP; p.mu_a = mu_a_d; p.mu_b = mu_b_d;
  p.C = C_d; p.ldc = ldc; p.alpha = alpha; p.accumulate = accumulate;
  p.a_tri = a_tri; p.b_tri = b_tri; p.lower_only = lower_only;
  p.epilogue = rowsq_d ? GPK_EPI_ROWSUMSQ : GPK_EPI_STORE;
  p.rowsq = rowsq_d; p.rowsq_ld = N / I8_BN;
  const int grid = (M / I8_BM) * (N / I8_BN);
  if (KB == 64) gpk_i8_gemm_kernel<64><<<grid, 256, smem, (cudaStream_t)stream>>>(ma, mb, p);
  else          gpk_i8_gemm_kernel<32><<<grid, 256, smem, (cudaStream_t)stream>>>(ma, mb, p);
  GPK_LAUNCHED();
  GPK_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------- opt-in use on the path
// Nothing below runs unless gpk_i8_configure() has registered scratch on the current device (engine.py does that
// only when GPK_I8 is set).  mode bit 1: K^-1 = L^-T L^-1 (LAUUM) of gpk_lml_eval; bit 2: predict's Ks L^-T;
// bit 4: right-most-path inverse products inside gpk_lml_eval's factorisation.
static GpkI8State g_i8[GPK_MAX_DEV];

static inline size_t i8_align(size_t b) { return (b + 255) & ~(size_t)255; }
static inline size_t i8_square_bytes(int NP, int S) { return i8_align((size_t)S * NP * NP) + i8_align(sizeof(double) * (size_t)NP); }
static inline size_t i8_rows_bytes(int MP, int NP, int S) { return i8_align((size_t)S * MP * NP) + i8_align(sizeof(double) * (size_t)MP); }

size_t gpk_i8_workspace_bytes(int NP, int MPmax, int S) {
  if (NP < I8_BM || NP % I8_BM || MPmax < 0 || MPmax % I8_BM || S < 1 || S > I8_MAXS) return 0;
  return 2 * i8_square_bytes(NP, S) + i8_rows_bytes(MPmax, NP, S);
}

int gpk_i8_configure(void* ws_d, size_t bytes, int NP, int MPmax, int S, int mode) {
  GpkI8State& g = g_i8[gpk_current_device()];
  g = GpkI8State{};
  if (!ws_d || mode == 0) return 0;                              // disabled
  const size_t need = gpk_i8_workspace_bytes(NP, MPmax, S);
  if (need == 0) return -3;
  if (bytes < need) return -2;
  if ((long long)S * NP * 4096 >= (1ll << 31)) return -7;
  g.ws = (char*)ws_d; g.bytes = bytes; g.NP = NP; g.MPmax = MPmax; g.S = S; g.mode = mode;
  return 0;
}

int gpk_i8_mode() { return g_i8[gpk_current_device()].mode; }

void gpk_i8_invalidate() { g_i8[gpk_current_device()].prepared = nullptr; }

int gpk_i8_prepare_predict(const double* Mlow_d, int NP, void* stream) {
  GpkI8State& g = g_i8[gpk_current_device()];
  g.prepared = nullptr;
  if (!(g.mode & 2)) return 0;
  if (!Mlow_d) return -1;
  if (NP != g.NP) return -2;
  signed char* planes = (signed char*)g.ws;
  double* mu = (double*)(g.ws + i8_align((size_t)g.S * NP * NP));
  int rc = gpk_i8_slice(Mlow_d, NP, NP, NP, GPK_TRI_LOWER, g.S, planes, NP, NP, mu, stream);
  if (rc) return rc;
  g.prepared = Mlow_d;
  return 0;
}

}  // extern "C"

// K^-1 (lower tiles) = Mup Mup^T from the int8 planes of Mup = L^-T.  Returns 1 when it did the work, 0 when the
// int8 path is not configured for this shape (caller falls through to the DMMA kernel), < 0 on error.
int gpk_i8_try_lauum(const double* Mup, int ldm, double* Kinv, int ldk, int NP, cudaStream_t st) {
  GpkI8State& g = g_i8[gpk_current_device()];
  if (!(g.mode & 1) || NP != g.NP || ldm != NP) return 0;
  char* base = g.ws + i8_square_bytes(NP, g.S);                  // second square region (the first holds predict's L^-1)
  signed char* planes = (signed char*)base;
  double* mu = (double*)(base + i8_align((size_t)g.S * NP * NP));
  int rc = gpk_i8_slice(Mup, ldm, NP, NP, GPK_TRI_UPPER, g.S, planes, NP, NP, mu, st);
  if (rc) return rc;
  rc = gpk_i8_gemm_nt(planes, mu, NP, planes, mu, NP, NP, g.S, Kinv, ldk, 1.0, 0, GPK_TRI_UPPER, GPK_TRI_UPPER, 1, nullptr, st);
  return rc ? rc : 1;
}

// Right-most-path inverse products of the recursive factorisation (mode bit 4; gpk_chol.cu): C (M x N) = alpha A B^T
// from FP64 blocks with leading dimensions, planes carved from the second square region at region_off (one
// region per recursion depth, so the product a side stream still runs never shares planes with a deeper level;
// LAUUM re-uses the region afterwards, stream-ordered behind all of them).  Same return convention.
static int i8_tail_min() {
  static int v = 0;
  if (!v) { const char* e = getenv("GPK_I8_TAIL_MIN"); v = (e && atoi(e) >= I8_BM) ? atoi(e) : 1024; }
  return v;
}

size_t gpk_i8_pair_bytes(int M, int N, int K) {
  const GpkI8State& g = g_i8[gpk_current_device()];
  if (!(g.mode & 4)) return 0;
  return i8_align((size_t)g.S * M * K) + i8_align(sizeof(double) * (size_t)M) + i8_align((size_t)g.S * N * K) +
         i8_align(sizeof(double) * (size_t)N);
}

int gpk_i8_try_gemm(const double* A, int lda, int a_tri, const double* B, int ldb, int b_tri, int M, int N, int K,
                    double alpha, double* C, int ldc, size_t region_off, cudaStream_t st) {
  GpkI8State& g = g_i8[gpk_current_device()];
  if (!(g.mode & 4) || M % I8_BM || N % I8_BN || K % I8_KPAD) return 0;
  if (M < i8_tail_min() || N < i8_tail_min() || K < i8_tail_min()) return 0;
  const size_t need = gpk_i8_pair_bytes(M, N, K);
  if (region_off + need > i8_square_bytes(g.NP, g.S)) return 0;
  char* base = g.ws + i8_square_bytes(g.NP, g.S) + region_off;
  signed char* pa = (signed char*)base;
  double* mua = (double*)(base + i8_align((size_t)g.S * M * K));
  char* bb = (char*)mua + i8_align(sizeof(double) * (size_t)M);
  signed char* pb = (signed char*)bb;
  double* mub = (double*)(bb + i8_align((size_t)g.S * N * K));
  int rc = gpk_i8_slice(A, lda, M, K, a_tri, g.S, pa, M, K, mua, st);
  if (rc) return rc;
  rc = gpk_i8_slice(B, ldb, N, K, b_tri, g.S, pb, N, K, mub, st);
  if (rc) return rc;
  rc = gpk_i8_gemm_nt(pa, mua, M, pb, mub, N, K, g.S, C, ldc, alpha, 0, a_tri, b_tri, 0, nullptr, st);
  return rc ? rc : 1;
}

// rowsq (MP x NP/64) = per-tile row sums of squares of Ks L^-T through the prepared planes of L^-1.  Same return
// convention as gpk_i8_try_lauum.
int gpk_i8_try_predict_rowsq(const double* Ks, int MP, const double* Mlow, int NP, double* rowsq, cudaStream_t st) {
  GpkI8State& g = g_i8[gpk_current_device()];
  if (!(g.mode & 2) || g.prepared != Mlow || NP != g.NP || MP > g.MPmax || MP % I8_BM) return 0;
  const signed char* lp = (const signed char*)g.ws;
  const double* lmu = (const double*)(g.ws + i8_align((size_t)g.S * NP * NP));
  char* base = g.ws + 2 * i8_square_bytes(NP, g.S);
  signed char* kp = (signed char*)base;
  double* kmu = (double*)(base + i8_align((size_t)g.S * MP * NP));
  int rc = gpk_i8_slice(Ks, NP, MP, NP, GPK_TRI_NONE, g.S, kp, MP, NP, kmu, st);
  if (rc) return rc;
  rc = gpk_i8_gemm_nt(kp, kmu, MP, lp, lmu, NP, NP, g.S, nullptr, 0, 1.0, 0, GPK_TRI_NONE, GPK_TRI_LOWER, 0, rowsq, st);
  return rc ? rc : 1;
}
