// gpk_chol.cu -- recursive blocked FP64 Cholesky fused with the triangular inverse.
//
// Replaces scipy.linalg.cholesky (dpotrf; sklearn _gpr.py:352,591) and the N-right-hand-side
// cho_solve(L, I) sklearn uses to form K^-1 (_gpr.py:631, 2N^3 flops) by
//   potrf + trtri (2N^3/3) followed by lauum (N^3/3), all as large-K NT GEMMs on the FP64 tensor pipe.
//
// Buffers (NP x NP, NP % 128 == 0):
//   A    : in  = lower triangle of K (+ identity padding), out = lower Cholesky factor L; the strictly upper
//          128-blocks are scratch (each level parks its Wt product there)
//   Mlow : out = L^-1 (lower; diagonal 128-blocks carry explicit zeros above the diagonal)
//   Mup  : out = L^-T (upper; = Mlow transposed)
// Recursion on [off, off+n):  n1 | n2 split at a multiple of 128
//   rec(11);  L21 = A21 * M11^T;  A22 -= L21 L21^T;  rec(22);  M21 = -M22 * L21 * M11  (details at chol_rec).
#include "gpk_common.cuh"

// ---------------------------------------------------------------------------------------------------------
// Leaf: one CTA factors a 128x128 diagonal block and inverts the factor (L, L^-1, L^-T out).
// v4: blocked at 32 inside shared memory (the block never leaves the SM between load and store):
//   for kb = 0..3:  P1  warp 0 factors the 32x32 diagonal block in registers (lane = row; the column of
//                       multipliers is broadcast with shuffles, no barrier inside the 32 column steps)
//                   P2  one thread per row below: forward substitution x L_kk^T = a  (panel, in place)
//                   P3  all threads: trailing update A22 -= P P^T on 16x16-cyclic register tiles
//   I1  warps 0..3 invert the four diagonal blocks (lane = column, axpy form)
//   I2  off-diagonal blocks of L^-1 by block forward substitution, level by level:
//       M_ij = -M_ii (sum_{j<=k<i} L_ik M_kj)
// Shared memory: S (128 x 129) holds A -> L in its lower triangle and the off-diagonal blocks of L^-1
// TRANSPOSED in its upper triangle; diagonal blocks of L^-1 live in Mkk.  Row stride 129 makes both row and
// column walks bank-conflict free for 64-bit accesses.
// info: first failing global pivot index + 1 (LAPACK convention), 0 if none.
#define LEAF_THREADS 256
#define LF_LD 129
struct LeafSmem {
  double S[GPK_TILE * LF_LD];
  double Mkk[4][32 * 33];
  double LT[32 * 32];      // LT[c][c2] = L_kk[c2][c] (transposed copy of the current diagonal block)
  double W[3][32 * 33];
  double dinv[GPK_TILE];   // 1 / L_ii
  double col[2][64];       // P1: raw pivot column, double-buffered (reads of dead columns run up to index 63)
};
#define LEAF_SMEM_BYTES ((int)sizeof(LeafSmem))

__device__ __forceinline__ void leaf_record(int* info, int want) {
  int old = atomicCAS(info, 0, want);
  if (old != 0 && old > want) atomicMin(info, want);
}

// P1: potrf of the 32x32 block at (k0,k0) by FOUR warps.  In every warp lane i owns row i; warp w keeps the
// columns c = w (mod 4) in 8 registers.  Step j: the owner warp (j mod 4) publishes the raw column through a
// double-buffered shared vector, one 128-thread named barrier, then every warp scales by 1/sqrt(pivot) (1-ulp rsqrt,
// no division) and updates its own live columns with r_c -= (a_i a_c)/d.  The code is kept SMALL (rolled over groups
// of four steps): a fully unrolled sweep runs at instruction-fetch speed.  The owner rotates its registers as it
// updates (the FMA writes r[t-1]), so the live column is always r[0] and the indexing stays static.
// One warp did this in 356 cycles per column (8 k cycles of FMA issue + the serial chain); four take the chain only.
__device__ __forceinline__ void leaf_diag_potrf4(LeafSmem& sm, int k0, int lane, int w, int off, int* info) {
  double r[8];
  const double* row = sm.S + (k0 + lane) * LF_LD + k0 + w;
#pragma unroll
  for (int t = 0; t < 8; ++t) r[t] = row[4 * t];
#pragma unroll 1
  for (int jj = 0; jj < 8; ++jj) {
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const int j = 4 * jj + s;
      double* buf = sm.col[j & 1];
      if (w == s) buf[lane] = r[0];
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const double raw = buf[lane];
      double d = buf[j];
      const double* bc = buf + 4 * (jj + (w < s ? 1 : 0)) + w;   // rows that match this warp's live columns
      double bl[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) bl[t] = bc[4 * t];
      if (!(d > 0.0)) {  // non-positive or NaN pivot: record and poison
        if (w == 0 && lane == 0) leaf_record(info, off + k0 + j + 1);
        d = __longlong_as_double(0x7ff8000000000000LL);
      }
      const double dinv = rsqrt(d);
      const double ljj = d * dinv;
      const double li = (lane > j) ? raw * dinv : (lane == j ? ljj : 0.0);   // finished rows contribute nothing
      const double sc = -li * dinv;
      if (w == s) {
        sm.S[(k0 + lane) * LF_LD + k0 + j] = li;
        sm.LT[j * 32 + lane] = li;
        if (lane == j) sm.dinv[k0 + j] = dinv;
#pragma unroll
        for (int t = 1; t < 8; ++t) r[t - 1] = fma(sc, bl[t], r[t]);
      } else {
#pragma unroll
        for (int t = 0; t < 8; ++t) r[t] = fma(sc, bl[t], r[t]);
      }
    }
  }
}

// P2: rows below the diagonal block, one thread per row: x L_kk^T = a by forward substitution (axpy form)
__device__ __forceinline__ void leaf_panel_row(LeafSmem& sm, int k0, int i) {
  double a[32];
  double* row = sm.S + i * LF_LD + k0;
#pragma unroll
  for (int c = 0; c < 32; ++c) a[c] = row[c];
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    const double x = a[c] * sm.dinv[k0 + c];
    a[c] = x;
#pragma unroll
    for (int c2 = c + 1; c2 < 32; ++c2) a[c2] = fma(-x, sm.LT[c * 32 + c2], a[c2]);
  }
#pragma unroll
  for (int c = 0; c < 32; ++c) row[c] = a[c];
}

// P3: A[i][c] -= sum_k P[i][k] P[c][k] over the trailing rows/cols (NRB 32-blocks), lower 16x16 sub-tiles only.
// Thread (ty,tx) of a 16x16 grid owns rows base+ty+16a, cols base+tx+16b.
template <int NRB>
__device__ __forceinline__ void leaf_trailing(LeafSmem& sm, int k0, int base, int tx, int ty) {
  constexpr int NA = 2 * NRB;
  double acc[NA][NA];
#pragma unroll
  for (int a = 0; a < NA; ++a)
#pragma unroll
    for (int b = 0; b < NA; ++b) acc[a][b] = 0.0;
  const double* pr = sm.S + (base + ty) * LF_LD + k0;
  const double* pc = sm.S + (base + tx) * LF_LD + k0;
#pragma unroll 4
  for (int k = 0; k < 32; ++k) {
    double rv[NA], cv[NA];
#pragma unroll
    for (int a = 0; a < NA; ++a) { rv[a] = pr[a * 16 * LF_LD + k]; cv[a] = pc[a * 16 * LF_LD + k]; }
#pragma unroll
    for (int a = 0; a < NA; ++a)
#pragma unroll
      for (int b = 0; b <= a; ++b) acc[a][b] = fma(rv[a], cv[b], acc[a][b]);
  }
#pragma unroll
  for (int a = 0; a < NA; ++a)
#pragma unroll
    for (int b = 0; b <= a; ++b) {
      const int i = base + ty + 16 * a, c = base + tx + 16 * b;
      if (c <= i) sm.S[i * LF_LD + c] -= acc[a][b];
    }
}

// I1: inverse of the 32x32 lower-triangular diagonal block b; lane c computes column c (zeros above the diagonal)
__device__ __forceinline__ void leaf_diag_inverse(LeafSmem& sm, int b, int lane) {
  const int k0 = 32 * b;
  double t[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) t[i] = (i == lane) ? 1.0 : 0.0;
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    const double mk = t[k] * sm.dinv[k0 + k];
    t[k] = mk;
    const double* lcol = sm.S + (k0 + k) * LF_LD + k0 + k;   // L[k0+i][k0+k] = lcol[(i-k)*LF_LD]
#pragma unroll
    for (int i = k + 1; i < 32; ++i) t[i] = fma(-lcol[(i - k) * LF_LD], mk, t[i]);
  }
#pragma unroll
  for (int i = 0; i < 32; ++i) sm.Mkk[b][i * 33 + lane] = t[i];
}

// cycle stamps of the last leaf launch (thread 0): load, P1, P2, P3 (summed over kb), I1, I2, store
__device__ long long g_leaf_cycles[8];
#define LEAF_STAMP(slot) do { if (tid == 0) { long long t__ = clock64(); cyc[slot] += t__ - tprev; tprev = t__; } } while (0)

// M[k][c] for k >= c in different 32-blocks is stored transposed in the upper triangle: S[c][k]
__global__ void __launch_bounds__(LEAF_THREADS, 1)
gpk_leaf_potrf_inv_kernel(double* __restrict__ A, int lda, double* __restrict__ Mlow, double* __restrict__ Mup,
                          int ldm, int off, int* __restrict__ info) {
  extern __shared__ __align__(16) unsigned char leaf_smem_raw[];
  LeafSmem& sm = *reinterpret_cast<LeafSmem*>(leaf_smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tx = tid & 15, ty = tid >> 4;
  double* Ab = A + (size_t)off * lda + off;
  long long cyc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
  // 16-byte loads, 8 in flight per thread (the block is 128 KB: latency-bound on one SM otherwise)
#pragma unroll 1
  for (int it = 0; it < 32; it += 8) {
    double2 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int idx = tid + (it + u) * LEAF_THREADS, i = idx >> 6, c = (idx & 63) * 2;
      v[u] = (c <= i) ? *reinterpret_cast<const double2*>(Ab + (size_t)i * lda + c) : make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int idx = tid + (it + u) * LEAF_THREADS, i = idx >> 6, c = (idx & 63) * 2;
      sm.S[i * LF_LD + c] = v[u].x;
      sm.S[i * LF_LD + c + 1] = (c + 1 <= i) ? v[u].y : 0.0;
    }
  }
  if (tid < 128) sm.col[0][tid] = 0.0;
  __syncthreads();
  LEAF_STAMP(0);
  // ---- factorisation ------------------------------------------------------------------------------
#pragma unroll 1
  for (int kb = 0; kb < 4; ++kb) {
    const int k0 = 32 * kb, base = k0 + 32;
    if (warp < 4) leaf_diag_potrf4(sm, k0, lane, warp, off, info);
    __syncthreads();
    LEAF_STAMP(1);
    if (kb == 3) break;
    if (tid < GPK_TILE - base) leaf_panel_row(sm, k0, base + tid);
    __syncthreads();
    LEAF_STAMP(2);
    if (kb == 0) leaf_trailing<3>(sm, k0, base, tx, ty);
    else if (kb == 1) leaf_trailing<2>(sm, k0, base, tx, ty);
    else leaf_trailing<1>(sm, k0, base, tx, ty);
    __syncthreads();
    LEAF_STAMP(3);
  }
  // ---- inverse --------------------------------------------------------------------------------------
  if (warp < 4) leaf_diag_inverse(sm, warp, lane);
  __syncthreads();
  LEAF_STAMP(4);
  // level d handles the 4-d blocks (ib = jb + d); 64 threads per block, each a 4x4 register tile
  // (rows ty4+8a, cols tx4+8b): 8 shared loads per 16 FMAs
  {
    const int bi = tid >> 6, t64 = tid & 63, tx4 = t64 & 7, ty4 = t64 >> 3;
#pragma unroll 1
    for (int d = 1; d < 4; ++d) {
      const bool active = bi < 4 - d;
      const int ib = bi + d, jb = bi, i0 = 32 * ib, j0 = 32 * jb;
      if (active) {
        // W[r][c] = sum_{k=j0}^{i0-1} L[i0+r][k] M[k][j0+c]
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        const double* lrow = sm.S + (i0 + ty4) * LF_LD;
        const double* mk = sm.Mkk[jb] + tx4;
#pragma unroll 2
        for (int k = 0; k < 32; ++k) {                               // k inside block jb: diagonal block of M
          double lv[4], mv[4];
#pragma unroll
          for (int a = 0; a < 4; ++a) { lv[a] = lrow[a * 8 * LF_LD + j0 + k]; mv[a] = mk[k * 33 + 8 * a]; }
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fma(lv[a], mv[b], acc[a][b]);
        }
        const double* mt = sm.S + (j0 + tx4) * LF_LD;
#pragma unroll 2
        for (int k = j0 + 32; k < i0; ++k) {                         // later blocks: transposed storage
          double lv[4], mv[4];
#pragma unroll
          for (int a = 0; a < 4; ++a) { lv[a] = lrow[a * 8 * LF_LD + k]; mv[a] = mt[a * 8 * LF_LD + k]; }
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fma(lv[a], mv[b], acc[a][b]);
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) sm.W[bi][(ty4 + 8 * a) * 33 + tx4 + 8 * b] = acc[a][b];
      }
      __syncthreads();
      if (active) {
        // M_ij[r][c] = -sum_{k<=r} M_ii[r][k] W[k][c]  -> stored transposed at S[j0+c][i0+r]
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        const double* mi = sm.Mkk[ib] + ty4 * 33;
        const double* wk = sm.W[bi] + tx4;
#pragma unroll 2
        for (int k = 0; k < 32; ++k) {
          double lv[4], mv[4];
#pragma unroll
          for (int a = 0; a < 4; ++a) { lv[a] = mi[a * 8 * 33 + k]; mv[a] = wk[k * 33 + 8 * a]; }
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fma(lv[a], mv[b], acc[a][b]);
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) sm.S[(j0 + tx4 + 8 * b) * LF_LD + i0 + ty4 + 8 * a] = -acc[a][b];
      }
      __syncthreads();
    }
  }
  LEAF_STAMP(5);
  // ---- store L (lower), L^-1 (lower, zeros above), L^-T (upper, zeros below) -------------------------
  double* Ml = Mlow + (size_t)off * ldm + off;
  double* Mu = Mup + (size_t)off * ldm + off;
  // each thread produces 2 adjacent columns of one row per iteration (16-byte stores)
#pragma unroll 4
  for (int it = 0; it < 32; ++it) {
    const int idx = tid + it * LEAF_THREADS, i = idx >> 6, c = (idx & 63) * 2;   // row i, columns c, c+1
    const int bi = i >> 5, bc = c >> 5;
    if (c <= i) {
      double2 l = make_double2(sm.S[i * LF_LD + c], sm.S[i * LF_LD + c + 1]);
      if (c + 1 <= i) *reinterpret_cast<double2*>(Ab + (size_t)i * lda + c) = l;
      else Ab[(size_t)i * lda + c] = l.x;
    }
    double2 m = make_double2(0.0, 0.0), u = make_double2(0.0, 0.0);
    if (bc == bi) {            // diagonal 32-block: Mkk holds explicit zeros above the diagonal
      const double* mk = sm.Mkk[bi];
      m = make_double2(mk[(i & 31) * 33 + (c & 31)], mk[(i & 31) * 33 + (c & 31) + 1]);
      u = make_double2(mk[(c & 31) * 33 + (i & 31)], mk[((c & 31) + 1) * 33 + (i & 31)]);
    } else if (bc < bi) {      // Mlow[i][c] = M[i][c] = S[c][i]
      m = make_double2(sm.S[c * LF_LD + i], sm.S[(c + 1) * LF_LD + i]);
    } else {                   // Mup[i][c] = M[c][i] = S[i][c]
      u = make_double2(sm.S[i * LF_LD + c], sm.S[i * LF_LD + c + 1]);
    }
    *reinterpret_cast<double2*>(Ml + (size_t)i * ldm + c) = m;
    *reinterpret_cast<double2*>(Mu + (size_t)i * ldm + c) = u;
  }
  LEAF_STAMP(6);
  if (tid == 0) {
#pragma unroll
    for (int q = 0; q < 8; ++q) g_leaf_cycles[q] = cyc[q];
  }
}

cudaError_t gpk_leaf_cycles(long long* out_h) {
  return cudaMemcpyFromSymbol(out_h, g_leaf_cycles, sizeof(long long) * 8);
}

static cudaError_t leaf_configure() {
  static bool done[GPK_MAX_DEV] = {false};
  const int dev = gpk_current_device();
  if (done[dev]) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(gpk_leaf_potrf_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       LEAF_SMEM_BYTES);
  if (e == cudaSuccess) done[dev] = true;
  return e;
}


// Side streams for the one product of each level that nothing in the sub-recursion depends on
// (Wt = M11^T-part * L21^T, needed only by M21 after rec(22)): it fills the SMs the small GEMMs and the
// latency-bound leaves of rec(22) leave idle.  One fork event (re-recorded per level: a wait captures the state at
// enqueue time) and one join event per recursion depth (one outstanding product per depth).
#define GPK_CHOL_MAXDEPTH 24
struct CholSide { cudaStream_t side[GPK_CHOL_MAXDEPTH]; cudaEvent_t join[GPK_CHOL_MAXDEPTH]; cudaEvent_t fork; bool ready; };
static CholSide g_cs[GPK_MAX_DEV];   // one per device; a parent's large product must not queue ahead of its children's
static cudaError_t chol_side_init(CholSide& cs) {
  if (cs.ready) return cudaSuccess;
  cudaError_t e;
  for (int i = 0; i < GPK_CHOL_MAXDEPTH; ++i) {
    if ((e = cudaStreamCreateWithFlags(&cs.side[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&cs.join[i], cudaEventDisableTiming)) != cudaSuccess) return e;
  }
  if ((e = cudaEventCreateWithFlags(&cs.fork, cudaEventDisableTiming)) != cudaSuccess) return e;
  cs.ready = true;
  return cudaSuccess;
}

struct CholCtx { double* A; int lda; double* Ml; double* Mu; int ldm; int* info; cudaStream_t st; cudaError_t err;
                 CholSide* side; bool keep_L; };

// One node of the recursion on [off, off+n), split n1 | n2 at a multiple of 128:
//   rec(11);  L21 = A21 M11^T -> Mlow21;  [side stream] Wt = M11^T-part * L21^T -> A12 (scratch: the strictly upper
//   part of A is never read);  A22 -= L21 L21^T;  rec(22);  M21 = -M22 Wt^T -> Mlow21 AND, transposed by the same
//   GEMM's epilogue, Mup12.
// L21 is consumed where the GEMM left it (Mlow21) and is copied over A21 only when the caller wants L itself
// (keep_L: the final factorisation, whose L_ the estimator may download); an objective evaluation needs only L's
// diagonal (log-det), which the leaves leave in A.
static void chol_rec(CholCtx& c, int off, int n, int depth = 0) {
  if (c.err != cudaSuccess) return;
  if (n <= GPK_TILE) {
    if ((c.err = leaf_configure()) != cudaSuccess) return;
    gpk_leaf_potrf_inv_kernel<<<1, LEAF_THREADS, LEAF_SMEM_BYTES, c.st>>>(c.A, c.lda, c.Ml, c.Mu, c.ldm, off, c.info); GPK_LAUNCHED();
    c.err = cudaGetLastError();
    return;
  }
  const int n1 = GPK_TILE * ((n / GPK_TILE + 1) / 2), n2 = n - n1, o2 = off + n1;
  chol_rec(c, off, n1, depth + 1);
  if (c.err != cudaSuccess) return;
  double* A21 = c.A + (size_t)o2 * c.lda + off;
  double* A12 = c.A + (size_t)off * c.lda + o2;      // scratch for Wt (n1 x n2)
  double* Ml21 = c.Ml + (size_t)o2 * c.ldm + off;
  double* Mu12 = c.Mu + (size_t)off * c.ldm + o2;
  GpkGemmArgs g{};
  // L21 = A21 * M11^T -> Mlow21
  g.A = A21; g.lda = c.lda; g.B = c.Ml + (size_t)off * c.ldm + off; g.ldb = c.ldm; g.C = Ml21; g.ldc = c.ldm;
  g.M = n2; g.N = n1; g.K = n1; g.alpha = 1.0; g.accumulate = 0; g.a_tri = GPK_TRI_NONE; g.b_tri = GPK_TRI_LOWER;
  g.lower_only = 0; g.epilogue = GPK_EPI_STORE; g.rowsq = nullptr;
  if ((c.err = gpk_launch_gemm_nt(g, c.st)) != cudaSuccess) return;
  if (c.keep_L && (c.err = gpk_launch_copy_block(Ml21, c.ldm, A21, c.lda, n2, n1, c.st)) != cudaSuccess) return;
  // Wt = Mt11 * L21^T -> A12: independent of the A22 update and of rec(22) -> side stream
  const bool fork = c.side != nullptr && depth < GPK_CHOL_MAXDEPTH;
  cudaStream_t ws = fork ? c.side->side[depth] : c.st;
  if (fork) {
    if ((c.err = cudaEventRecord(c.side->fork, c.st)) != cudaSuccess) return;    // after L21 above
    if ((c.err = cudaStreamWaitEvent(ws, c.side->fork, 0)) != cudaSuccess) return;
  }
  GpkGemmArgs w{};
  w.A = c.Mu + (size_t)off * c.ldm + off; w.lda = c.ldm; w.B = Ml21; w.ldb = c.ldm; w.C = A12; w.ldc = c.lda;
  w.M = n1; w.N = n2; w.K = n1; w.alpha = 1.0; w.a_tri = GPK_TRI_UPPER; w.epilogue = GPK_EPI_STORE;
  if (fork) {
    if ((c.err = gpk_launch_gemm_nt(w, ws)) != cudaSuccess) return;
    if ((c.err = cudaEventRecord(c.side->join[depth], ws)) != cudaSuccess) return;
  }
  // A22 -= L21 L21^T (lower tiles)
  g = GpkGemmArgs{};
  g.A = Ml21; g.lda = c.ldm; g.B = Ml21; g.ldb = c.ldm; g.C = c.A + (size_t)o2 * c.lda + o2; g.ldc = c.lda;
  g.M = n2; g.N = n2; g.K = n1; g.alpha = -1.0; g.accumulate = 1; g.lower_only = 1; g.epilogue = GPK_EPI_STORE;
  if ((c.err = gpk_launch_gemm_nt(g, c.st)) != cudaSuccess) return;
  chol_rec(c, o2, n2, depth + 1);
  if (c.err != cudaSuccess) return;
  if (fork) {
    if ((c.err = cudaStreamWaitEvent(c.st, c.side->join[depth], 0)) != cudaSuccess) return;
  } else {
    if ((c.err = gpk_launch_gemm_nt(w, c.st)) != cudaSuccess) return;
  }
  // M21 = -M22 * W : C[m,j] = -sum_k M22[m,k] Wt[j,k] -> Mlow21, and its transpose -> Mup12 from the same epilogue
  g = GpkGemmArgs{};
  g.A = c.Ml + (size_t)o2 * c.ldm + o2; g.lda = c.ldm; g.B = A12; g.ldb = c.lda; g.C = Ml21; g.ldc = c.ldm;
  g.M = n2; g.N = n1; g.K = n2; g.alpha = -1.0; g.a_tri = GPK_TRI_LOWER; g.epilogue = GPK_EPI_STORE;
  g.Ct = Mu12; g.ldct = c.ldm;
  c.err = gpk_launch_gemm_nt(g, c.st);
}

// The side streams and the fork/join events are per device: one enqueue at a time per device (two host threads
// factoring on the same GPU would otherwise re-record the fork event between the other's record and wait).
#include <mutex>
static std::mutex g_chol_mu[GPK_MAX_DEV];

// keep_L = false: objective evaluations (gpk_lml_eval) -- A ends with L's diagonal 128-blocks only.
cudaError_t gpk_launch_potrf_inv_ex(double* A, int lda, double* Mlow, double* Mup, int ldm, int NP, int* info,
                                    cudaStream_t st, bool keep_L) {
  if (NP % GPK_TILE) return cudaErrorInvalidValue;
  std::lock_guard<std::mutex> lock(g_chol_mu[gpk_current_device()]);
  cudaError_t e = cudaMemsetAsync(info, 0, sizeof(int), st);
  if (e != cudaSuccess) return e;
  // the side stream is bypassed while per-launch GEMM timing is on (overlapped durations would not add up)
  CholSide* side = nullptr;
  if (!g_gpk_prof.enabled && NP > 2 * GPK_TILE) {
    side = &g_cs[gpk_current_device()];
    if ((e = chol_side_init(*side)) != cudaSuccess) return e;
  }
  CholCtx c{A, lda, Mlow, Mup, ldm, info, st, cudaSuccess, side, keep_L};
  chol_rec(c, 0, NP);
  return c.err;
}

cudaError_t gpk_launch_potrf_inv(double* A, int lda, double* Mlow, double* Mup, int ldm, int NP, int* info,
                                 cudaStream_t st) {
  return gpk_launch_potrf_inv_ex(A, lda, Mlow, Mup, ldm, NP, info, st, true);
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
// Pure streaming read of N^2/2 doubles: every lane keeps four 16-byte loads in flight per trip (a warp with one
// 8-byte load outstanding per lane leaves the short rows latency-bound: 3.3-4.0 TB/s measured with that form).
template <int TT>
__global__ void __launch_bounds__(256) gpk_trimv_kernel(const double* __restrict__ Mt, int ldm, int NP,
                                                        const double* __restrict__ v, int ldv, double* __restrict__ out,
                                                        int ldo, int upper, int t0, int T) {
  int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= NP) return;
  const int kb = upper ? (row & ~1) : 0, ke = upper ? NP : row + 1;   // [kb, ke), kb even; elements outside the
  const int lo = upper ? row : 0;                                      // triangle (k < lo or k >= ke) are masked
  double acc[TT];
#pragma unroll
  for (int t = 0; t < TT; ++t) acc[t] = 0.0;
  const double* mr = Mt + (size_t)row * ldm;
  for (int k0 = kb + 2 * lane; k0 < ke; k0 += 256) {
    double2 m[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + 64 * u;
      m[u] = (k < ke) ? *reinterpret_cast<const double2*>(mr + k) : make_double2(0.0, 0.0);   // k + 1 < NP: NP is even
      if (k < lo) m[u].x = 0.0;
      if (k + 1 >= ke) m[u].y = 0.0;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + 64 * u;
      if (k < ke) {
        const double* vk = v + (size_t)k * ldv + t0;
#pragma unroll
        for (int t = 0; t < TT; ++t)
          if (t0 + t < T) acc[t] = fma(m[u].x, vk[t], fma(m[u].y, (k + 1 < ke) ? vk[ldv + t] : 0.0, acc[t]));
      }
    }
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
