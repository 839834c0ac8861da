"""Numerical study (CPU, NumPy only): how many int8 slices would an Ozaki-scheme FP64 GEMM on the tcgen05 INT8 pipe
need before the GP path's own tolerances hold?  ("FP64 emulation on tcgen05".)

Nothing here is on the product path and nothing here runs on the GPU: it restates the splitting arithmetic exactly
(slice products of small integers are exact in float64, as they are in int32 on the tensor pipe) and pushes the
three contractions of the path through it:
  predict   V = Ks L^-T, var = prior - rowsum(V^2)             tolerance: std 1e-8 absolute (north star)
  inverse   K^-1 = L^-T L^-1 (LAUUM) feeding the LML gradient   tolerance: gradient 1e-6 relative
  factor    every GEMM of the recursive Cholesky+inverse        tolerance: LML 1e-8 rel, L 1e-10 abs, alpha 1e-7 rel

Splitting: row i of an operand is scaled by mu_i = the power of two just above max|a_i.|; slice 1 keeps 6 bits + sign, every
further slice 7 bits + sign (all fit int8); slice products with s + t <= S + 1 are kept (S(S+1)/2 products) and
summed per s + t in one exact integer accumulator, then combined in float64 from the smallest scale up.

  python tools/ozaki_study.py [--n 2048] [--slices 4 5 6 7 8]
"""
import argparse
import json
import time

import numpy as np
from scipy.linalg import cholesky, solve_triangular
from scipy.spatial.distance import cdist


def split(A, S):
    """rows of A -> (mu (rows,), [q_1..q_S] integer-valued float64 arrays, shifts[s] = bits kept through slice s)"""
    amax = np.max(np.abs(A), axis=1)
    _, e = np.frexp(amax)                     # amax = f * 2^e, f in [0.5, 1)  (csrc/gpk_i8gemm.cu does the same)
    mu = np.where(amax > 0, np.ldexp(1.0, e), 1.0)
    r = A / mu[:, None]                       # |r| < 1, exact (power of two)
    qs, shifts, bits = [], [], 0
    for s in range(S):
        b = 6 if s == 0 else 7
        r = r * 2.0 ** b
        q = np.rint(r)
        r = r - q                             # |r| <= 0.5, exact
        bits += b
        qs.append(q)
        shifts.append(bits)
    assert max(np.max(np.abs(q)) for q in qs) <= 64
    return mu, qs, shifts


def gemm_nt_ozaki(A, B, S):
    """A (M,K) B (N,K) -> A B^T through S int8 slices per operand"""
    mua, qa, sa = split(A, S)
    mub, qb, sb = split(B, S)
    groups = {}
    for s in range(S):
        for t in range(S):
            if s + t <= S - 1:                # 0-based: s + t <= S - 1  <=>  1-based s + t <= S + 1
                groups.setdefault(sa[s] + sb[t], []).append((s, t))
    C = np.zeros((A.shape[0], B.shape[0]))
    for scale in sorted(groups, reverse=True):        # smallest contribution first
        P = np.zeros_like(C)
        for s, t in groups[scale]:
            P += qa[s] @ qb[t].T                      # exact: |entries| <= 64*64*K*S < 2^53
        C += P * 2.0 ** (-scale)
    return C * mua[:, None] * mub[None, :]


def make_gemm(S):
    if S == 0:
        return lambda A, B: A @ B.T
    return lambda A, B: gemm_nt_ozaki(A, B, S)


def leaf(A):
    L = cholesky(A, lower=True, check_finite=False)
    M = solve_triangular(L, np.eye(len(A)), lower=True, check_finite=False)
    return L, M


def chol_inv(A, gemm, nb=128, gemm_inv=None, gemm_tail=None, on_tail=True):
    """same recursion as csrc/gpk_chol.cu chol_rec: returns (L, M = L^-1).  The factor's products (L21, trailing
    update) go through gemm(A, B) = A B^T; the inverse's (Wt, M21) through gemm_inv (default: the same), except on
    the right-most path of the recursion (root, root.22, root.22.22, ...), whose inverse blocks feed no later
    L21 -- only K^-1 and predict -- and go through gemm_tail (default: gemm_inv).  That path holds 6/7 of the
    inverse's flops (N^3/4 at the root, an eighth of that per level down)."""
    gemm_inv = gemm_inv or gemm
    gemm_tail = gemm_tail or gemm_inv
    n = len(A)
    if n <= nb:
        return leaf(A)
    h = (n // 2 + nb - 1) // nb * nb
    L11, M11 = chol_inv(A[:h, :h], gemm, nb, gemm_inv, gemm_tail, False)
    L21 = gemm(A[h:, :h], M11)                         # A21 M11^T
    L22, M22 = chol_inv(A[h:, h:] - gemm(L21, L21), gemm, nb, gemm_inv, gemm_tail, on_tail)
    gi = gemm_tail if on_tail else gemm_inv
    Wt = gi(M11.T.copy(), L21)                         # (M11^T) L21^T = (L21 M11)^T
    M21 = -gi(M22, Wt)                                 # -M22 (L21 M11)
    L = np.zeros((n, n)); M = np.zeros((n, n))
    L[:h, :h], L[h:, :h], L[h:, h:] = L11, L21, L22
    M[:h, :h], M[h:, :h], M[h:, h:] = M11, M21, M22
    return L, M


def gram(X, ls, c, noise):
    K = c * np.exp(-0.5 * cdist(X / ls, X / ls, "sqeuclidean"))
    K[np.diag_indices_from(K)] += noise
    return K


def gradient(X, ls, c, noise, alpha, Kinv):
    T = alpha.shape[1]
    Kc = c * np.exp(-0.5 * cdist(X / ls, X / ls, "sqeuclidean"))
    W = alpha @ alpha.T - T * Kinv
    WK = W * Kc
    Xs = X / ls
    g = [0.5 * np.sum(WK * (Xs[:, d][:, None] - Xs[:, d][None, :]) ** 2) for d in range(X.shape[1])]
    return np.array([0.5 * np.sum(WK)] + g + [0.5 * noise * np.trace(W)])


def case(name, X, Y, ls, c, noise, jitter, Xq, slices):
    N = len(X)
    K = gram(X, ls, c, noise + jitter)
    out = []
    ref = None
    for S in [0] + slices:
        t0 = time.time()
        g = make_gemm(S)
        try:
            L, M = chol_inv(K, g)
        except np.linalg.LinAlgError:                  # truncation error above the smallest eigenvalue
            row = {"case": name, "N": N, "slices": S, "products": S * (S + 1) // 2, "all_gemms": "not positive definite"}
            out.append(row)
            print(json.dumps(row), flush=True)
            continue
        u = M @ Y
        alpha = M.T @ u
        lml = float(np.sum(-0.5 * np.einsum("ik,ik->k", Y, alpha)) - Y.shape[1] * np.log(np.diag(L)).sum()
                    - Y.shape[1] * N / 2 * np.log(2 * np.pi))
        Kinv = g(M.T.copy(), M.T.copy())               # L^-T (L^-T)^T
        grad = gradient(X, ls, c, noise, alpha, Kinv)
        Ks = c * np.exp(-0.5 * cdist(Xq / ls, X / ls, "sqeuclidean"))
        V = g(Ks, M)                                   # Ks L^-T
        var = (c + noise) - np.sum(V * V, axis=1)
        std = np.sqrt(np.maximum(var, 0.0))
        cur = dict(L=L, alpha=alpha, lml=lml, grad=grad, std=std)
        if S == 0:
            ref = cur
            # also: FP64 factor, emulated contraction only (what a predict-only or LAUUM-only use would see)
            ref["M"], ref["Ks"] = M, Ks
            continue
        rel = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
        Vp = g(ref["Ks"], ref["M"])
        std_p = np.sqrt(np.maximum((c + noise) - np.sum(Vp * Vp, axis=1), 0.0))
        Kinv_p = g(ref["M"].T.copy(), ref["M"].T.copy())
        grad_p = gradient(X, ls, c, noise, ref["alpha"], Kinv_p)
        # FP64 factor products, emulated inverse products + LAUUM (2/3 of the step's flops)
        Lm, Mm = chol_inv(K, make_gemm(0), gemm_inv=g)
        am = Mm.T @ (Mm @ Y)
        lml_m = float(np.sum(-0.5 * np.einsum("ik,ik->k", Y, am)) - Y.shape[1] * np.log(np.diag(Lm)).sum()
                      - Y.shape[1] * N / 2 * np.log(2 * np.pi))
        grad_m = gradient(X, ls, c, noise, am, g(Mm.T.copy(), Mm.T.copy()))
        Vm = g(ref["Ks"], Mm)
        std_m = np.sqrt(np.maximum((c + noise) - np.sum(Vm * Vm, axis=1), 0.0))
        # FP64 everywhere the factor reads; emulated right-most-path inverse blocks + LAUUM + predict.
        # alpha comes from the FP64 factor (substitution over L, as the product would do it), not from M.
        f64 = make_gemm(0)
        Lt, Mt = chol_inv(K, f64, gemm_inv=f64, gemm_tail=g)
        grad_t = gradient(X, ls, c, noise, ref["alpha"], g(Mt.T.copy(), Mt.T.copy()))
        Vt = g(ref["Ks"], Mt)
        std_t = np.sqrt(np.maximum((c + noise) - np.sum(Vt * Vt, axis=1), 0.0))
        assert np.array_equal(Lt, ref["L"])
        # alpha through the tail-emulated M, then one refinement step over the FP64 factor:
        #   r = y - L (L^T a0),  a = a0 + M^T (M r)
        a0 = Mt.T @ (Mt @ Y)
        a1 = a0 + Mt.T @ (Mt @ (Y - Lt @ (Lt.T @ a0)))
        amax = np.max(np.abs(ref["alpha"]))
        lml_of = lambda al: float(np.sum(-0.5 * np.einsum("ik,ik->k", Y, al)) - Y.shape[1] * np.log(np.diag(Lt)).sum()
                                  - Y.shape[1] * N / 2 * np.log(2 * np.pi))
        row = {
            "case": name, "N": N, "slices": S, "products": S * (S + 1) // 2,
            "tail_lauum_grad_rel": rel(grad_t, ref["grad"]),
            "tail_lauum_predict_std_abs": float(np.max(np.abs(std_t - ref["std"]))),
            "tail_alpha_rel": float(np.max(np.abs(a0 - ref["alpha"])) / amax),
            "tail_alpha_refined_rel": float(np.max(np.abs(a1 - ref["alpha"])) / amax),
            "tail_lml_rel": abs(lml_of(a0) - ref["lml"]) / abs(ref["lml"]),
            "tail_lml_refined_rel": abs(lml_of(a1) - ref["lml"]) / abs(ref["lml"]),
            "inv_lauum_L_abs": float(np.max(np.abs(Lm - ref["L"]))),
            "inv_lauum_alpha_rel": float(np.max(np.abs(am - ref["alpha"])) / np.max(np.abs(ref["alpha"]))),
            "inv_lauum_lml_rel": abs(lml_m - ref["lml"]) / abs(ref["lml"]),
            "inv_lauum_grad_rel": rel(grad_m, ref["grad"]),
            "inv_lauum_predict_std_abs": float(np.max(np.abs(std_m - ref["std"]))),
            "predict_only_std_abs": float(np.max(np.abs(std_p - ref["std"]))),
            "lauum_only_grad_rel": rel(grad_p, ref["grad"]),
            "all_gemms_L_abs": float(np.max(np.abs(L - ref["L"]))),
            "all_gemms_alpha_rel": float(np.max(np.abs(alpha - ref["alpha"])) / np.max(np.abs(ref["alpha"]))),
            "all_gemms_lml_rel": abs(lml - ref["lml"]) / abs(ref["lml"]),
            "all_gemms_grad_rel": rel(grad, ref["grad"]),
            "all_gemms_std_abs": float(np.max(np.abs(std - ref["std"]))),
            "seconds": round(time.time() - t0, 1),
        }
        out.append(row)
        print(json.dumps(row), flush=True)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=2048)
    ap.add_argument("--slices", type=int, nargs="+", default=[4, 5, 6, 7, 8])
    a = ap.parse_args()
    N = a.n
    rng = np.random.default_rng(20251129)                          # BASELINE.json config 2 generator
    X = rng.standard_normal((N, 9))
    w = rng.standard_normal((9, 3)) / 3
    Y = np.sin(X @ w) + 0.1 * rng.standard_normal((N, 3))
    Xq = np.vstack([X[:256], X[:256] + 0.05 * rng.standard_normal((256, 9))])   # training rows: worst cancellation
    # (i) near the config-2 optimum; (ii) the corner L-BFGS-B visits: long length-scales, smallest noise
    case("config2_optimum", X, Y, np.full(9, 3.0), 1.0, 0.01, 1e-6, Xq, a.slices)
    case("config2_corner_ls10_noise1e-5", X, Y, np.full(9, 10.0), 1.0, 1e-5, 1e-6, Xq, a.slices)
    # (iii) the reference trainer's production kernel shape: RBF(0.0916)+White(0.282), alpha 1e-4, D=10, T=6,
    #       inputs on the scale of its CSV (positions +-6, velocities < 1)
    X10 = np.hstack([rng.uniform(-6, 6, (N, 2)), 3.09 + 0.04 * rng.standard_normal((N, 1)),
                     0.5 * rng.standard_normal((N, 3)), 0.3 * rng.standard_normal((N, 4))])
    Y6 = 0.05 * rng.standard_normal((N, 6))
    Xq10 = np.vstack([X10[:256], X10[:256] + 0.01 * rng.standard_normal((256, 10))])
    case("reference_kernel_iso", X10, Y6, np.full(10, 0.0916), 1.0, 0.282, 1e-4, Xq10, a.slices)


if __name__ == "__main__":
    main()
