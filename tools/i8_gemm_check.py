"""First thing to run on a B200 for the experimental tcgen05 int8 contraction (csrc/gpk_i8gemm.cu):

  timeout 300 python tools/i8_gemm_check.py            # correctness ladder, smallest case first, then timing
  GPK_I8_EXPERIMENT=1 python -m pytest tests/test_i8_gemm.py -m gpu -x -q

Every step prints one JSON line and the script stops at the first failure, so a wrong descriptor shows up on the
128 x 64 x 64 single-slice case instead of inside a 20k contraction.  Run it under `timeout`: a wrong mbarrier
protocol hangs instead of failing.
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from test_i8_gemm import gemm_nt_sliced, split  # noqa: E402  (NumPy statement of the arithmetic; no oracle involved)

from unmanned_aerial_vehicles_b200 import _lib  # noqa: E402


def sliced(lib, A, S, row_mult, tri=0):
    rows, K = A.shape
    rowsP, KP = -(-rows // row_mult) * row_mult, -(-K // 64) * 64
    planes = torch.empty((S, rowsP, KP), dtype=torch.int8, device="cuda")
    mu = torch.empty(rowsP, dtype=torch.float64, device="cuda")
    rc = lib.gpk_i8_slice(A.data_ptr(), A.stride(0), rows, K, tri, S, planes.data_ptr(), rowsP, KP, mu.data_ptr(),
                          torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    return planes, mu, rowsP, KP


def main():
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(0)
    ok = True
    for m, n, k, S in [(128, 64, 64, 1), (128, 64, 128, 1), (128, 64, 64, 2), (256, 128, 512, 4), (512, 512, 4096, 7),
                       (1024, 1024, 20096, 8)]:
        A, B = rng.standard_normal((m, k)), rng.standard_normal((n, k))
        Ad, Bd = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
        pa, mua, MP, KP = sliced(lib, Ad, S, 128)
        pb, mub, NP, _ = sliced(lib, Bd, S, 64)
        mu_h, qs, _ = split(A, S)
        planes_ok = bool(np.array_equal(mua.cpu().numpy()[:m], mu_h)) and all(
            np.array_equal(pa[s, :m, :k].cpu().numpy(), qs[s].astype(np.int8)) for s in range(S))
        C = torch.full((MP, NP), float("nan"), dtype=torch.float64, device="cuda")
        rc = lib.gpk_i8_gemm_nt(pa.data_ptr(), mua.data_ptr(), MP, pb.data_ptr(), mub.data_ptr(), NP, KP, S,
                                C.data_ptr(), NP, 1.0, 0, 0, 0, 0, None, st)
        torch.cuda.synchronize()
        got = C.cpu().numpy()[:m, :n]
        want = gemm_nt_sliced(A, B, S)
        ref = A @ B.T
        row = {"case": [m, n, k, S], "rc": rc, "planes_bit_identical": planes_ok,
               "product_bit_identical": bool(np.array_equal(got, want)),
               "max_abs_vs_numpy_statement": float(np.nanmax(np.abs(got - want))) if np.isfinite(got).any() else None,
               "nan_left": int(np.isnan(got).sum()),
               "max_rel_vs_fp64": float(np.max(np.abs(want - ref)) / np.max(np.abs(ref)))}
        print(json.dumps(row), flush=True)
        ok = ok and rc == 0 and planes_ok and row["product_bit_identical"]
        if not ok:
            sys.exit(1)

    # timing: LAUUM-shaped (upper x upper, lower tiles only) and predict-shaped (full x lower, row sums of squares)
    for n, S in [(8192, 7), (20096, 7)]:
        U = torch.triu(torch.randn(n, n, dtype=torch.float64, device="cuda"))
        t0 = time.time()
        pu, mu, NP, KP = sliced(lib, U, S, 128, tri=2)
        torch.cuda.synchronize()
        t_slice = time.time() - t0
        C = torch.empty((NP, NP), dtype=torch.float64, device="cuda")
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        for it in range(3):
            if it == 1:
                ev[0].record()
            rc = lib.gpk_i8_gemm_nt(pu.data_ptr(), mu.data_ptr(), NP, pu.data_ptr(), mu.data_ptr(), NP, KP, S,
                                    C.data_ptr(), NP, 1.0, 0, 2, 2, 1, None, st)
            assert rc == 0, rc
        ev[1].record()
        torch.cuda.synchronize()
        sec = ev[0].elapsed_time(ev[1]) / 2e3
        print(json.dumps({"shape": "lauum", "n": n, "slices": S, "slice_seconds_first_call": round(t_slice, 4),
                          "seconds": sec, "fp64_equivalent_tflops": n ** 3 / 3 / sec / 1e12,
                          "int8_tops": n ** 3 / 3 * S * (S + 1) / 2 / sec / 1e12}), flush=True)
        del U, C, pu
        Mq = 8192
        Ks = torch.rand(Mq, n, dtype=torch.float64, device="cuda")
        L = torch.tril(torch.randn(n, n, dtype=torch.float64, device="cuda"))
        pq, mq, MP, _ = sliced(lib, Ks, S, 128)
        pl, ml, NP, KP = sliced(lib, L, S, 64, tri=1)
        rowsq = torch.empty((MP, NP // 64), dtype=torch.float64, device="cuda")
        for it in range(3):
            if it == 1:
                ev[0].record()
            rc = lib.gpk_i8_gemm_nt(pq.data_ptr(), mq.data_ptr(), MP, pl.data_ptr(), ml.data_ptr(), NP, KP, S,
                                    None, 0, 1.0, 0, 0, 1, 0, rowsq.data_ptr(), st)
            assert rc == 0, rc
        ev[1].record()
        torch.cuda.synchronize()
        sec = ev[0].elapsed_time(ev[1]) / 2e3
        print(json.dumps({"shape": "predict_rowsumsq", "n": n, "rows": Mq, "slices": S, "seconds": sec,
                          "pts_per_s": Mq / sec, "fp64_equivalent_tflops": Mq * n * n / sec / 1e12}), flush=True)
        del Ks, L, pq, pl, rowsq


if __name__ == "__main__":
    main()
