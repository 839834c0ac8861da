"""gpk_predict_mean_fused_static -- the one entry point that allocates nothing (scratch in a file-scope device array;
no counterpart in the reference): called through the C ABI and checked
against the existing entry point, gpk_predict_mean_fused, on two calls in a row that share the static array.

Kept in a file of its own that pytest collects LAST (after test_zz_late_additions_gpu.py):
a failure here cannot hide any other test.  Self-contained: passes alone in a fresh process and after
the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _p(t):
    return C.c_void_p(t.data_ptr())


def _model(n, d, t, seed):
    """a fitted model's device-resident pieces through the engine (Gram + factor + alpha on the GPU)"""
    from unmanned_aerial_vehicles_b200.engine import DeviceGP
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    Y = np.sin(X @ (rng.standard_normal((d, t)) / np.sqrt(d))) + 0.1 * rng.standard_normal((n, t))
    eng = DeviceGP(X, Y, 1e-6)
    inv_ls = 1.0 / np.linspace(0.7, 2.0, d)
    assert eng.factor(inv_ls, 1.3, 0.02) == 0
    return eng, X, inv_ls


def _both(lib, _lib, eng, inv_ls, Xq):
    """(existing entry point with caller scratch, new entry point with the static array) on the same inputs"""
    M, D, T, N = Xq.shape[0], eng.D, eng.T, eng.N
    MP = -(-M // 128) * 128
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    xq = torch.zeros((MP, D), dtype=torch.float64, device="cuda")
    xq[:M] = torch.as_tensor(Xq, device="cuda")
    ym = torch.tensor(np.linspace(-0.5, 0.5, T), dtype=torch.float64, device="cuda")
    ys = torch.tensor(np.linspace(0.8, 1.7, T), dtype=torch.float64, device="cuda")
    inv = _lib.darr(inv_ls)
    ws = torch.empty(int(lib.gpk_predict_mean_workspace_doubles(MP, N, D, T)), dtype=torch.float64, device="cuda")
    want = torch.full((MP, T), np.nan, dtype=torch.float64, device="cuda")
    got = torch.full((MP, T), np.nan, dtype=torch.float64, device="cuda")
    rc = lib.gpk_predict_mean_fused(_p(xq), M, MP, _p(eng.X_d), N, D, inv, 1.3, _p(eng.alpha_d), T, _p(ym), _p(ys),
                                    _p(ws), _p(want), st)
    assert rc == 0
    torch.cuda.synchronize()
    mem0 = torch.cuda.memory_allocated()
    rc = lib.gpk_predict_mean_fused_static(_p(xq), M, MP, _p(eng.X_d), N, D, inv, 1.3, _p(eng.alpha_d), T, _p(ym),
                                           _p(ys), _p(got), st)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == mem0          # nothing went through torch's allocator either
    return want[:M].cpu().numpy(), got[:M].cpu().numpy()


def test_static_scratch_entry_point_matches_the_existing_one_on_two_calls_in_a_row():
    from unmanned_aerial_vehicles_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(21)
    # call 1: many query rows, one split of the training rows; call 2 (same array, stale contents of call 1 in it):
    # few rows, many splits, another model with other D and T
    eng1, X1, inv1 = _model(1500, 9, 3, seed=1)
    Xq1 = np.vstack([X1[:400] + 1e-3, 1.5 * rng.standard_normal((20300, 9))])
    eng2, X2, inv2 = _model(900, 10, 6, seed=2)
    Xq2 = X2[:25] + 0.05
    for eng, inv_ls, Xq in ((eng1, inv1, Xq1), (eng2, inv2, Xq2), (eng1, inv1, Xq1[:130])):
        want, got = _both(lib, _lib, eng, inv_ls, Xq)
        assert np.isfinite(want).all() and np.isfinite(got).all()
        scale = np.abs(want).max()
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-12 * scale)   # the tolerance of the mean-only tests ...
        np.testing.assert_array_equal(got, want)                            # ... and in fact the same bits
    # and against the estimator-level predict (host arrays through gpk_predict), mean 1e-6 relative (north star)
    m = eng2.predict(Xq2, np.linspace(-0.5, 0.5, 6), np.linspace(0.8, 1.7, 6), False)[0]
    want, got = _both(lib, _lib, eng2, inv2, Xq2)
    np.testing.assert_allclose(got, m, rtol=1e-6, atol=1e-9 * np.abs(m).max())


def test_static_scratch_entry_point_rejects_shapes_beyond_its_array():
    from unmanned_aerial_vehicles_b200 import _lib
    lib = _lib.load()
    eng, X, inv_ls = _model(300, 4, 3, seed=3)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ym = torch.zeros(3, dtype=torch.float64, device="cuda")
    ys = torch.ones(3, dtype=torch.float64, device="cuda")
    small = torch.zeros((128, 4), dtype=torch.float64, device="cuda")
    out = torch.full((128, 3), 7.0, dtype=torch.float64, device="cuda")
    MP = 262144                                    # one split: 262144 * 3 = 786432 doubles > GPK_STATIC_SCRATCH_DOUBLES
    assert MP * 3 > _lib.STATIC_SCRATCH_DOUBLES
    # rejected before anything is launched: the (too small) buffers are never touched
    rc = lib.gpk_predict_mean_fused_static(_p(small), MP, MP, _p(eng.X_d), eng.N, 4, _lib.darr(inv_ls), 1.0,
                                           _p(eng.alpha_d), 3, _p(ym), _p(ys), _p(out), st)
    assert rc == _lib.ERR_STATIC_SCRATCH_TOO_SMALL
    torch.cuda.synchronize()
    assert float(out.min()) == 7.0 and float(out.max()) == 7.0
    assert lib.gpk_predict_mean_fused_static(None, 1, 128, _p(eng.X_d), eng.N, 4, _lib.darr(inv_ls), 1.0,
                                             _p(eng.alpha_d), 3, _p(ym), _p(ys), _p(out), st) == -1
    # the largest one-split shape that fits is accepted: 174720 rows * 3 targets = 524160 doubles
    MPok = (_lib.STATIC_SCRATCH_DOUBLES // 3) // 128 * 128
    xq = torch.zeros((MPok, 4), dtype=torch.float64, device="cuda")
    big = torch.empty((MPok, 3), dtype=torch.float64, device="cuda")
    assert lib.gpk_predict_mean_fused_static(_p(xq), MPok, MPok, _p(eng.X_d), eng.N, 4, _lib.darr(inv_ls), 1.0,
                                             _p(eng.alpha_d), 3, _p(ym), _p(ys), _p(big), st) == 0
    torch.cuda.synchronize()
    assert bool(torch.isfinite(big).all())
