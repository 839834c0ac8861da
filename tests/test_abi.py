"""The C-ABI library loads and exports exactly what include/gpk.h declares (no compute calls here)."""
import ctypes
import os
import re

from unmanned_aerial_vehicles_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "gpk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gpk_[a-z_0-9]+)\s*\(", src)))


def test_header_matches_binding_table():
    assert header_symbols() == sorted(_lib.SYMBOLS)


def test_library_exports_every_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_pure_host_entry_points():
    lib = _lib.load()
    assert lib.gpk_abi_version() == 1
    assert lib.gpk_pad(1) == 128 and lib.gpk_pad(128) == 128 and lib.gpk_pad(129) == 256
    assert lib.gpk_pad(20000) == 20096
    assert lib.gpk_lml_out_len(9, 3) == 2 + 3 + 1 + 9 + 1
    assert lib.gpk_grad_workspace_doubles(1000, 10) == (1024 // 64) ** 2 * 12 + 1024   # D=9,10 -> padded to 10


def test_no_cpu_fallback_without_cuda():
    import numpy as np
    import pytest
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from sklearn.gaussian_process.kernels import RBF

    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, GpkUnavailable

    g = GaussianProcessRegressor(kernel=RBF(1.0), optimizer=None)
    with pytest.raises(GpkUnavailable):
        g.fit(np.random.rand(10, 2), np.random.rand(10))


def test_workspace_queries_are_pure_host():
    lib = _lib.load()
    NP, MP, D, T = 20096, 8192, 9, 3
    full = lib.gpk_predict_workspace_doubles(MP, NP, D, T)
    mean_only = lib.gpk_predict_mean_workspace_doubles(MP, 20000, D, T)
    assert full >= MP * NP + MP * D + 2 * MP * T and full >= mean_only      # one scratch serves both branches
    assert mean_only >= MP * D + 2 * MP * T and mean_only < full // 100      # no M x N term in the fused branch
    # few query rows: the training rows are split over enough CTAs to fill the GPU, one partial sum per split
    assert lib.gpk_predict_mean_workspace_doubles(128, 20000, D, T) >= 128 * D + 128 * T * (1 + 148)
    assert lib.gpk_predict_mean_workspace_doubles(100, 20000, D, T) == 0     # MP must be a multiple of 64


def test_stacked_predict_argument_block_is_marshalled_field_by_field():
    """gpk_predict_stacked validates every model's arguments before its first CUDA call, so the ctypes mirror of
    GpkStackedModel can be checked without a GPU: break one field at a time and read the code that field's check
    returns (same codes as gpk_predict)."""
    lib = _lib.load()
    fake = 0x1000                                                             # never dereferenced: validation only
    inv = _lib.darr([1.0] * 9)

    def model(**kw):
        arr = (_lib.StackedModel * 1)()
        m = arr[0]
        m.Xq_h, m.D, m.X_d, m.N = fake, 9, fake, 1000
        m.inv_ls_h, m.c, m.noise = ctypes.cast(inv, ctypes.c_void_p), 1.0, 0.1
        m.alpha_d, m.T, m.Mlow_d, m.NP = fake, 3, fake, 1024
        m.y_mean_d, m.y_std_d, m.mean_h, m.std_h, m.neg_h = fake, fake, fake, fake, fake
        m.ws_d, m.ws_doubles = fake, 0                                         # too little scratch: -19 when all else is fine
        for k, v in kw.items():
            setattr(m, k, v)
        return arr

    call = lambda arr, n=1, mq=25: lib.gpk_predict_stacked(arr, n, mq, None)
    assert call(None) == -1
    assert call(model(), n=0) == -2 and call(model(), n=_lib.MAX_STACKED + 1) == -2
    assert call(model(), mq=0) == -3
    assert call(model()) == -19                                               # every other field passed its check
    need = lib.gpk_predict_workspace_doubles(128, 1024, 9, 3)
    assert call(model(ws_doubles=need - 1)) == -19
    assert call(model(mean_h=None)) == -1 and call(model(Xq_h=None)) == -1 and call(model(ws_d=None)) == -1
    assert call(model(NP=1000)) == -12 and call(model(NP=896)) == -12         # not a multiple of 128 / smaller than N
    assert call(model(T=0)) == -10
    assert call(model(D=33, ws_doubles=1 << 40)) == -3                        # more than 32 features
    assert call(model(inv_ls_h=None, ws_doubles=1 << 40)) == -3


def test_null_and_shape_arguments_are_refused_before_any_launch():
    """include/gpk.h: a bad argument returns -k (k = its 1-based position) instead of launching a kernel on it -- a
    NULL device pointer would otherwise fault the CUDA context for the whole process, and the reference's callers
    only survive ordinary exceptions (simple_gp.py:199-201)."""
    lib = _lib.load()
    f, inv = 0x1000, _lib.darr([1.0] * 9)                                     # f: never dereferenced
    N, NP, D, T = 1000, 1024, 9, 3
    assert lib.gpk_gram(None, N, D, inv, 1.0, 0.1, 0.0, f, NP, 1, None) == -1
    assert lib.gpk_gram(f, N, D, None, 1.0, 0.1, 0.0, f, NP, 1, None) == -3
    assert lib.gpk_gram(f, N, D, inv, 1.0, 0.1, 0.0, None, NP, 1, None) == -8
    assert lib.gpk_gram(f, N, D, inv, 1.0, 0.1, 0.0, f, NP + 128, 1, None) == -9
    assert lib.gpk_potrs(f, f, NP, None, f, f, T, None) == -4
    assert lib.gpk_potrs(f, f, NP, f, None, f, T, None) == -5
    assert lib.gpk_potrs(f, f, NP, f, f, None, T, None) == -6
    ev = [f, N, D, inv, 1.0, 0.1, 1e-6, f, T, f, f, f, f, f, NP, 1, f, f, f, None]
    for pos in (1, 8, 10, 11, 12, 13, 14, 17, 18, 19):
        a = list(ev); a[pos - 1] = None
        assert lib.gpk_lml_eval(*a) == -pos, pos
    a = list(ev); a[1] = 0
    assert lib.gpk_lml_eval(*a) == -2
    a = list(ev); a[8] = 0
    assert lib.gpk_lml_eval(*a) == -9
    a = list(ev); a[14] = NP + 128
    assert lib.gpk_lml_eval(*a) == -15
    gr = [f, N, D, inv, 1.0, 0.1, f, f, f, T, f, NP, 1, f, f, None]
    for pos in (1, 7, 8, 9, 11, 14, 15):
        a = list(gr); a[pos - 1] = None
        assert lib.gpk_lml_grad(*a) == -pos, pos
    cg = [f, 25, f, N, D, inv, 1.0, f, 128, NP, None]
    for pos in (1, 3, 8):
        a = list(cg); a[pos - 1] = None
        assert lib.gpk_cross_gram(*a) == -pos, pos
    pm = [f, 25, N, NP, f, T, f, f, f, None]
    for pos in (1, 5, 7, 8, 9):
        a = list(pm); a[pos - 1] = None
        assert lib.gpk_predict_mean(*a) == -pos, pos
    ps = [f, 25, 128, f, NP, 1.1, f, T, f, f, f, None]
    for pos in (1, 4, 7, 9, 10, 11):
        a = list(ps); a[pos - 1] = None
        assert lib.gpk_predict_std(*a) == -pos, pos
    gm = [f, 256, f, 256, f, 256, 256, 256, 256, 1.0, 0, 0, 0, 0, None]
    for pos in (1, 3, 5):
        a = list(gm); a[pos - 1] = None
        assert lib.gpk_gemm_nt(*a) == -pos, pos
    for pos in (2, 4, 6):
        a = list(gm); a[pos - 1] = 128                                         # leading dimension shorter than the row
        assert lib.gpk_gemm_nt(*a) == -pos, pos
    a = list(gm); a[6] = 200
    assert lib.gpk_gemm_nt(*a) == -7
    assert lib.gpk_gemv(f, 128, 128, 128, None, T, f, 1.0, 0.0, 0, None) == -5
    assert lib.gpk_gemv(f, 128, 128, 128, f, T, None, 1.0, 0.0, 0, None) == -7
    assert lib.gpk_trimv(f, NP, None, f, T, 0, None) == -3 and lib.gpk_trimv(f, NP, f, None, T, 0, None) == -4
    assert lib.gpk_fix_diag(None, NP, NP, N, 0.1, 1.0, None) == -1
    assert lib.gpk_logdiag(None, NP, NP, f, None) == -1 and lib.gpk_logdiag(f, NP, NP, None, None) == -4
    assert lib.gpk_extract_lower(None, NP, f, N, None) == -1 and lib.gpk_extract_lower(f, NP, None, N, None) == -3
    assert lib.gpk_extract_lower(f, NP, f, NP + 1, None) == -4
    assert lib.gpk_pad_copy(None, 10, 10, f, 128, 128, None) == -1 and lib.gpk_pad_copy(f, 10, 10, None, 128, 128, None) == -4
    assert lib.gpk_dmma_probe(296, 10, None, None, None) == -3
