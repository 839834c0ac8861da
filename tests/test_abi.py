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
    assert lib.gpk_grad_workspace_doubles(1000, 10) == (1024 // 64) ** 2 * 14 + 1024


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


def test_int8_opt_in_registration_is_pure_host_and_off_by_default():
    lib = _lib.load()
    assert lib.gpk_i8_mode() == 0                                             # nothing enabled unless asked
    al = lambda b: (b + 255) & ~255
    NP, MP, S = 20096, 16384, 7
    want = 2 * (al(S * NP * NP) + al(8 * NP)) + al(S * MP * NP) + al(8 * MP)
    assert lib.gpk_i8_workspace_bytes(NP, MP, S) == want
    assert lib.gpk_i8_workspace_bytes(NP + 64, MP, S) == 0                    # NP must be a multiple of 128
    assert lib.gpk_i8_workspace_bytes(NP, MP, 9) == 0                         # at most 8 slices (512 TMEM columns)
    fake = ctypes.c_void_p(0x1000)                                            # never dereferenced by configure
    assert lib.gpk_i8_configure(fake, want - 1, NP, MP, S, 3) == -2 and lib.gpk_i8_mode() == 0
    assert lib.gpk_i8_configure(fake, want, NP, MP, 8, 3) != 0 and lib.gpk_i8_mode() == 0
    assert lib.gpk_i8_configure(fake, want, NP, MP, S, 3) == 0 and lib.gpk_i8_mode() == 3
    assert lib.gpk_i8_configure(None, 0, 0, 0, 0, 0) == 0 and lib.gpk_i8_mode() == 0
