"""B200-native exact Gaussian-process residual model (drop-in for the sklearn estimator the
Grandediw/Unmanned_Aerial_Vehicles GP-MPC path uses).  See DESIGN.md."""
from ._lib import GpkError, GpkUnavailable, LIB_PATH  # noqa: F401
from .gpr import GaussianProcessRegressor  # noqa: F401

__all__ = ["GaussianProcessRegressor", "GpkError", "GpkUnavailable", "LIB_PATH"]
