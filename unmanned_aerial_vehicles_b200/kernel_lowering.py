"""Lower a scikit-learn kernel object to the flat parameter set libgpk evaluates.

Supported algebra (what the reference builds: src/px4/simple_gp.py:167 ``RBF(0.5)+WhiteKernel(0.1)``,
src/px4/gp_trainer.py:163-166 ``ConstantKernel(1,"fixed")*RBF([1]*D)+WhiteKernel(0.01)``):

    [ConstantKernel *] RBF(iso | ARD)  [+ WhiteKernel]        (operands in either order)

Anything else raises ``NotImplementedError`` -- there is deliberately no CPU fallback.
Ordering of theta follows sklearn (kernels.py:290-360): hyper-parameters of k1 before k2, fixed ones
dropped, all in log space.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
from sklearn.gaussian_process.kernels import RBF, ConstantKernel, Product, Sum, WhiteKernel


@dataclass
class Lowered:
    c: float
    length_scale: np.ndarray  # (D,)
    noise: float
    iso: bool
    # one entry per element of theta: ("c",0) | ("ls",d) | ("ls_iso",0) | ("noise",0)
    theta_map: list

    def with_theta(self, theta) -> "Lowered":
        """Same kernel structure at another theta (log space, sklearn's order) without touching sklearn's kernel
        objects: what ``lower(kernel.clone_with_theta(theta), D)`` returns, at a fraction of its host cost (the
        optimiser calls this once per objective evaluation; sklearn's theta setter walks ``dir()`` of every node)."""
        v = np.exp(np.asarray(theta, dtype=np.float64))
        if v.shape != (len(self.theta_map),):
            raise ValueError(f"theta has shape {v.shape}, kernel has {len(self.theta_map)} free hyper-parameters")
        c, noise, ls = self.c, self.noise, self.length_scale.copy()
        for k, (kind, d) in enumerate(self.theta_map):
            if kind == "c":
                c = float(v[k])
            elif kind == "noise":
                noise = float(v[k])
            elif kind == "ls":
                ls[d] = v[k]
            else:
                ls[:] = v[k]
        return Lowered(c=c, length_scale=ls, noise=noise, iso=self.iso, theta_map=self.theta_map)


def _split(kernel):
    """-> (const_kernel|None, rbf, white|None) or raise."""
    const = white = None
    k = kernel
    if isinstance(k, Sum):
        a, b = k.k1, k.k2
        if isinstance(a, WhiteKernel) and not isinstance(b, WhiteKernel):
            white, k = a, b
        elif isinstance(b, WhiteKernel):
            white, k = b, a
        else:
            raise NotImplementedError(f"unsupported kernel sum: {kernel}")
    if isinstance(k, Product):
        a, b = k.k1, k.k2
        if isinstance(a, ConstantKernel) and isinstance(b, RBF):
            const, k = a, b
        elif isinstance(b, ConstantKernel) and isinstance(a, RBF):
            const, k = b, a
        else:
            raise NotImplementedError(f"unsupported kernel product: {kernel}")
    if not isinstance(k, RBF) or type(k) is not RBF:
        raise NotImplementedError(
            f"libgpk evaluates [ConstantKernel*]RBF[+WhiteKernel] only, got {kernel!r} (no CPU fallback)")
    return const, k, white


def lower(kernel, n_features: int) -> Lowered:
    const, rbf, white = _split(kernel)
    ls = np.atleast_1d(np.asarray(rbf.length_scale, dtype=np.float64))
    iso = not rbf.anisotropic
    if iso:
        ls_full = np.full(n_features, float(ls[0]))
    else:
        if ls.shape[0] != n_features:
            raise ValueError(
                f"Anisotropic kernel must have the same number of dimensions as data ({ls.shape[0]}!={n_features})")
        ls_full = ls.copy()
    c = float(const.constant_value) if const is not None else 1.0
    noise = float(white.noise_level) if white is not None else 0.0
    theta_map = []
    for hp in kernel.hyperparameters:  # sklearn's own ordering
        if hp.fixed:
            continue
        if hp.name.endswith("constant_value"):
            theta_map.append(("c", 0))
        elif hp.name.endswith("noise_level"):
            theta_map.append(("noise", 0))
        elif hp.name.endswith("length_scale"):
            if iso:
                theta_map.append(("ls_iso", 0))
            else:
                theta_map.extend(("ls", d) for d in range(n_features))
        else:  # pragma: no cover
            raise NotImplementedError(hp.name)
    if len(theta_map) != kernel.n_dims:
        raise NotImplementedError(f"could not map theta of {kernel!r}")
    return Lowered(c=c, length_scale=ls_full, noise=noise, iso=iso, theta_map=theta_map)


def assemble_gradient(low: Lowered, g_c: float, g_ls: np.ndarray, g_noise: float) -> np.ndarray:
    """device outputs -> d lml / d theta in sklearn's order"""
    out = np.empty(len(low.theta_map))
    for k, (kind, d) in enumerate(low.theta_map):
        if kind == "c":
            out[k] = g_c
        elif kind == "noise":
            out[k] = g_noise
        elif kind == "ls":
            out[k] = g_ls[d]
        else:
            out[k] = float(np.sum(g_ls))
    return out
