"""The sharded paths on real GPUs over NCCL (skipped unless the box has at least two): what tests/test_parallel_gloo.py
and tests/test_dist_chol.py prove on CPU with gloo and an oracle-backed engine, here through libgpk.so --
restart sharding from the shared queue == the serial fit, row-sharded predict with the device-side gather == the
single-GPU predict, and the 2-D block-cyclic factorisation (grids 1x2 and 2x1) == the single-GPU engine."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _data(n, d, t, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    Y = np.sin(X @ (rng.standard_normal((d, t)) / np.sqrt(d))) + 0.1 * rng.standard_normal((n, t))
    return X, Y


def _kernel(d):
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
    return CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * d, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))


def _worker(rank, world, port, q):
    import warnings
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, parallel
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    from unmanned_aerial_vehicles_b200.engine import DeviceGP

    out = {}
    warnings.simplefilter("ignore")
    # (1) optimiser runs claimed from the shared queue: every rank ends with the serial fit
    X, Y = _data(600, 4, 2, seed=1)
    mk = lambda: GaussianProcessRegressor(kernel=_kernel(4), alpha=1e-6, n_restarts_optimizer=4, random_state=3)
    serial = mk().fit(X, Y)
    sharded = parallel.fit_sharded_restarts(mk(), X, Y)
    stats = parallel.gather_stats()
    out["theta_equal"] = bool(np.array_equal(serial.kernel_.theta, sharded.kernel_.theta))
    out["lml_equal"] = serial.log_marginal_likelihood_value_ == sharded.log_marginal_likelihood_value_
    out["runs"] = sorted(i for st in stats for i in st["items"])
    out["evals_here"] = int(sharded.n_lml_evals_)
    # (2) per-output models sharded over ranks: only (theta, lml) travels, the other rank rebuilds and factors lazily
    models = parallel.fit_sharded_outputs(lambda t: GaussianProcessRegressor(kernel=_kernel(4), alpha=1e-6,
                                                                             n_restarts_optimizer=1, random_state=t),
                                          X, Y)
    want = [GaussianProcessRegressor(kernel=_kernel(4), alpha=1e-6, n_restarts_optimizer=1, random_state=t).fit(X, Y[:, t])
            for t in range(2)]
    out["outputs_equal"] = all(np.array_equal(a.kernel_.theta, b.kernel_.theta) and
                               np.array_equal(a.predict(X[:7] + 0.1), b.predict(X[:7] + 0.1)) for a, b in zip(models, want))
    # (3) query rows sharded, outputs gathered with all_gather_into_tensor on device buffers
    Xq = np.random.default_rng(5).standard_normal((1001, 4))
    m1, s1 = serial.predict(Xq, return_std=True)
    m2, s2 = parallel.predict_sharded(serial, Xq, return_std=True)
    out["predict_mean_equal"] = bool(np.allclose(m1, m2, rtol=0, atol=1e-12))   # block heights differ -> summation order
    out["predict_std_equal"] = bool(np.allclose(s1, s2, rtol=0, atol=1e-12))
    own = parallel.predict_sharded(serial, Xq, gather=False)
    lo, hi = parallel.row_block(len(Xq), rank, world)
    out["own_block"] = bool(np.allclose(own, serial.predict(Xq)[lo:hi], rtol=0, atol=1e-12))
    # (4) a failure on one rank surfaces on both
    def boom(v):
        if v == 1:
            raise np.linalg.LinAlgError("boom")
        return v
    try:
        parallel.run_sharded(boom, [0, 1, 2])
        out["error_propagates"] = False
    except parallel.ShardedWorkError as e:
        out["error_propagates"] = "LinAlgError" in str(e)
    # (5) block-cyclic Cholesky over NCCL broadcasts against the single-GPU engine, both 2-rank grids
    Xd, Yd = _data(3000, 9, 3, seed=2)
    inv_ls, noise = 1.0 / np.linspace(0.8, 2.0, 9), 0.02
    eng = DeviceGP(Xd, Yd, 1e-6)
    ref_lml, _, gls_ref, gn_ref = eng.lml(inv_ls, 1.0, noise, True)
    assert eng.factor(inv_ls, 1.0, noise) == 0
    a_ref = eng.alpha_host()
    Xq9 = np.random.default_rng(6).standard_normal((300, 9))
    mr, sr, _, _ = eng.predict(Xq9, 0.0, 1.0, True)
    for grid in ((1, 2), (2, 1)):
        gp = DistGP(Xd, Yd, 1e-6, NB=512, grid=grid)
        gp.ops.trace = []
        assert gp.factor(inv_ls, 1.0, noise) == 0
        gp.solve()
        lml = gp.lml()
        md, sd = gp.predict(Xq9, return_std=True)
        lml_g, g_c, g_ls, g_noise = gp.lml_grad(inv_ls, 1.0, noise)         # distributed TRTRI + LAUUM + gradient
        phases = gp.ops.trace_summary()
        out[f"dist_{grid[0]}x{grid[1]}"] = {
            "lml_rel": abs(lml - ref_lml) / abs(ref_lml), "alpha_err": float(np.abs(gp.alpha_host() - a_ref).max()),
            "mean_err": float(np.abs(md - mr).max()), "std_err": float(np.abs(sd - sr).max()),
            "bcast_bytes": int(gp.ops.bcast_bytes), "phases": sorted(phases),
            "grad_err": float(max(np.max(np.abs(g_ls - gls_ref)) / np.max(np.abs(gls_ref)), abs(g_noise - gn_ref) / abs(gn_ref),
                                  abs(lml_g - ref_lml) / abs(ref_lml)))}
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs (NCCL)")
def test_sharded_paths_over_nccl_on_two_gpus():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    outs = dict(q.get(timeout=900) for _ in range(2))
    [p.join(120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank in (0, 1):
        o = outs[rank]
        assert o["theta_equal"] and o["lml_equal"] and o["runs"] == [0, 1, 2, 3, 4], o
        assert o["outputs_equal"] and o["predict_mean_equal"] and o["predict_std_equal"] and o["own_block"], o
        assert o["error_propagates"], o
        for g in ("dist_1x2", "dist_2x1"):
            d = o[g]
            assert d["lml_rel"] < 1e-11 and d["alpha_err"] < 1e-8 and d["mean_err"] < 1e-9 and d["std_err"] < 1e-8, d
            assert d["bcast_bytes"] > 0 and "update_trailing" in d["phases"] and "diag_potrf_inv" in d["phases"], d
            assert d["grad_err"] < 1e-7 and "lauum_update" in d["phases"] and "trtri_update" in d["phases"], d
    assert outs[0]["evals_here"] + outs[1]["evals_here"] > 10
