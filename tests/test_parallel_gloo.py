"""World-size-2 gloo tests of the sharding logic (host side; the objective here is the CPU oracle because this
container has no GPU -- the sharding code is identical on NCCL)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


class Fake:
    """duck-typed estimator: the sharding helpers must not care what computes"""

    def __init__(self, t=None):
        self.t, self.seen = t, 0

    def fit(self, X, y):
        self.mu = float(y.mean())
        return self

    def predict(self, Xq, return_std=False):
        self.seen += len(Xq)
        m = Xq @ np.array([1.0, -2.0])
        return (m, np.abs(m) + 1.0) if return_std else m


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import scipy.optimize
    from oracle import gp_oracle as orc
    from unmanned_aerial_vehicles_b200 import parallel

    rng = np.random.default_rng(0)
    X = rng.standard_normal((60, 2)); y = np.sin(X[:, 0]) + 0.05 * rng.standard_normal(60)
    calls = []

    def run_one(t0):
        calls.append(tuple(t0))
        f = lambda th: (lambda l, g: (-l, -np.array([g["ls"].sum(), g["noise"]])))(
            *orc.lml_and_grad(X, y, np.exp(th[0]), 1.0, np.exp(th[1]), 1e-8))
        r = scipy.optimize.minimize(f, t0, method="L-BFGS-B", jac=True, bounds=[(-3, 3), (-8, 1)])
        return r.x, r.fun
    starts = [np.array([0.0, -2.0])] + [np.random.RandomState(1).uniform([-3, -8], [3, 1]) for _ in range(4)]
    res = parallel.run_sharded(run_one, starts)                      # dynamic queue: first come, first served
    n_dyn = len(calls)
    assert parallel.last_stats["schedule"] == "dynamic" and len(parallel.last_stats["items"]) == n_dyn
    res_static = parallel.run_sharded(run_one, starts, schedule="static")
    assert len(calls) - n_dyn == (3 if rank == 0 else 2)             # 5 runs round-robin over 2 ranks
    assert all(np.array_equal(a[0], b[0]) and a[1] == b[1] for a, b in zip(res, res_static))
    stats = parallel.gather_stats()
    assert sorted(i for st in stats for i in st["items"]) == [0, 1, 2, 3, 4]

    def boom(v):                                                     # a failure on ONE rank surfaces on EVERY rank
        if v == 3:
            raise np.linalg.LinAlgError("not positive definite")
        return v
    for sched in ("static", "dynamic"):
        try:
            parallel.run_sharded(boom, [0, 1, 2, 3], schedule=sched)
            raise AssertionError("expected ShardedWorkError")
        except parallel.ShardedWorkError as e:
            assert "LinAlgError" in str(e) and "work item 3" in str(e)
    lo, hi = parallel.row_block(11, rank, world)

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_parallel_gloo import Fake          # module-level: all_gather_object pickles the fitted models
    Xq = rng.standard_normal((11, 2))
    fk = Fake()
    m, sd = parallel.predict_sharded(fk, Xq, return_std=True)
    assert fk.seen == hi - lo and np.array_equal(m, Xq @ np.array([1.0, -2.0])) and np.array_equal(sd, np.abs(m) + 1.0)
    own = parallel.predict_sharded(Fake(), Xq, gather=False)
    assert np.array_equal(own, (Xq @ np.array([1.0, -2.0]))[lo:hi])

    class Fake2(Fake):                                               # multi-target outputs stay 2-D through the gather
        def predict(self, Xq, return_std=False):
            m = np.stack([Xq[:, 0], Xq[:, 1], Xq.sum(1)], 1)
            return (m, m * m) if return_std else m
    m2, s2 = parallel.predict_sharded(Fake2(), Xq, return_std=True)
    assert m2.shape == (11, 3) and np.array_equal(m2[:, 2], Xq.sum(1)) and np.array_equal(s2, m2 * m2)
    one = parallel.predict_sharded(Fake2(), Xq[:1])                  # rank 1's block is empty
    assert one.shape == (1, 3) and np.array_equal(one[0], [Xq[0, 0], Xq[0, 1], Xq[0].sum()])
    Y3 = np.stack([y, 2 * y, y + 1.0], 1)
    models = parallel.fit_sharded_outputs(lambda t: Fake(t), X, Y3)
    assert [mm.t for mm in models] == [0, 1, 2] and np.allclose([mm.mu for mm in models], Y3.mean(0))
    q.put((rank, n_dyn, [(list(map(float, a)), float(b)) for a, b in res], (lo, hi)))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_restarts_match_serial():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    outs = sorted(q.get(timeout=240) for _ in range(2))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    (r0, n0, res0, blk0), (r1, n1, res1, blk1) = outs
    assert n0 + n1 == 5                   # 5 runs claimed from the shared counter, each by exactly one rank
    assert res0 == res1 and len(res0) == 5  # every rank holds every result, in start order
    assert blk0 == (0, 6) and blk1 == (6, 11)
    best = int(np.argmin([f for _, f in res0]))
    assert np.isfinite(res0[best][1])


def test_shard_helpers_single_process():
    from unmanned_aerial_vehicles_b200 import parallel
    assert parallel.shard_indices(10, 1, 4) == [1, 5, 9]
    assert [parallel.row_block(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert parallel.run_sharded(lambda v: v * 2, [1, 2, 3]) == [2, 4, 6]


def _trainer_case():
    rng = np.random.default_rng(4)
    X = rng.standard_normal((70, 4)) * np.array([1.0, 3.0, 0.5, 2.0])
    Y = np.stack([np.sin(X[:, 0]), np.cos(X[:, 1] / 3), X[:, 2] * X[:, 3]], 1) + 0.05 * rng.standard_normal((70, 3))
    return X, Y


def _run_trainer(tmp):
    """callers.GPTrainer with the oracle-backed engine double (no GPU here); stats + one prediction per component"""
    import warnings
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_estimator_host_logic import OracleEngine
    from unmanned_aerial_vehicles_b200 import callers, gpr as gpr_mod
    real, gpr_mod.DeviceGP = gpr_mod.DeviceGP, OracleEngine
    try:
        X, Y = _trainer_case()
        tr = callers.GPTrainer(data_dir=os.path.join(tmp, "d"), model_dir=os.path.join(tmp, "m"), random_state=0)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            stats = tr.train_gp_models(X, Y)
        pg = callers.PerOutputGP(tr.gp_models, tr.scalers_X, tr.scalers_y, names=list(stats))
        mean, std = pg.predict_batch(X[:4] + 0.01)
    finally:
        gpr_mod.DeviceGP = real
    return {n: (s["log_marginal_likelihood"], s["rmse"], s["kernel"]) for n, s in stats.items()}, mean.tolist(), std.tolist()


def _trainer_worker(rank, world, port, q, tmp):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    q.put((rank,) + _run_trainer(os.path.join(tmp, str(rank))))
    dist.barrier()
    dist.destroy_process_group()


def test_trainer_components_sharded_over_two_ranks_equal_the_serial_loop(tmp_path):
    """gp_trainer.py:140-176: the per-component fits are independent; sharded over 2 ranks (component i -> rank i % 2,
    fitted models pickled to the other rank as host NumPy) every rank must hold the serial loop's models and scores."""
    sys.path.insert(0, ROOT)
    serial = _run_trainer(str(tmp_path / "serial"))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_trainer_worker, args=(r, 2, port, q, str(tmp_path))) for r in range(2)]
    [p.start() for p in procs]
    outs = sorted(q.get(timeout=300) for _ in range(2))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for out in outs:
        assert out[1].keys() == serial[0].keys() and len(serial[0]) == 3
        for n in serial[0]:
            assert out[1][n][2] == serial[0][n][2]                                   # same kernel_ string
            np.testing.assert_allclose(out[1][n][:2], serial[0][n][:2], rtol=1e-12)
        np.testing.assert_allclose(out[2], serial[1], rtol=1e-10, atol=1e-13)
        np.testing.assert_allclose(out[3], serial[2], rtol=1e-10, atol=1e-13)


def _run_config4_shape():
    """three single-output ARD models x 4 optimiser runs each through parallel.fit_outputs_sharded_restarts (engine
    double: no GPU here); returns theta*, LML and a prediction per model"""
    import warnings
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
    from test_estimator_host_logic import OracleEngine
    from unmanned_aerial_vehicles_b200 import gpr as gpr_mod, parallel
    real, gpr_mod.DeviceGP = gpr_mod.DeviceGP, OracleEngine
    try:
        X, Y = _trainer_case()
        k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 4, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
        make = lambda t: gpr_mod.GaussianProcessRegressor(kernel=k, alpha=1e-6, n_restarts_optimizer=3, random_state=10 + t)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            models = parallel.fit_outputs_sharded_restarts(make, X, Y)
            assert all(m._engine is None for m in models)                      # device buffers released per output
            out = [(m.kernel_.theta.tolist(), float(m.log_marginal_likelihood_value_),
                    np.asarray(m.predict(X[:3] + 0.01)).tolist()) for m in models]
    finally:
        gpr_mod.DeviceGP = real
    return out


def _config4_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    q.put((rank, _run_config4_shape()))
    dist.barrier()
    dist.destroy_process_group()


def test_outputs_in_turn_with_restarts_sharded_equal_the_serial_fits():
    """BASELINE config 4's sharding (outputs one after the other, each model's optimiser runs over all ranks): every
    rank ends with the serial fits' theta*, LML and predictions."""
    sys.path.insert(0, ROOT)
    serial = _run_config4_shape()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_config4_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    outs = sorted(q.get(timeout=300) for _ in range(2))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for _, got in outs:
        assert len(got) == 3
        for (th, lml, pred), (th0, lml0, pred0) in zip(got, serial):
            np.testing.assert_allclose(th, th0, rtol=0, atol=1e-12)
            assert lml == pytest.approx(lml0, rel=1e-12)
            np.testing.assert_allclose(pred, pred0, rtol=1e-10, atol=1e-13)


def _run_pilot_case(pilot_rows):
    """one ARD model, 5 optimiser runs, through parallel.fit_sharded_restarts with the pilot claim order (engine
    double: no GPU here); returns theta*, LML, the order used and a prediction"""
    import warnings
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
    from test_estimator_host_logic import OracleEngine
    from unmanned_aerial_vehicles_b200 import gpr as gpr_mod, parallel
    real, gpr_mod.DeviceGP = gpr_mod.DeviceGP, OracleEngine
    try:
        X, Y = _trainer_case()
        k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 4, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            g = gpr_mod.GaussianProcessRegressor(kernel=k, alpha=1e-6, n_restarts_optimizer=4, random_state=7)
            parallel.fit_sharded_restarts(g, X, Y[:, 0], pilot_rows=pilot_rows)
            unseeded = gpr_mod.GaussianProcessRegressor(kernel=k, alpha=1e-6, n_restarts_optimizer=1)
            assert parallel.pilot_order(unseeded, X, Y[:, 0], 40) is None          # its pilot would draw other starts
            out = (g.kernel_.theta.tolist(), float(g.log_marginal_likelihood_value_), g.pilot_order_,
                   np.asarray(g.predict(X[:3] + 0.01)).tolist(), list((parallel.last_stats or {}).get("items", [])))
    finally:
        gpr_mod.DeviceGP = real
    return out


def _pilot_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    q.put((rank, _run_pilot_case(40)))
    dist.barrier()
    dist.destroy_process_group()


def test_pilot_claim_order_changes_who_runs_what_never_the_result():
    """parallel.pilot_order: rank 0 guesses the chain lengths on a subsample and broadcasts the claim order (longest
    first); the sharded fit must still equal the serial fit exactly, every rank must hold the same order, and every
    run must be executed exactly once."""
    sys.path.insert(0, ROOT)
    serial = _run_pilot_case(None)                        # single process, no pilot: the reference result
    assert serial[2] is None
    single = _run_pilot_case(40)                          # single process with the pilot: same result, an order
    assert sorted(single[2]) == [0, 1, 2, 3, 4] and single[:2] == serial[:2] and single[3] == serial[3]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_pilot_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    outs = dict(q.get(timeout=300) for _ in range(2))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert outs[0][2] == outs[1][2] == single[2]          # the same order everywhere (rank 0's, broadcast)
    assert sorted(outs[0][4] + outs[1][4]) == [0, 1, 2, 3, 4]
    for r in (0, 1):
        np.testing.assert_allclose(outs[r][0], serial[0], rtol=0, atol=1e-12)
        assert outs[r][1] == pytest.approx(serial[1], rel=1e-12)
        np.testing.assert_allclose(outs[r][3], serial[3], rtol=1e-10, atol=1e-13)
    from unmanned_aerial_vehicles_b200 import parallel
    assert parallel.run_sharded(lambda v: v * 2, [1, 2, 3], order=[2, 0, 1]) == [2, 4, 6]     # results in ITEM order
    with pytest.raises(ValueError):
        parallel.run_sharded(lambda v: v, [1, 2, 3], order=[0, 0, 1])
