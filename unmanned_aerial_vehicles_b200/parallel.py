"""Multi-GPU sharding of the GP hot path (one process per GPU, torch.distributed; NCCL on GPUs, gloo in tests).

SURVEY.md section 8(e): the path shards over independent units, so there is NO collective in the data path:
  * optimiser restarts (x outputs): `_gpr.py:329-333` / `gp_trainer.py:140` loops are independent L-BFGS-B
    runs of UNEQUAL length -> a dynamic queue (a fetch-and-add counter in the rendezvous store: whichever rank
    finishes a run claims the next start), one all_gather_object of (theta*, -lml*) tuples, argmin everywhere;
  * query rows: `_gpr.py:446-500` has no cross-row term -> contiguous row blocks per rank, every rank holds
    the factor (recomputed from theta: 0.19 s at N=20k, cheaper than broadcasting 3.2 GB); the optional gather of
    the outputs is one `all_gather_into_tensor` per output array (device buffers under NCCL), nothing is pickled.
Starting points are drawn from the same seeded stream on every rank (identical to sklearn's serial draw) and the
results are merged in start order, so the sharded fit returns the serial fit's answer whoever ran which start.
Only hyper-parameters travel between ranks: a fitted estimator is rebuilt on the other ranks from (theta*, lml*)
and factors lazily on first use, so no N x N matrix is ever pickled.
An exception inside a work item is caught, gathered as a marker and re-raised on EVERY rank (a rank that raised
alone would leave the others blocked in the collective).
"""
from __future__ import annotations

import time
import traceback

import numpy as np

_SEQ = {}          # per-group call counter: every rank of a group calls run_sharded in the same order
last_stats = None  # per-rank bookkeeping of the most recent run_sharded on this rank (items run, busy/idle seconds)


def _dist():
    import torch.distributed as dist

    return dist if (dist.is_available() and dist.is_initialized()) else None


def rank_world(group=None):
    d = _dist()
    return (d.get_rank(group), d.get_world_size(group)) if d else (0, 1)


def shard_indices(n_items: int, rank: int, world: int):
    """round-robin ownership (item r -> rank r % world): the static schedule"""
    return list(range(rank, n_items, world))


class ShardedWorkError(RuntimeError):
    """A work item raised on some rank; raised on every rank of the group with the original traceback text."""


class _Failed:
    def __init__(self, rank, item, text):
        self.rank, self.item, self.text = rank, item, text


def _queue_key(group):
    d = _dist()
    ranks = tuple(d.get_process_group_ranks(group)) if group is not None else ("world",)
    n = _SEQ.get(ranks, 0)
    _SEQ[ranks] = n + 1
    return "gpk/queue/%s/%d" % ("-".join(map(str, ranks)), n)


def run_sharded(work_fn, items, group=None, schedule="dynamic", order=None):
    """Run work_fn(item) over the ranks of `group`; return the results of ALL items, in item order, on every rank
    (one all_gather_object of small host objects).

    order: optional permutation of range(len(items)) -- the order in which items are handed out (longest first when
    the caller can guess lengths); it must be THE SAME on every rank.  Results stay in item order whatever it is.

    schedule="dynamic": items are claimed first-come through an atomic counter in the rendezvous store, so a rank
    that drew short runs takes the leftovers instead of idling; "static": round-robin (item i -> rank i % world).
    Results do not depend on the schedule (each item is computed by exactly one rank, merged in item order)."""
    global last_stats
    d = _dist()
    rank, world = rank_world(group)
    items = list(items)
    order = list(range(len(items))) if order is None else [int(i) for i in order]
    if sorted(order) != list(range(len(items))):
        raise ValueError("order must be a permutation of the item indices")
    t_begin = time.perf_counter()
    mine, busy, failed = {}, 0.0, None

    def run(i):
        nonlocal busy, failed
        t0 = time.perf_counter()
        try:
            mine[i] = work_fn(items[i])
        except Exception:          # noqa: BLE001 -- re-raised on every rank below
            failed = _Failed(rank, i, traceback.format_exc())
        busy += time.perf_counter() - t0

    if world == 1:
        for i in order:
            run(i)
            if failed:
                raise ShardedWorkError(f"work item {failed.item} failed:\n{failed.text}")
        last_stats = {"rank": 0, "items": list(range(len(items))), "busy_s": busy, "idle_s": 0.0, "schedule": "serial"}
        return [mine[i] for i in range(len(items))]
    if schedule == "dynamic":
        from torch.distributed.distributed_c10d import _get_default_store

        store, key = _get_default_store(), _queue_key(group)
        while failed is None:
            i = int(store.add(key, 1)) - 1
            if i >= len(items):
                break
            run(order[i])
    else:
        for i in shard_indices(len(items), rank, world):
            run(order[i])
            if failed:
                break
    t_done = time.perf_counter()
    gathered = [None] * world
    d.all_gather_object(gathered, failed if failed is not None else mine, group=group)
    t_end = time.perf_counter()
    last_stats = {"rank": rank, "items": sorted(mine), "busy_s": busy, "idle_s": t_end - t_done,
                  "wall_s": t_end - t_begin, "schedule": schedule}
    bad = [p for p in gathered if isinstance(p, _Failed)]
    if bad:
        raise ShardedWorkError("; ".join(f"rank {b.rank}, work item {b.item}" for b in bad) + " failed:\n" + bad[0].text)
    merged = {}
    for part in gathered:
        merged.update(part)
    return [merged[i] for i in range(len(items))]


def gather_stats(group=None):
    """last_stats of every rank, on every rank (for bench / tools: per-rank run counts and idle time)"""
    d = _dist()
    rank, world = rank_world(group)
    if world == 1:
        return [last_stats]
    out = [None] * world
    d.all_gather_object(out, last_stats, group=group)
    return out


def pilot_order(estimator, X, y, pilot_rows, group=None, seed=0):
    """A claim order for the optimiser runs of ``estimator.fit(X, y)``: longest chain first, the lengths GUESSED by
    running the same fit (same kernel, same seeded starts) on a seeded subsample of ``pilot_rows`` rows -- a fit that
    costs about a second on the GPU at 2000 rows.  On the BASELINE config-2 data the pilot's three longest chains at 500
    and at 2000 rows contain the two 59-evaluation chains of the N=20 000 fit.  Rank 0 runs the
    pilot and broadcasts the order, so every rank holds the same one; anything that goes wrong (an unseeded estimator,
    whose pilot would draw other starts; an exception in the pilot) yields None = the plain queue order.  The order
    changes who runs what and when, never the result."""
    import numbers

    d = _dist()
    rank, world = rank_world(group)
    order = None
    if rank == 0:
        try:
            n = len(X)
            if isinstance(getattr(estimator, "random_state", None), numbers.Integral) and 0 < pilot_rows < n \
                    and getattr(estimator, "n_restarts_optimizer", 0) > 0:
                from sklearn.base import clone

                idx = np.sort(np.random.default_rng(seed).choice(n, int(pilot_rows), replace=False))
                pilot = clone(estimator)
                lens = []

                def serial(run_one, starts):
                    out = []
                    for t0 in starts:
                        k0 = pilot.n_lml_evals_
                        out.append(run_one(t0))
                        lens.append(pilot.n_lml_evals_ - k0)
                    return out
                pilot._restart_runner = serial
                import warnings
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    pilot.fit(np.asarray(X)[idx], np.asarray(y)[idx])
                pilot._engine = None
                order = [int(i) for i in np.argsort(-np.asarray(lens), kind="stable")]
        except Exception:          # noqa: BLE001 -- a heuristic: never in the way of the fit itself
            order = None
    if d is not None and world > 1:
        box = [order]
        d.broadcast_object_list(box, src=d.get_global_rank(group, 0) if group is not None else 0, group=group)
        order = box[0]
    return order


def fit_sharded_restarts(estimator, X, y, group=None, schedule="dynamic", pilot_rows=None):
    """estimator.fit(X, y) with its 1 + n_restarts_optimizer L-BFGS-B runs sharded over the ranks of `group`.
    Every rank ends with the same fitted estimator (same kernel_, L_, alpha_).  pilot_rows: hand the runs out longest
    first, lengths guessed by `pilot_order` (off by default)."""
    t0 = time.perf_counter()
    order = pilot_order(estimator, X, y, pilot_rows, group) if pilot_rows else None
    estimator.pilot_order_, estimator.pilot_seconds_ = order, time.perf_counter() - t0

    def runner(run_one, starts):
        use = order if (order is not None and len(order) == len(starts)) else None
        return run_sharded(run_one, starts, group, schedule, use)
    estimator._restart_runner = runner
    try:
        return estimator.fit(X, y)
    finally:
        del estimator._restart_runner


def _fit_state(est):
    """what has to travel for another rank to hold the same fitted estimator: hyper-parameters and the LML only"""
    if hasattr(est, "kernel_") and hasattr(est, "log_marginal_likelihood_value_"):
        return ("theta", np.array(est.kernel_.theta, dtype=np.float64), float(est.log_marginal_likelihood_value_))
    return ("object", est)            # duck-typed estimators (tests): travel whole


def _rebuild(make_estimator, t, state, X, y):
    if state[0] == "object":
        return state[1]
    return make_estimator(t).fit_from_theta(X, y, state[1], state[2])


def fit_sharded_outputs(make_estimator, X, Y, group=None, schedule="dynamic"):
    """The gp_trainer.py:140-176 loop (one independent single-output model per column of Y) sharded over ranks:
    work item = output column.  Returns the list of fitted estimators on every rank; only (theta*, lml*) crosses
    ranks, the non-owners rebuild the estimator from them (`fit_from_theta`: factorisation deferred to first use)."""
    cols = list(range(Y.shape[1]))
    kept = {}

    def fit_col(t):
        kept[t] = make_estimator(t).fit(X, Y[:, t])
        return _fit_state(kept[t])
    states = run_sharded(fit_col, cols, group, schedule)
    return [kept[t] if t in kept else _rebuild(make_estimator, t, states[t], X, Y[:, t]) for t in cols]


def fit_outputs_sharded_restarts(make_estimator, X, Y, group=None, release_device=True, schedule="dynamic"):
    """BASELINE config 4 (3 outputs x 32 restarts over 8 GPUs at N=50k): one single-output model per column of Y,
    fitted one after the other, each with its optimiser runs sharded over ALL ranks (32 runs / 8 ranks, claimed
    dynamically -- sharding the three outputs instead would leave five GPUs idle).  Every rank ends
    with every fitted estimator.  ``release_device`` drops each model's device buffers (three NP x NP matrices:
    60 GB at N=50k) once it is fitted, so the next output's factorisation fits beside it; predict re-creates them
    lazily from ``kernel_`` (one Gram + factorisation)."""
    models = []
    for t in range(Y.shape[1]):
        est = fit_sharded_restarts(make_estimator(t), X, Y[:, t], group, schedule)
        if release_device:
            est._engine = None
            est._low = None
        models.append(est)
    return models


def row_block(n_rows: int, rank: int, world: int):
    """contiguous [lo, hi) block of query rows owned by `rank`"""
    base, rem = divmod(n_rows, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _all_gather_rows(arr, n_rows, n_cols, group):
    """concatenate the ranks' row blocks of one float64 output: one all_gather_into_tensor of equal-sized (padded)
    blocks -- device buffers under NCCL, host tensors under gloo; no pickling"""
    import torch

    d = _dist()
    rank, world = rank_world(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if d.get_backend(group) == "nccl" else torch.device("cpu")
    blk = -(-n_rows // world)
    send = torch.zeros((blk, n_cols), dtype=torch.float64, device=dev)
    if arr is not None and len(arr):
        send[: len(arr)].copy_(torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64).reshape(len(arr), n_cols)))
    recv = torch.empty((world * blk, n_cols), dtype=torch.float64, device=dev)
    d.all_gather_into_tensor(recv, send, group=group)
    recv = recv.cpu().numpy().reshape(world, blk, n_cols)
    sizes = [hi - lo for lo, hi in (row_block(n_rows, r, world) for r in range(world))]
    return np.concatenate([recv[r, : sizes[r]] for r in range(world)])


def predict_sharded(estimator, Xq, return_std=False, gather=True, group=None):
    """predict() over this rank's row block; gather=True returns the full result on every rank
    (all_gather_into_tensor of the float64 blocks), gather=False leaves the outputs sharded."""
    rank, world = rank_world(group)
    lo, hi = row_block(len(Xq), rank, world)
    out = estimator.predict(Xq[lo:hi], return_std=return_std) if hi > lo else None
    if world == 1 or not gather:
        return out
    # output width: the estimator squeezes single-target results to 1-D (sklearn _gpr.py:453-454)
    import torch

    d = _dist()
    probe = out[0] if (return_std and out is not None) else out
    width = 0 if probe is None else (1 if probe.ndim == 1 else probe.shape[1])
    dev = torch.device("cuda", torch.cuda.current_device()) if d.get_backend(group) == "nccl" else torch.device("cpu")
    meta = torch.tensor([width, int(probe is not None and probe.ndim == 1)], dtype=torch.int64, device=dev)
    d.all_reduce(meta, op=d.ReduceOp.MAX, group=group)     # ranks with an empty block learn the shape from the others
    ncol, sq = int(meta[0]), bool(int(meta[1]))
    n = len(Xq)
    mean = _all_gather_rows(out[0] if (return_std and out is not None) else out, n, ncol, group)
    mean = mean[:, 0] if sq else mean
    if not return_std:
        return mean
    std = _all_gather_rows(out[1] if out is not None else None, n, ncol, group)
    return mean, (std[:, 0] if sq else std)
