"""Multi-GPU sharding of the GP hot path (one process per GPU, torch.distributed; NCCL on GPUs, gloo in tests).

SURVEY.md section 8(e): the path shards over independent units, so there is NO collective in the data path:
  * optimiser restarts (x outputs): `_gpr.py:329-333` / `gp_trainer.py:140` loops are independent L-BFGS-B
    runs -> round-robin over ranks, one all_gather_object of (theta*, -lml*) tuples, argmin everywhere;
  * query rows: `_gpr.py:446-500` has no cross-row term -> contiguous row blocks per rank, every rank holds
    the factor (recomputed from theta: 0.24 s at N=20k, cheaper than broadcasting 3.2 GB).
Starting points are drawn from the same seeded stream on every rank (identical to sklearn's serial draw), so
the sharded fit returns the serial fit's answer.
"""
from __future__ import annotations

import numpy as np


def _dist():
    import torch.distributed as dist

    return dist if (dist.is_available() and dist.is_initialized()) else None


def rank_world(group=None):
    d = _dist()
    return (d.get_rank(group), d.get_world_size(group)) if d else (0, 1)


def shard_indices(n_items: int, rank: int, world: int):
    """round-robin ownership (restart r -> rank r % world)"""
    return list(range(rank, n_items, world))


def run_sharded(work_fn, items, group=None):
    """Run work_fn(item) for the items this rank owns; return the results of ALL items, in item order, on every
    rank (one all_gather_object; results must be picklable host objects)."""
    rank, world = rank_world(group)
    mine = {i: work_fn(items[i]) for i in shard_indices(len(items), rank, world)}
    if world == 1:
        return [mine[i] for i in range(len(items))]
    gathered = [None] * world
    _dist().all_gather_object(gathered, mine, group=group)
    merged = {}
    for part in gathered:
        merged.update(part)
    return [merged[i] for i in range(len(items))]


def fit_sharded_restarts(estimator, X, y, group=None):
    """estimator.fit(X, y) with its 1 + n_restarts_optimizer L-BFGS-B runs sharded over the ranks of `group`.
    Every rank ends with the same fitted estimator (same kernel_, L_, alpha_)."""
    estimator._restart_runner = lambda run_one, starts: run_sharded(run_one, starts, group)
    try:
        return estimator.fit(X, y)
    finally:
        del estimator._restart_runner


def fit_sharded_outputs(make_estimator, X, Y, group=None):
    """The gp_trainer.py:140-176 loop (one independent single-output model per column of Y) sharded over ranks:
    work item = output column.  Returns the list of fitted estimators (pickled across ranks: host NumPy)."""
    cols = list(range(Y.shape[1]))
    return run_sharded(lambda t: make_estimator(t).fit(X, Y[:, t]), cols, group)


def fit_outputs_sharded_restarts(make_estimator, X, Y, group=None, release_device=True):
    """BASELINE config 4 (3 outputs x 32 restarts over 8 GPUs at N=50k): one single-output model per column of Y,
    fitted one after the other, each with its optimiser runs sharded over ALL ranks (32 runs / 8 ranks = 4 each, so
    12 runs per GPU in total -- sharding the three outputs instead would leave five GPUs idle).  Every rank ends
    with every fitted estimator.  ``release_device`` drops each model's device buffers (three NP x NP matrices:
    60 GB at N=50k) once it is fitted, so the next output's factorisation fits beside it; predict re-creates them
    lazily from ``kernel_`` (one Gram + factorisation)."""
    models = []
    for t in range(Y.shape[1]):
        est = fit_sharded_restarts(make_estimator(t), X, Y[:, t], group)
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


def predict_sharded(estimator, Xq, return_std=False, gather=True, group=None):
    """predict() over this rank's row block; gather=True returns the full result on every rank."""
    rank, world = rank_world(group)
    lo, hi = row_block(len(Xq), rank, world)
    out = estimator.predict(Xq[lo:hi], return_std=return_std) if hi > lo else None
    if world == 1 or not gather:
        return out
    parts = [None] * world
    _dist().all_gather_object(parts, out, group=group)
    parts = [p for p in parts if p is not None]
    if return_std:
        return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts])
    return np.concatenate(parts)
