"""Node-range sharding of a fleet across ranks (one process per GPU).

Every per-node output depends only on that node's row, the pods scheduled onto it and its series,
so rank g owns the contiguous node range [g*N/G, (g+1)*N/G): its pods are routed by node id, pods
without a node go round-robin (they only touch fleet totals), per-pod outputs stay with the owner.
The single exchange step is the all-reduce of the fleet totals vector; inside libigpu_rollup.so it
runs over NCCL right after the kernels (igpu_comm_init), and `allreduce_totals` below is the same
reduction through torch.distributed for host-side tests (gloo) and for callers without NCCL.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

from . import columns as K


def node_range(n_nodes: int, rank: int, world: int) -> Tuple[int, int]:
    """Equal node counts per rank: right for the series (every node carries the same number of samples)."""
    return (rank * n_nodes) // world, ((rank + 1) * n_nodes) // world


def balanced_node_ranges(pods_per_node: np.ndarray, world: int) -> np.ndarray:
    """Contiguous node ranges with (nearly) equal POD counts: bounds[g] .. bounds[g + 1] is rank g's range.

    Pods are not spread evenly over nodes (a few busy nodes carry most of them); ranges of equal node count then leave
    one rank with a multiple of the others' pods, and the rollup takes as long as that rank.  The cut points are where
    the running pod count passes g / world of the total; a node is never split (its pods must meet on one GPU)."""
    cum = np.concatenate([[0], np.cumsum(np.asarray(pods_per_node, np.int64))])
    targets = cum[-1] * np.arange(1, world, dtype=np.float64) / world
    inner = np.searchsorted(cum, targets, side="left")
    bounds = np.concatenate([[0], inner, [len(cum) - 1]]).astype(np.int64)
    return np.maximum.accumulate(np.minimum(bounds, len(cum) - 1))


def shard_fleet(cols: K.FleetColumns, rank: int, world: int, bounds=None):
    """Returns (shard columns, pod_rows) where pod_rows[i] is the fleet-wide row of shard pod i.  `bounds` (world + 1 node
    indices, e.g. from balanced_node_ranges) replaces the equal-node-count ranges."""
    lo, hi = (int(bounds[rank]), int(bounds[rank + 1])) if bounds is not None else node_range(cols.n_nodes, rank, world)
    pn = cols.pod_node
    unsched = pn == K.NODE_NONE
    mine = (~unsched & (pn >= lo) & (pn < hi)) | (unsched & (np.arange(cols.n_pods) % world == rank))
    rows = np.nonzero(mine)[0]
    counts = np.diff(cols.pod_entry_off.astype(np.int64))[rows]
    off = np.zeros(len(rows) + 1, np.uint32)
    np.cumsum(counts, out=off[1:])
    # gather each kept pod's entry run
    starts = cols.pod_entry_off[rows].astype(np.int64)
    idx = np.repeat(starts - off[:-1].astype(np.int64), counts) + np.arange(int(off[-1]), dtype=np.int64)
    entries = cols.entries[idx].copy()
    entries["pod"] = np.repeat(np.arange(len(rows), dtype=np.uint32), counts)
    new_node = np.where(unsched[rows], np.int64(K.NODE_NONE), pn[rows].astype(np.int64) - lo).astype(np.uint32)
    shard = K.FleetColumns(
        node_flags=cols.node_flags[lo:hi],
        node_cap_i915=cols.node_cap_i915[lo:hi],
        node_cap_xe=cols.node_cap_xe[lo:hi],
        node_alloc_i915=cols.node_alloc_i915[lo:hi],
        node_alloc_xe=cols.node_alloc_xe[lo:hi],
        node_alloc_own=cols.node_alloc_own[lo:hi],
        pod_node=new_node,
        pod_phase=cols.pod_phase[rows],
        pod_entry_off=off,
        entries=entries,
        n_resources=cols.n_resources,
        options=cols.options,
        resource_names=cols.resource_names,
        node_names=cols.node_names[lo:hi] if cols.node_names else None,
    )
    return shard, rows


def finalize_totals(totals: np.ndarray) -> np.ndarray:
    """Recomputes the non-additive slots after totals were summed over shards
    (OverviewPage.tsx:115-116: Math.round(allocated / capacity * 100), 0 unless capacity > 0)."""
    t = totals.copy()
    cap_ok = t[K.T_CAP_NAN] == 0 and t[K.T_CAPACITY] > 0
    if cap_ok:
        x = float(t[K.T_ALLOCATED]) / float(t[K.T_CAPACITY]) * 100.0
        f = np.floor(x)
        t[K.T_UTIL_PCT] = int(f + 1 if x - f >= 0.5 else f)
    else:
        t[K.T_UTIL_PCT] = 0
    return t


def allreduce_totals(local_totals: np.ndarray) -> np.ndarray:
    """Sum of the additive totals over all ranks via torch.distributed (any backend), then finalize."""
    import torch
    import torch.distributed as dist

    t = torch.from_numpy(local_totals.astype(np.int64).copy())
    for s in K.T_NON_ADDITIVE:
        t[s] = 0
    if dist.is_initialized() and dist.get_world_size() > 1:
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        t = t.cpu()
    return finalize_totals(t.numpy())


def merge_fleet_vectors(vectors) -> np.ndarray:
    """Combines per-rank series fleet vectors [sum, count, min, max, nodes, ...]."""
    v = np.array(vectors, dtype=np.float64)
    out = np.zeros(K.FLEET_LEN)
    out[K.F_SUM] = v[:, K.F_SUM].sum()
    out[K.F_COUNT] = v[:, K.F_COUNT].sum()
    out[K.F_NODES] = v[:, K.F_NODES].sum()
    has = v[:, K.F_NODES] > 0
    out[K.F_MIN] = v[has, K.F_MIN].min() if has.any() else np.nan
    out[K.F_MAX] = v[has, K.F_MAX].max() if has.any() else np.nan
    return out
