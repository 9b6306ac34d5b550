"""CPU suite, world_size 2 over gloo: node-range sharding + all-reduce of the fleet totals gives the
unsharded answer (the N>1 path of SURVEY §8(e) without a GPU; the oracle does the per-rank compute)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist

    from headlamp_intel_gpu_plugin_b200 import columns as K
    from headlamp_intel_gpu_plugin_b200 import shard, synth
    from oracle import columnar as oracle

    dist.init_process_group("gloo", rank=rank, world_size=world)
    cols = synth.synth_fleet(301, 9000, seed=77)  # odd node count: uneven ranges
    mine, rows = shard.shard_fleet(cols, rank, world)
    mine.validate()
    out = oracle.alloc_rollup(mine)
    totals = shard.allreduce_totals(out.fleet_totals)
    # series fleet vector: sharded by node range too
    x = synth.synth_series(301 * 2, 500, seed=3)
    lo, hi = shard.node_range(301, rank, world)
    off = synth.even_node_offsets(hi - lo, 2)
    _, _, fleet = oracle.quantile_by_node(x[lo * 2:hi * 2], off, 0.95)
    gathered = [None] * world
    dist.all_gather_object(gathered, fleet)
    merged = shard.merge_fleet_vectors(gathered)
    if rank == 0:
        full = oracle.alloc_rollup(cols)
        _, _, ffleet = oracle.quantile_by_node(x, synth.even_node_offsets(301, 2), 0.95)
        q.put((np.array_equal(totals, full.fleet_totals), np.array_equal(out.pod_req, full.pod_req[:, rows]),
               np.array_equal(out.node_pod_count, full.node_pod_count[lo:hi]),
               bool(np.allclose(merged[:5], ffleet[:5], rtol=1e-12)), int(totals[K.T_GPU_PODS])))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_over_gloo():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[:4] == (True, True, True, True), res
    assert res[4] > 0


def test_shard_partition_is_exact():
    from headlamp_intel_gpu_plugin_b200 import shard, synth

    cols = synth.synth_fleet(100, 3000, seed=5)
    seen = np.zeros(cols.n_pods, int)
    for world in (1, 3, 8):
        seen[:] = 0
        n_nodes = 0
        for r in range(world):
            s, rows = shard.shard_fleet(cols, r, world)
            seen[rows] += 1
            n_nodes += s.n_nodes
        assert (seen == 1).all() and n_nodes == cols.n_nodes


@pytest.mark.parametrize("cap,alloc,nan,want", [(100, 250, 0, 250), (0, 5, 0, 0), (10, 5, 1, 0), (3, 1, 0, 33)])
def test_finalize_totals_rounding(cap, alloc, nan, want):
    from headlamp_intel_gpu_plugin_b200 import columns as K
    from headlamp_intel_gpu_plugin_b200 import shard

    t = np.zeros(K.TOTALS_LEN, np.int64)
    t[K.T_CAPACITY], t[K.T_ALLOCATED], t[K.T_CAP_NAN] = cap, alloc, nan
    assert shard.finalize_totals(t)[K.T_UTIL_PCT] == want


@pytest.mark.parametrize("world", [2, 3, 8])
def test_ranged_generator_with_unscheduled_share_equals_shard_of_whole_fleet(world):
    """bench.py builds every rank's C3 shard directly (node range + its round-robin share of the pods
    without a node); that must be the shard `shard_fleet` cuts out of the whole fleet."""
    sys.path.insert(0, ROOT)
    from headlamp_intel_gpu_plugin_b200 import shard, synth

    n_nodes, n_pods = 203, 7000
    whole = synth.synth_fleet(n_nodes, n_pods, seed=5)
    for rank in range(world):
        want, _ = shard.shard_fleet(whole, rank, world)
        got = synth.synth_fleet(n_nodes, n_pods, seed=5, node_range=shard.node_range(n_nodes, rank, world), unscheduled_share=(rank, world))
        for f in "node_flags node_cap_i915 node_cap_xe node_alloc_i915 node_alloc_xe node_alloc_own pod_node pod_phase pod_entry_off entries".split():
            assert np.array_equal(getattr(got, f), getattr(want, f)), (rank, f)


def test_balanced_node_ranges_even_out_the_pods_and_lose_nothing(oracle_lib):
    """Node ranges of equal POD count (shard.balanced_node_ranges): a partition of the nodes, much flatter than equal node
    counts on the generator's skewed fleet, and the shards' totals still add up to the whole fleet's."""
    sys.path.insert(0, ROOT)
    from headlamp_intel_gpu_plugin_b200 import columns as K
    from headlamp_intel_gpu_plugin_b200 import shard, synth

    n_nodes, n_pods, world = 400, 30_000, 8
    whole = synth.synth_fleet(n_nodes, n_pods, seed=9)
    hist = synth.pods_per_node(n_pods, n_nodes, seed=9)
    assert np.array_equal(hist, np.bincount(whole.pod_node[whole.pod_node != K.NODE_NONE], minlength=n_nodes))
    bounds = shard.balanced_node_ranges(hist, world)
    assert bounds[0] == 0 and bounds[-1] == n_nodes and np.all(np.diff(bounds) >= 0)
    per_rank = [int(hist[bounds[g]:bounds[g + 1]].sum()) for g in range(world)]
    even = [int(hist[g * n_nodes // world:(g + 1) * n_nodes // world].sum()) for g in range(world)]
    assert max(per_rank) < 1.15 * np.mean(per_rank) and max(even) > 2 * np.mean(even)
    want = oracle_lib.alloc_rollup(whole)
    totals = np.zeros(K.TOTALS_LEN, np.int64)
    for g in range(world):
        mine, rows = shard.shard_fleet(whole, g, world, bounds=bounds)
        mine.validate()
        got = oracle_lib.alloc_rollup(mine)
        t = got.fleet_totals.copy()
        for s_ in K.T_NON_ADDITIVE:
            t[s_] = 0
        totals += t
        assert np.array_equal(got.pod_req, want.pod_req[:, rows])
        assert np.array_equal(got.node_pod_count, want.node_pod_count[bounds[g]:bounds[g + 1]])
        # the generator can build the same shard directly
        direct = synth.synth_fleet(n_nodes, n_pods, seed=9, node_range=(int(bounds[g]), int(bounds[g + 1])), unscheduled_share=(g, world))
        assert np.array_equal(direct.entries, mine.entries) and np.array_equal(direct.pod_node, mine.pod_node)
    assert np.array_equal(shard.finalize_totals(totals), want.fleet_totals)
    # degenerate inputs: all pods on one node, no pods at all
    assert shard.balanced_node_ranges(np.array([5, 0, 0, 0]), 3).tolist() == [0, 1, 1, 4]
    assert shard.balanced_node_ranges(np.zeros(4, int), 2).tolist()[0] == 0
