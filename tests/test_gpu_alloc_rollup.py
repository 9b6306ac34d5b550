"""GPU parity (through the C ABI): allocation rollup + classification versus the oracle.
Bit-exact bar: every integer / class / flag column must be identical."""
import json
import os

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import k8s, native, synth

from .dirty import dirty_fleet
from .parity import assert_outputs_equal, check_against_objects

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
VECTORS = json.load(open(os.path.join(HERE, "golden", "k8s_test_vectors.json")))["vectors"]

FN = {
    "isIntelGpuNode": k8s.isIntelGpuNode,
    "filterIntelGpuNodes": lambda items: [n["metadata"]["name"] for n in k8s.filterIntelGpuNodes(items)],
    "getNodeGpuType": k8s.getNodeGpuType,
    "getNodeGpuCount": k8s.getNodeGpuCount,
    "isNodeReady": k8s.isNodeReady,
    "isGpuRequestingPod": k8s.isGpuRequestingPod,
    "filterGpuRequestingPods": lambda items: [p["metadata"]["name"] for p in k8s.filterGpuRequestingPods(items)],
    "getPodGpuRequests": k8s.getPodGpuRequests,
}


@pytest.mark.parametrize("v", VECTORS, ids=[f'{v["fn"]}@{v["ref"].split(":")[-1]}' for v in VECTORS])
def test_reference_known_answers_through_cuda(ctx, v):
    """The reference's own k8s.test.ts cases, answered by the frozen-signature helpers on the GPU."""
    assert FN[v["fn"]](v["input"]) == v["expect"], v["ref"]


@pytest.mark.parametrize("n_nodes,n_pods,seed,plain_bp", [(50, 2000, 20261106, 0), (7, 300, 5, 3000), (1, 40, 9, 0), (300, 1, 2, 500)])
def test_generated_fleet_matches_oracles(ctx, oracle_lib, n_nodes, n_pods, seed, plain_bp):
    nodes, pods = synth.fleet_objects(n_nodes, n_pods, seed, non_gpu_share_bp=plain_bp)
    cols = k8s.marshalFleet(nodes, pods)
    out = ctx.alloc_rollup(cols)
    assert_outputs_equal(out, oracle_lib.alloc_rollup(cols), "cuda vs columnar oracle")
    check_against_objects(nodes, pods, cols, out)
    r = k8s.rollupFleet(nodes, pods, ctx, container_rows=True)  # the staged form + igpu_fleet_container_rows
    assert np.array_equal(r.container_rows, oracle_lib.container_rows(r.cols))
    check_against_objects(nodes, pods, r.cols, r.out, r.order, r.pos, container_rows=r.container_rows)


@pytest.mark.parametrize("seed", [7, 8, 9, 10, 11])
def test_dirty_fleet_matches_oracles(ctx, oracle_lib, seed):
    nodes, pods = dirty_fleet(40, 400, seed)
    cols = k8s.marshalFleet(nodes, pods)  # > 8 resource ids: exercises the wide-R kernel configuration
    out = ctx.alloc_rollup(cols)
    assert_outputs_equal(out, oracle_lib.alloc_rollup(cols), "cuda vs columnar oracle")
    check_against_objects(nodes, pods, cols, out)
    for grouped in (False, True):  # per-container key rows (PodsPage.tsx:49-88, PodDetailSection.tsx:48-83) off the device
        r = k8s.rollupFleet(nodes, pods, ctx, group_by_node=grouped, container_rows=True)
        assert np.array_equal(r.container_rows, oracle_lib.container_rows(r.cols))
        check_against_objects(nodes, pods, r.cols, r.out, r.order, r.pos, container_rows=r.container_rows)


def test_empty_and_ragged(ctx, oracle_lib):
    for nodes, pods in ([], []), (synth.fleet_objects(3, 0)[0], []), ([], synth.fleet_objects(3, 25)[1]):
        cols = k8s.marshalFleet(nodes, pods)
        assert_outputs_equal(ctx.alloc_rollup(cols), oracle_lib.alloc_rollup(cols))
    # pods with zero entries interleaved with heavy pods (ragged CSR)
    cols = synth.synth_fleet(20, 5000, seed=1)
    keep = np.ones(cols.n_entries, bool)
    starts, ends = cols.pod_entry_off[:-1], cols.pod_entry_off[1:]
    for p in range(0, 5000, 3):
        keep[starts[p]:ends[p]] = False
    per = np.array([keep[s:e].sum() for s, e in zip(starts, ends)])
    off = np.zeros(5001, np.uint32)
    np.cumsum(per, out=off[1:])
    ent = cols.entries[keep].copy()
    ent["pod"] = np.repeat(np.arange(5000, dtype=np.uint32), per)
    ragged = K.FleetColumns(node_flags=cols.node_flags, node_cap_i915=cols.node_cap_i915, node_cap_xe=cols.node_cap_xe,
                            node_alloc_i915=cols.node_alloc_i915, node_alloc_xe=cols.node_alloc_xe, node_alloc_own=cols.node_alloc_own,
                            pod_node=cols.pod_node, pod_phase=cols.pod_phase, pod_entry_off=off, entries=ent,
                            n_resources=cols.n_resources, options=cols.options)
    ragged.validate()
    assert_outputs_equal(ctx.alloc_rollup(ragged), oracle_lib.alloc_rollup(ragged))


def test_options_without_limits(ctx, oracle_lib):
    cols = synth.synth_fleet(40, 3000, seed=4, options=0)
    out = ctx.alloc_rollup(cols)
    assert out.pod_lim is None and out.node_lim is None
    assert_outputs_equal(out, oracle_lib.alloc_rollup(cols))


def test_extreme_values_stay_exact(ctx, oracle_lib):
    cols = synth.synth_fleet(5, 200, seed=6)
    cols.entries["value"] = (cols.entries["value"].astype(np.int64) * ((1 << 27) + 12345)) * np.where(np.arange(cols.n_entries) % 7 == 0, -1, 1)
    cols.validate()
    assert_outputs_equal(ctx.alloc_rollup(cols), oracle_lib.alloc_rollup(cols))


def test_extreme_values_stay_exact_when_grouped(ctx, oracle_lib):
    """The warp reduction of grouped pods splits 64-bit sums into 22-bit limbs: large and negative values must come
    out exactly like per-pod atomics do."""
    cols = synth.synth_fleet(6, 3000, seed=8)
    cols.entries["value"] = (cols.entries["value"].astype(np.int64) * ((1 << 25) + 777)) * np.where(np.arange(cols.n_entries) % 5 == 0, -1, 1)
    cols, _ = cols.grouped_by_node()
    cols.validate()
    assert_outputs_equal(ctx.alloc_rollup(cols), oracle_lib.alloc_rollup(cols))


def test_config_c2_full_size_grouped_bit_exact(ctx, oracle_lib):
    """C2 with the pod columns grouped by node (the bench's layout): the warp-reduction path at full size."""
    cols, order = synth.synth_fleet(10_000, 1_000_000).grouped_by_node()
    out = ctx.alloc_rollup(cols)
    assert_outputs_equal(out, oracle_lib.alloc_rollup(cols, threads=8), "C2 grouped")
    assert np.array_equal(np.sort(order), np.arange(cols.n_pods))


def test_config_c2_full_size_bit_exact(ctx, oracle_lib):
    """BASELINE config 2: 10k nodes / 1M pods on one GPU, every output column identical."""
    cols = synth.synth_fleet(10_000, 1_000_000)
    out = ctx.alloc_rollup(cols)
    assert_outputs_equal(out, oracle_lib.alloc_rollup(cols, threads=8), "C2")
    # size-independent properties: totals equal the sums of the per-node / per-pod columns
    gpu = out.pod_is_gpu.astype(bool)
    run = gpu & (cols.pod_phase == K.PHASE_RUNNING)
    assert out.fleet_totals[K.T_GPU_PODS] == gpu.sum()
    assert out.fleet_totals[K.T_ALLOCATED] == out.pod_req[0, run].sum() + out.pod_req[1, run].sum()
    assert out.fleet_totals[K.T_SCHEDULED_PODS] == out.node_pod_count.sum()
    for r in range(cols.n_resources):
        assert out.fleet_totals[K.T_REQ0 + r] == out.pod_req[r, run].sum()
        sched_run = run & (cols.pod_node != K.NODE_NONE)
        assert out.node_req[r].sum() == out.pod_req[r, sched_run].sum()


def test_config_c3_shard_bit_exact(ctx, oracle_lib):
    """BASELINE config 3 (10 M pods / 100 k nodes over 8 GPUs): one rank's node-range shard — 12,500 nodes and the
    ≈ 0.93 M pods scheduled onto them — in API order and grouped by node, every column identical to the oracle."""
    from headlamp_intel_gpu_plugin_b200 import shard

    lo, hi = shard.node_range(100_000, 3, 8)
    cols = synth.synth_fleet(100_000, 10_000_000, node_range=(lo, hi))
    assert cols.n_nodes == 12_500 and 800_000 < cols.n_pods < 1_100_000
    want = oracle_lib.alloc_rollup(cols, threads=8)
    assert_outputs_equal(ctx.alloc_rollup(cols), want, "C3 shard")
    grouped, order = cols.grouped_by_node()
    got = ctx.alloc_rollup(grouped)
    for f in ("node_pod_count", "node_inuse", "node_req", "node_lim", "fleet_totals", "node_bar_pct", "node_util_pct"):
        assert np.array_equal(getattr(got, f), getattr(want, f)), f
    assert np.array_equal(got.pod_req[:, np.argsort(order)], want.pod_req)


def test_staged_api_and_repeatability(ctx, oracle_lib):
    cols = synth.synth_fleet(500, 40_000, seed=12)
    want = oracle_lib.alloc_rollup(cols)
    fleet = ctx.fleet_create(cols.n_nodes, cols.n_pods, cols.n_entries, cols.n_resources, cols.options)
    pinned = native.pinned_copy_of_columns(cols)
    fleet.upload(pinned)
    for _ in range(3):  # accumulators are re-zeroed per rollup
        fleet.rollup()
        assert_outputs_equal(fleet.download(), want)
    t = ctx.timing()
    assert t["kernel_ms"] > 0 and t["kernel_launches"] >= 2
    fleet.destroy()


def test_error_paths(ctx):
    cols = synth.synth_fleet(4, 10)
    bad = K.FleetColumns(**{f: getattr(cols, f) for f in ("node_flags node_cap_i915 node_cap_xe node_alloc_i915 node_alloc_xe "
                                                          "node_alloc_own pod_node pod_phase pod_entry_off entries").split()},
                         n_resources=40)
    with pytest.raises(native.IgpuError) as e:
        ctx.alloc_rollup(bad, K.FleetOutputs.allocate(4, 10, 32))
    assert e.value.status == -6
    fleet = ctx.fleet_create(2, 2, 8, 5)
    with pytest.raises(native.IgpuError) as e:
        fleet.upload(cols)
    assert e.value.status == -6
    with pytest.raises(native.IgpuError) as e:
        fleet.rollup()
    assert e.value.status == -7
    fleet.destroy()


@pytest.mark.parametrize("n_nodes,n_pods", [(7, 300), (300, 20_000), (40, 70_000), (2_000, 9_000)])
def test_grouped_pods_take_the_run_reduction_and_change_nothing(ctx, oracle_lib, n_nodes, n_pods):
    """Pods grouped by node (FleetColumns.grouped_by_node) make the tiled kernel sum each node run with a warp and
    issue one set of atomics per run; API-order pods keep per-pod atomics.  Both must equal the oracle bit for bit,
    and equal each other after undoing the permutation."""
    cols = synth.synth_fleet(n_nodes, n_pods, seed=n_pods)
    grouped, order = cols.grouped_by_node()
    grouped.validate()
    inv = np.argsort(order)
    got_api, got_grp, want_grp = ctx.alloc_rollup(cols), ctx.alloc_rollup(grouped), oracle_lib.alloc_rollup(grouped)
    for f in K.FleetOutputs.FIELDS:
        a, g, w = getattr(got_api, f), getattr(got_grp, f), getattr(want_grp, f)
        if a is None:
            continue
        assert np.array_equal(g, w), f
        if f.startswith("pod_"):
            g = g.reshape(-1, n_pods)[:, inv].reshape(g.shape)
        assert np.array_equal(a, g), f


def test_grouped_dirty_fleet_and_legacy_kernel_agree(ctx, oracle_lib, switch):
    nodes, pods = dirty_fleet(60, 4000, 5)
    cols = k8s.marshalFleet(nodes, pods)
    grouped, _ = cols.grouped_by_node()
    want = oracle_lib.alloc_rollup(grouped)
    got = ctx.alloc_rollup(grouped)
    switch("IGPU_K1_POD_PARALLEL")
    pod_parallel = ctx.alloc_rollup(grouped)
    switch("IGPU_K1_LEGACY")
    legacy = ctx.alloc_rollup(grouped)
    for f in K.FleetOutputs.FIELDS:
        if getattr(want, f) is not None:
            assert np.array_equal(getattr(got, f), getattr(want, f)), f
            assert np.array_equal(getattr(pod_parallel, f), getattr(want, f)), f
            assert np.array_equal(getattr(legacy, f), getattr(want, f)), f


@pytest.mark.parametrize("grouped", [False, True])
def test_tiles_outside_the_32_bit_domain_take_the_64_bit_path(ctx, oracle_lib, grouped):
    """The entry-parallel kernel sums in 32 bits where a tile's values allow it; a tile holding a value outside [0, 2^20)
    or more than 2,048 entries is folded in 64 bits.  Mixed tiles, same integers as the oracle either way."""
    cols = synth.synth_fleet(300, 40_000, seed=12)
    e = cols.entries.copy()
    rng = np.random.default_rng(5)
    for i in rng.integers(0, len(e), 40):  # ~40 of ~160 tiles leave the fast domain
        e["value"][i] = int(rng.choice([1 << 20, (1 << 40) + 3, -1, -(1 << 33), (1 << 20) - 1, 1 << 31, 1 << 32]))
    # one pod with 3,000 entries: its tile is beyond the entry bound of the 32-bit path and of the staging buffer
    off = cols.pod_entry_off.astype(np.int64)
    at, big = int(off[20_000]), 3000
    extra = np.zeros(big, K.ENTRY_DTYPE)
    extra["pod"], extra["res"], extra["flags"], extra["container"] = 20_000, np.arange(big) % 5, K.ENTRY_TRUTHY, 2
    extra["value"] = rng.integers(0, 1 << 19, big)
    e = np.concatenate([e[:at], extra, e[at:]])
    off[20_001:] += big
    # the generated entries of that pod follow the inserted run; containers 0..2 precede it in a valid stream, so keep order valid:
    own = slice(at + big, int(off[20_001]))
    e["container"][own] = np.maximum(e["container"][own], 2)
    e["flags"][own] &= ~np.uint8(K.ENTRY_INIT)
    cols = K.FleetColumns(node_flags=cols.node_flags, node_cap_i915=cols.node_cap_i915, node_cap_xe=cols.node_cap_xe,
                          node_alloc_i915=cols.node_alloc_i915, node_alloc_xe=cols.node_alloc_xe, node_alloc_own=cols.node_alloc_own,
                          pod_node=cols.pod_node, pod_phase=cols.pod_phase, pod_entry_off=off.astype(np.uint32), entries=e,
                          n_resources=cols.n_resources, options=cols.options)
    cols.validate()
    if grouped:
        cols, _ = cols.grouped_by_node()
    assert_outputs_equal(ctx.alloc_rollup(cols), oracle_lib.alloc_rollup(cols), "mixed 32 / 64-bit tiles")


@pytest.mark.parametrize("seed", [1, 2])
def test_rollup_fleet_grouped_keeps_the_callers_pod_indices(ctx, seed):
    """rollupFleet(group_by_node=True): columns regrouped for the device, views still indexed by the caller's pods;
    every reference-visible quantity equals the per-object restatement."""
    nodes, pods = dirty_fleet(30, 600, seed)
    r = k8s.rollupFleet(nodes, pods, ctx, group_by_node=True)
    assert r.order is not None and not np.array_equal(r.order, np.arange(len(r.order)))
    check_against_objects(nodes, pods, r.cols, r.out, r.order, r.pos)
    plain = k8s.rollupFleet(nodes, pods, ctx, group_by_node=False)
    assert r.overview() == plain.overview() or str(r.overview()) == str(plain.overview())
    assert np.array_equal(r.gpu_pod_rows(), plain.gpu_pod_rows())


def test_rollup_fleet_json_text_to_views(ctx):
    """JSON text -> native marshal (grouped by node in C++) -> GPU -> views in the caller's pod order."""
    nodes, pods = dirty_fleet(30, 700, 11)
    nj, pj = json.dumps(nodes), json.dumps(pods)
    r = k8s.rollupFleetJson(nj, pj, ctx, group_by_node=True, threads=3)
    assert r.order is not None
    check_against_objects(nodes, pods, r.cols, r.out, r.order, r.pos)
    rc = k8s.rollupFleetJson(nj, pj, ctx, group_by_node=True, threads=2, container_rows=True)
    check_against_objects(nodes, pods, rc.cols, rc.out, rc.order, rc.pos, container_rows=rc.container_rows)
    plain = k8s.rollupFleetJson(nj, pj, ctx, group_by_node=False)
    assert plain.order is None and str(plain.overview()) == str(r.overview())
    assert np.array_equal(plain.gpu_pod_rows(), r.gpu_pod_rows())


@pytest.mark.parametrize("legacy", [False, True])
@pytest.mark.parametrize("what", ["offsets_descend", "offset_past_stream", "resource_id", "node_id", "entry_pod"])
def test_broken_columns_are_rejected_not_followed(ctx, oracle_lib, switch, what, legacy):
    """Columns from a caller that skipped validate(): the device counts and skips what breaks the invariants (no wild
    reads), the call returns IGPU_ERR_ARG, and the context keeps working."""
    if legacy:
        switch("IGPU_K1_LEGACY")
    cols = synth.synth_fleet(40, 3000, seed=21)
    good = oracle_lib.alloc_rollup(cols)
    off, ent, node = cols.pod_entry_off.copy(), cols.entries.copy(), cols.pod_node.copy()
    if what == "offsets_descend":
        off[1500] = off[1400]
        off[1499] = off[1400] + 3 if off[1400] + 3 <= off[-1] else off[1400]
        off[1500] = off[1499] - 2
    elif what == "offset_past_stream":
        off[700] = 0xFFFFFFF0
    elif what == "resource_id":
        ent["res"][123] = 7
        ent["res"][4567] = 0xFFFF
    elif what == "entry_pod":
        ent["pod"][2000] += 1          # a neighbour in the same tile
        ent["pod"][7000] = 0xFFFFFF00  # nowhere
    else:
        node[5] = cols.n_nodes + 5
        node[2999] = 0x7FFFFFFF
    bad = K.FleetColumns(node_flags=cols.node_flags, node_cap_i915=cols.node_cap_i915, node_cap_xe=cols.node_cap_xe,
                         node_alloc_i915=cols.node_alloc_i915, node_alloc_xe=cols.node_alloc_xe, node_alloc_own=cols.node_alloc_own,
                         pod_node=node, pod_phase=cols.pod_phase, pod_entry_off=off, entries=ent, n_resources=cols.n_resources,
                         options=cols.options)
    with pytest.raises(ValueError):
        bad.validate()
    with pytest.raises(native.IgpuError) as e:
        ctx.alloc_rollup(bad)
    assert e.value.status == -1 and "column invariants" in str(e.value)
    assert_outputs_equal(ctx.alloc_rollup(cols), good)  # the context is not poisoned



def test_upstream_pod_detail_fixtures_through_cuda(ctx):
    """The reference's own PodDetailSection fixtures (PodDetailSection.test.tsx:40-78): "GPU Containers" is 1 for the
    trainer pod (:131-137), the limits-only pod still renders (:119-124) with a request row whose value is absent."""
    from oracle import ref_objects as ro

    golden = json.load(open(os.path.join(HERE, "golden", "pod_detail_vectors.json")))
    pods = [v["pod"] for v in golden["vectors"]]
    r = k8s.rollupFleet([], pods, ctx, container_rows=True)
    for i, v in enumerate(golden["vectors"]):
        assert r.pod_gpu_container_count(i) == v["gpuContainers"], v["ref"]
        got = r.pod_detail_section(i, pods[i])
        assert got == ro.pod_detail_section(pods[i])
        assert (got is None) == v["rendersNothing"], v["ref"]
        if got is not None:
            assert [list(x) for x in got["rows"]] == v["rows"], v["ref"]


def test_container_rows_at_scale_match_the_columnar_oracle(ctx, oracle_lib):
    cols = synth.synth_fleet(500, 200_000, seed=4)  # ~100 scan blocks, rows in (pod, container) order
    fleet = ctx.fleet_create(cols.n_nodes, cols.n_pods, cols.n_entries, cols.n_resources, cols.options)
    fleet.upload(cols)
    fleet.rollup()
    out, rows = fleet.download(), fleet.container_rows()
    fleet.destroy()
    want = oracle_lib.container_rows(cols)
    assert np.array_equal(rows, want) and len(rows) == int(out.pod_gpu_containers.sum())
    assert np.array_equal(out.pod_gpu_containers, oracle_lib.alloc_rollup(cols).pod_gpu_containers)


@pytest.mark.parametrize("n_res", [4, 6, 7, 8])
def test_every_compiled_resource_count_on_grouped_input(ctx, oracle_lib, n_res):
    """The kernels are compiled per resource count (4..8).  Small values and pods grouped by node keep every tile on the
    32-bit path with its per-tile node table; with 8 resources that table has more words than the CTA has threads."""
    cols = synth.synth_fleet(400, 60_000, seed=21)
    e = cols.entries.copy()
    e["res"] = (e["res"].astype(np.uint32) * 3 + e["pod"] % 5) % n_res  # every id below n_res occurs, on every node
    cols = K.FleetColumns(node_flags=cols.node_flags, node_cap_i915=cols.node_cap_i915, node_cap_xe=cols.node_cap_xe,
                          node_alloc_i915=cols.node_alloc_i915, node_alloc_xe=cols.node_alloc_xe, node_alloc_own=cols.node_alloc_own,
                          pod_node=cols.pod_node, pod_phase=cols.pod_phase, pod_entry_off=cols.pod_entry_off, entries=e,
                          n_resources=n_res, options=cols.options)
    cols, _ = cols.grouped_by_node()
    cols.validate()
    want = oracle_lib.alloc_rollup(cols)
    assert_outputs_equal(ctx.alloc_rollup(cols), want, f"R = {n_res}")
    assert (want.node_req[n_res - 1] != 0).sum() > 300  # the last plane is populated on (almost) every node
