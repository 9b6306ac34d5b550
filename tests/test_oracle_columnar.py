"""CPU suite: the columnar C++ oracle (rollup_oracle.cpp) and the host marshal layer agree with the
object-level restatement of the reference on generated and adversarial fleets; the window / quantile
oracles agree with their numpy definitions; the three generators (numpy, C++, CUDA's twin) agree."""
import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import k8s, metrics, synth
from oracle import ref_objects as ro

from .dirty import dirty_fleet
from .parity import assert_outputs_equal, check_against_objects, nan_eq


def test_marshal_reproduces_generated_columns():
    cols = synth.synth_fleet(50, 2000)  # config C1
    nodes, pods = synth.fleet_objects(50, 2000)
    m = k8s.marshalFleet(nodes, pods)
    for f in ("node_flags node_cap_i915 node_cap_xe node_alloc_i915 node_alloc_xe node_alloc_own pod_node pod_phase "
              "pod_entry_off entries").split():
        assert np.array_equal(getattr(m, f), getattr(cols, f)), f
    assert 3.0 < cols.n_entries / cols.n_pods < 4.5  # E ~ 4P


@pytest.mark.parametrize("n_nodes,n_pods,seed,plain_bp", [(50, 2000, 20261106, 0), (7, 300, 5, 3000), (1, 40, 9, 0)])
def test_columnar_oracle_equals_object_oracle_generated(oracle_lib, n_nodes, n_pods, seed, plain_bp):
    nodes, pods = synth.fleet_objects(n_nodes, n_pods, seed, non_gpu_share_bp=plain_bp)
    cols = k8s.marshalFleet(nodes, pods)
    check_against_objects(nodes, pods, cols, oracle_lib.alloc_rollup(cols), container_rows=oracle_lib.container_rows(cols))


@pytest.mark.parametrize("seed", [7, 8, 9, 10])
def test_columnar_oracle_equals_object_oracle_dirty(oracle_lib, seed):
    nodes, pods = dirty_fleet(40, 400, seed)
    cols = k8s.marshalFleet(nodes, pods)
    assert cols.n_resources > 5  # custom resource names got dictionary ids
    out = oracle_lib.alloc_rollup(cols)
    rows = oracle_lib.container_rows(cols)
    check_against_objects(nodes, pods, cols, out, container_rows=rows)
    assert out.fleet_totals[K.T_CAP_NAN] > 0
    # the per-container views are exercised: both display forms, keys in limits only, a second GPU container
    assert (rows["same"] & rows["req_truthy"] & rows["lim_truthy"]).any() and (rows["lim_keys"] & ~rows["req_keys"]).any()
    assert out.pod_gpu_containers.max() >= 2 and len(rows) == int(out.pod_gpu_containers.sum())  # the NaN-poisoning path is exercised


def test_oracle_threads_are_bit_identical(oracle_lib):
    cols = synth.synth_fleet(300, 20000, seed=3)
    assert_outputs_equal(oracle_lib.alloc_rollup(cols, 1), oracle_lib.alloc_rollup(cols, 4), "1 vs 4 threads")


def test_empty_and_ragged(oracle_lib):
    empty = k8s.marshalFleet([], [])
    out = oracle_lib.alloc_rollup(empty)
    assert not out.fleet_totals.any()
    nodes, pods = synth.fleet_objects(3, 0)
    check_against_objects(nodes, pods, k8s.marshalFleet(nodes, pods), oracle_lib.alloc_rollup(k8s.marshalFleet(nodes, pods)))
    nodes, pods = [], synth.fleet_objects(3, 25)[1]
    check_against_objects(nodes, pods, k8s.marshalFleet(nodes, pods), oracle_lib.alloc_rollup(k8s.marshalFleet(nodes, pods)))


def test_phase_in_operator_quirk(oracle_lib):
    # OverviewPage.tsx:125 `phase in podPhaseCounts` is true for inherited names: such a pod is in no bucket
    pod = {"metadata": {"name": "p"}, "spec": {"containers": [{"name": "c", "resources": {"limits": {"gpu.intel.com/i915": "1"}}}]},
           "status": {"phase": "toString"}}
    cols = k8s.marshalFleet([], [pod])
    assert cols.pod_phase[0] == K.PHASE_PROTO
    t = oracle_lib.alloc_rollup(cols).fleet_totals
    assert t[K.T_GPU_PODS] == 1 and t[K.T_PHASE0:K.T_PHASE0 + 5].sum() == 0
    assert ro.overview_rollup([], [pod])["podPhaseCounts"] == {"Running": 0, "Pending": 0, "Succeeded": 0, "Failed": 0, "Other": 0}


def test_marshal_rejects_values_outside_exact_domain():
    pod = {"spec": {"containers": [{"resources": {"requests": {"gpu.intel.com/i915": "9007199254740993"}}}]}}
    with pytest.raises(ValueError):
        k8s.marshalFleet([], [pod])


# ---- metrics.ts join ---------------------------------------------------------------------------------


def test_power_join_equals_object_oracle(oracle_lib):
    ch, en, mx, un = synth.synth_power_results(25, 4)
    # a last-write-wins collision (metrics.ts:127-138 keys the maps by chip label only) + odd values
    en.append({"metric": {"chip": ch[0]["metric"]["chip"], "instance": "other"}, "value": [0, "7.5"]})
    en.append({"metric": {"chip": ch[1]["metric"]["chip"]}, "value": [0, "+Inf"]})
    mx.append({"metric": {"chip": ch[2]["metric"]["chip"]}, "value": [0, "NaN"]})
    en.append({"metric": {}, "value": [0, "1"]})
    un.append({"metric": {"instance": ch[5]["metric"]["instance"], "node": "renamed"}, "value": [0, "1"]})
    ch.append({"metric": {"chip": "lonely"}, "value": [0, "1"]})
    p = metrics.marshalMetrics(ch, en, mx, un)
    out = oracle_lib.power_join_sum(p)
    rows = metrics._decode_chips(p, out)
    ref = ro.join_gpu_metrics(ch, en, mx, un)
    assert nan_eq(rows, ref)
    assert rows[0]["powerWatts"] == 7.5
    fp = ro.fleet_power(ref)
    assert nan_eq(float(out["fleet"][0]), fp["totalWatts"]) and nan_eq(float(out["fleet"][1]), fp["maxTotalWatts"])


def test_power_bar_pct(oracle_lib):
    ch, en, mx, un = synth.synth_power_results(10, 2, missing_bp=0)
    p = metrics.marshalMetrics(ch, en, mx, un)
    out = oracle_lib.power_join_sum(p)
    fp = ro.fleet_power(ro.join_gpu_metrics(ch, en, mx, un))
    assert out["fleet"][2] == fp["pct"] and abs(out["fleet"][0] - fp["totalWatts"]) <= 1e-9 * fp["totalWatts"]


# ---- series: generators and A18 definitions ----------------------------------------------------------


def test_generators_bit_identical(oracle_lib):
    a = synth.synth_series(6, 3000, nan_per_million=2000, global_first=11)
    b = oracle_lib.generate_series(6, 3000, synth.DEFAULT_SEED, 11, 2000)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert np.isnan(a).sum() > 0 and np.nanmin(a) >= 0 and np.nanmax(a) <= 300


@pytest.mark.parametrize("T,W", [(1440, 60), (1000, 60), (999, 7), (64, 64), (10, 100)])
def test_window_stats_oracle_matches_numpy(oracle_lib, T, W):
    x = synth.synth_series(5, T, nan_per_million=20000)
    x[2, :] = np.nan
    mn, me, mx = oracle_lib.window_stats(x, W)
    nw = (T + W - 1) // W
    for s in range(5):
        for w in range(nw):
            seg = x[s, w * W:(w + 1) * W]
            v = seg[~np.isnan(seg)]
            if v.size == 0:
                assert np.isnan(mn[s, w]) and np.isnan(me[s, w]) and np.isnan(mx[s, w])
            else:
                assert mn[s, w] == v.min() and mx[s, w] == v.max()
                assert me[s, w] == np.float32(v.astype(np.float64).sum() / v.size) or abs(me[s, w] - v.astype(np.float64).mean()) < 1e-5 * abs(v.mean())


@pytest.mark.parametrize("q,mode", [(0.95, 0), (0.5, 0), (0.0, 0), (1.0, 0), (0.95, 1), (0.01, 1)])
def test_quantile_oracle_matches_numpy(oracle_lib, q, mode):
    x = synth.synth_series(9, 700, nan_per_million=30000)
    x[4, :] = np.nan
    off = np.array([0, 2, 4, 5, 5, 9], np.uint32)  # node 2 is all-NaN, node 3 has no series
    qv, st, fleet = oracle_lib.quantile_by_node(x, off, q, mode)
    for n in range(5):
        v = x[off[n]:off[n + 1]].reshape(-1)
        v = np.sort(v[~np.isnan(v)]).astype(np.float64)
        if v.size == 0:
            assert np.isnan(qv[n]) and st["count"][n] == 0
            continue
        if mode == 0:
            rank = q * (v.size - 1)
            lo = int(np.floor(rank)); hi = min(v.size - 1, lo + 1); w = rank - lo
            want = v[lo] * (1 - w) + v[hi] * w
        else:
            want = v[max(0, int(np.ceil(q * v.size)) - 1)]
        assert qv[n] == want
        assert st["count"][n] == v.size and st["min"][n] == v.min() and st["max"][n] == v.max()
        assert abs(st["sum"][n] - v.sum()) <= 1e-12 * v.sum()
    assert fleet[K.F_NODES] == 3 and fleet[K.F_COUNT] == st["count"].sum()
