"""CPU suite: the native JSON marshal (csrc/marshal.cpp, igpu_marshal_fleet_json) produces exactly the
columns of the Python marshal (k8s.marshalFleet), which tests/test_oracle_columnar.py ties to the
object-level restatement of the reference.  No GPU needed: the extractor is host code."""
import json

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import k8s, native, synth

from .dirty import dirty_fleet

# `entries` is compared as a whole record: pod, res, flags (incl. ENTRY_SAME), container ordinal, value
COLS = ("node_flags node_cap_i915 node_cap_xe node_alloc_i915 node_alloc_xe node_alloc_own pod_node pod_phase "
        "pod_entry_off entries").split()


def assert_same_columns(a: K.FleetColumns, b: K.FleetColumns):
    for f in COLS:
        x, y = getattr(a, f), getattr(b, f)
        assert x.shape == y.shape and np.array_equal(x, y), f
    assert a.n_resources == b.n_resources and a.resource_names == b.resource_names


@pytest.mark.parametrize("n_nodes,n_pods,seed,plain", [(50, 2000, 20261106, 0), (7, 300, 5, 3000), (1, 40, 9, 0), (3, 0, 1, 0)])
@pytest.mark.parametrize("threads", [1, 4])
def test_generated_fleet(n_nodes, n_pods, seed, plain, threads):
    nodes, pods = synth.fleet_objects(n_nodes, n_pods, seed, non_gpu_share_bp=plain)
    got = k8s.marshalFleetJson(json.dumps(nodes), json.dumps(pods), threads=threads)
    assert_same_columns(got, k8s.marshalFleet(nodes, pods))


@pytest.mark.parametrize("seed", [7, 8, 9, 10, 11, 12])
@pytest.mark.parametrize("threads", [1, 3, 16])
def test_dirty_fleet(seed, threads):
    nodes, pods = dirty_fleet(40, 400, seed)
    got = k8s.marshalFleetJson(json.dumps(nodes), json.dumps(pods), threads=threads)
    want = k8s.marshalFleet(nodes, pods)
    assert_same_columns(got, want)
    assert got.n_resources > 5  # custom resource names got the same dictionary ids in the same order


def test_kubelist_envelope_escapes_and_key_order():
    # real API servers wrap items in a List and emit limits before requests, initContainers before containers
    pod = {"metadata": {"name": "p"}, "status": {"phase": "Running"},
           "spec": {"nodeName": "n\\u00e9", "initContainers": [{"name": "i", "resources": {"limits": {"gpu.intel.com/xe": "3"}}}],
                    "containers": [{"name": "c", "resources": {"limits": {"gpu.intel.com\\/i915": "2"}, "requests": {"gpu.intel.com/i915": "\\u0032"}}}]}}
    node = {"metadata": {"name": "né", "labels": {"node-role.kubernetes.io/gpu": "true"}}, "status": {"capacity": {"gpu.intel.com/i915": " 4"}}}
    pods_text = '{"kind":"PodList","apiVersion":"v1","metadata":{"resourceVersion":"1"},"items":[' + json.dumps(pod).replace("\\\\", "\\") + ']}'
    nodes_text = '{"items": [' + json.dumps(node, ensure_ascii=False) + '], "kind": "NodeList"}'
    got = k8s.marshalFleetJson(nodes_text, pods_text)
    want = k8s.marshalFleet([json.loads(json.dumps(node))], [json.loads(json.dumps(pod).replace("\\\\", "\\"))])
    assert_same_columns(got, want)
    assert got.pod_node[0] == 0 and got.n_entries == 3
    e = got.entries
    # the request "\u0032" and the limit "2" are the same string once unescaped: req === lim (PodsPage.tsx:69)
    assert list(e["flags"]) == [K.ENTRY_TRUTHY | K.ENTRY_SAME, K.ENTRY_LIMIT | K.ENTRY_TRUTHY | K.ENTRY_SAME,
                                K.ENTRY_INIT | K.ENTRY_LIMIT | K.ENTRY_TRUTHY]
    assert list(e["container"]) == [0, 0, 0]
    assert list(e["value"]) == [2, 2, 3] and got.node_cap_i915[0] == 4


def test_js_number_quirks_in_json_numbers():
    # numbers instead of strings: parseInt(String(x)) — 2.9 -> 2, 1e-7 -> "1e-7" -> 1, 1e21 -> "1e+21" -> 1, 0 is falsy
    vals = [2.9, 1e-7, 1e21, 0, -0.5, 123456789, True, False, None, [], {}]
    pods = [{"spec": {"containers": [{"resources": {"requests": {"gpu.intel.com/i915": v}}}]}} for v in vals]
    got = k8s.marshalFleetJson("[]", json.dumps(pods))
    want = k8s.marshalFleet([], pods)
    assert_same_columns(got, want)
    assert list(got.entries["value"]) == [2, 1, 1, 0, 0, 123456789, 0, 0, 0, 0, 0]


def test_errors():
    with pytest.raises(native.IgpuError) as e:
        k8s.marshalFleetJson("[]", '[{"spec": {"containers": [')
    assert e.value.status == -1
    big = [{"spec": {"containers": [{"resources": {"requests": {"gpu.intel.com/i915": "9007199254740993"}}}]}}]
    with pytest.raises(native.IgpuError) as e:
        k8s.marshalFleetJson("[]", json.dumps(big))
    assert e.value.status == -6
    many = [{"spec": {"containers": [{"resources": {"requests": {f"gpu.intel.com/r{i}": "1" for i in range(40)}}}]}}]
    with pytest.raises(native.IgpuError) as e:
        k8s.marshalFleetJson("[]", json.dumps(many))
    assert e.value.status == -6
    assert k8s.marshalFleetJson("", "").n_pods == 0


def test_throughput_smoke():
    import time

    nodes, pods = synth.fleet_objects(200, 20000)
    nj, pj = json.dumps(nodes), json.dumps(pods)
    best = float("inf")
    for _ in range(5):  # best of five: a wall-clock bound must not fail because the box was busy for one of them
        t = time.perf_counter()
        got = k8s.marshalFleetJson(nj, pj)
        best = min(best, time.perf_counter() - t)
    assert got.n_pods == 20000
    assert 20000 / best > 120_000  # the Python marshal manages ~60k pods/s, json.loads alone ~57k pods/s; this one millions


def test_block_splitter_with_escapes_and_braces_in_strings():
    """The 64-byte block scanner must ignore braces / quotes inside strings, including escaped quotes and
    backslash runs that straddle block boundaries (kubectl's last-applied-configuration annotation is
    exactly such a string)."""
    import random

    rng = random.Random(5)
    nodes, pods = synth.fleet_objects(5, 300, seed=3)
    nasty = ['{"a": "}"}', '\\', '\\\\', '"', '\\"', ']}[{', 'x' * 61 + '\\', 'y' * 62 + '"', '\\\\"', '{"spec":{"nodeName":"node-000001"}}']
    for p in pods:
        p["metadata"]["annotations"] = {"kubectl.kubernetes.io/last-applied-configuration": json.dumps(p["spec"]) + rng.choice(nasty) * rng.randrange(1, 4),
                                        rng.choice(nasty): rng.choice(nasty)}
        p["metadata"]["pad"] = "z" * rng.randrange(0, 70)  # moves the block phase around
    for indent in (None, 1):
        text = json.dumps(pods, indent=indent)
        got = k8s.marshalFleetJson(json.dumps(nodes), text, threads=3)
        assert_same_columns(got, k8s.marshalFleet(nodes, pods))
    # the envelope form, too
    got = k8s.marshalFleetJson(json.dumps({"kind": "NodeList", "items": nodes}), json.dumps({"metadata": {"x": "[{"}, "items": pods}))
    assert_same_columns(got, k8s.marshalFleet(nodes, pods))


@pytest.mark.parametrize("threads", [1, 4])
def test_native_grouping_equals_numpy_grouping(threads):
    """OPT_GROUP_PODS_BY_NODE: the marshal's own counting sort gives exactly FleetColumns.grouped_by_node()."""
    from headlamp_intel_gpu_plugin_b200 import columns as K

    for nodes, pods in (synth.fleet_objects(30, 2500, 3), dirty_fleet(25, 900, 4), ([], []), (synth.fleet_objects(3, 0)[0], [])):
        nj, pj = json.dumps(nodes).encode(), json.dumps(pods).encode()
        plain = k8s.marshalFleetJson(nj, pj, threads=threads)
        got = k8s.marshalFleetJson(nj, pj, options=K.OPT_POD_LIMITS | K.OPT_GROUP_PODS_BY_NODE, threads=threads)
        want, order = plain.grouped_by_node()
        assert got.options == plain.options == K.OPT_POD_LIMITS
        if plain.n_pods:
            assert np.array_equal(got.pod_order, order)
        for f in ("pod_node", "pod_phase", "pod_entry_off", "entries", "node_flags", "node_cap_i915"):
            assert np.array_equal(getattr(got, f), getattr(want, f)), f
        got.validate()
        assert plain.pod_order is None


def test_mutated_json_is_accepted_or_rejected_never_worse():
    """A deterministic slice of scripts/fuzz_marshal.py: truncated, spliced and byte-flipped List / matrix JSON either
    marshals or raises IgpuError(ARG | RANGE).  (The same mutations run clean under ASAN + UBSAN: scripts/fuzz_marshal_asan.cpp.)"""
    import random

    from headlamp_intel_gpu_plugin_b200 import metrics

    rnd = random.Random(12345)
    nodes, pods = dirty_fleet(5, 30, 2)
    nj, pj = json.dumps(nodes).encode(), json.dumps(pods).encode()
    mx = json.dumps({"status": "success", "data": {"resultType": "matrix", "result": [
        {"metric": {"chip": 'a"b'}, "values": [[1_700_000_000 + 15 * i, repr(i * 1.5)] for i in range(20)]} for _ in range(3)]}}).encode()
    special = b'"\\{}[],:0-9eE.+ \n\ttruefalsn\xff\xc2\xa0'

    def mutate(b):
        b = bytearray(b)
        for _ in range(rnd.randint(1, 4)):
            if not b:
                break
            i, k = rnd.randrange(len(b)), rnd.random()
            if k < 0.3:
                b[i] = rnd.choice(special)
            elif k < 0.5:
                del b[i:i + rnd.randint(1, 20)]
            elif k < 0.7:
                b[i:i] = bytes(rnd.choice(special) for _ in range(rnd.randint(1, 6)))
            elif k < 0.85:
                b = b[:i]
            else:
                b[i:i] = b[max(0, i - rnd.randint(1, 40)):i]
        return bytes(b)

    outcomes = {"ok": 0, "rejected": 0}
    for it in range(600):
        try:
            if it % 3 == 0:
                k8s.marshalFleetJson(nj, mutate(pj), options=K.OPT_POD_LIMITS | (K.OPT_GROUP_PODS_BY_NODE if it % 2 else 0), threads=1 + it % 3)
            elif it % 3 == 1:
                k8s.marshalFleetJson(mutate(nj), pj, threads=2)
            else:
                metrics.marshalMatrixJson(mutate(mx), 1_700_000_000_000, 15_000, 20, labels=("chip",), threads=1 + it % 2)
            outcomes["ok"] += 1
        except native.IgpuError as e:
            assert e.status in (-1, -6), e
            outcomes["rejected"] += 1
    assert outcomes["ok"] > 20 and outcomes["rejected"] > 200, outcomes


def test_absurd_timestamps_are_dropped_not_overflowed():
    from headlamp_intel_gpu_plugin_b200 import metrics

    vals = [[99999999999999999999999, "1"], [-9223372036854775808, "2"], [1e300, "3"], [1_700_000_000, "4"], [1.7e9, "5"]]
    text = json.dumps({"status": "success", "data": {"result": [{"metric": {}, "values": vals}]}})
    v, _, dropped = metrics.marshalMatrixJson(text, 1_700_000_000_000, 15_000, 2)
    assert v[0, 0] == 5.0 and dropped == 3  # the exponent spelling of the same instant lands on the same step (last wins)



@pytest.mark.parametrize("item", ["0", "[]", '""', "{}"])
@pytest.mark.parametrize("threads", [1, 4])
def test_long_list_of_tiny_items_does_not_overrun_the_block_tables(item, threads):
    # The reference's JSON.parse + filter drops such elements (isGpuRequestingPod: k8s.ts:250-252).  Scalars are
    # items to the splitter too, and "0," is two bytes: the block tables used to be sized for three bytes per item.
    n = 300_000
    got = k8s.marshalFleetJson("[]", "[" + ",".join([item] * n) + "]", threads=threads)
    assert got.n_pods == n and got.n_entries == 0 and np.all(got.pod_node == K.NODE_NONE)
    assert np.all(got.pod_phase == K.PHASE_OTHER) or item == "{}"


def test_duplicate_keys_follow_json_parse_last_value_first_position():
    """JSON.parse keeps, for a key written twice, the position of the first and the value of the last."""
    pods_text = ('[{"spec":{"containers":[{"resources":{"requests":{"gpu.intel.com/i915":"1","gpu.intel.com/xe":"5","gpu.intel.com/i915":"3"},'
                 '"limits":{"gpu.intel.com/i915":"9"},"limits":{"gpu.intel.com/i915":"3"}}}]},"spec":{"nodeName":"n","containers":['
                 '{"name":"plain"},{"resources":{"requests":{"gpu.intel.com/i915":"1","gpu.intel.com/i915":"3"},"limits":{"gpu.intel.com/i915":3}}}]},'
                 '"status":{"phase":"Running"},"status":{}}]')
    got = k8s.marshalFleetJson('[{"metadata":{"name":"n"}}]', pods_text)
    want = k8s.marshalFleet([{"metadata": {"name": "n"}}], json.loads(pods_text))
    assert_same_columns(got, want)
    e = got.entries
    assert list(e["value"]) == [3, 3] and list(e["container"]) == [1, 1] and got.pod_phase[0] == K.PHASE_OTHER and got.pod_node[0] == 0
    assert list(e["flags"]) == [K.ENTRY_TRUTHY, K.ENTRY_LIMIT | K.ENTRY_TRUTHY]  # "3" === 3 is false


@pytest.mark.parametrize("req,lim,same", [('"2"', '"2"', True), ('"2"', '"02"', False), ("2", "2.0", True), ("2", '"2"', False),
                                          ("true", "true", True), ("null", "null", True), ("[]", "[]", False), ('""', '""', True)])
def test_same_flag_is_strict_equality_of_the_raw_values(req, lim, same):
    text = '[{"spec":{"containers":[{"resources":{"requests":{"gpu.intel.com/i915":%s},"limits":{"gpu.intel.com/i915":%s}}}]}}]' % (req, lim)
    got = k8s.marshalFleetJson("[]", text)
    assert_same_columns(got, k8s.marshalFleet([], json.loads(text)))
    assert [bool(f & K.ENTRY_SAME) for f in got.entries["flags"]] == [same, same]


def test_more_containers_than_the_ordinal_holds_is_a_range_error():
    plain = [{"name": f"c{i}"} for i in range(K.MAX_CONTAINERS)]
    ok = [{"spec": {"containers": plain[:-1] + [{"resources": {"limits": {"gpu.intel.com/xe": "1"}}}]}}]
    got = k8s.marshalFleetJson("[]", json.dumps(ok))
    assert got.entries["container"][0] == K.MAX_CONTAINERS - 1
    bad = [{"spec": {"containers": plain + [{"resources": {"limits": {"gpu.intel.com/xe": "1"}}}]}}]
    with pytest.raises(native.IgpuError) as ei:
        k8s.marshalFleetJson("[]", json.dumps(bad))
    assert ei.value.status == -6
    with pytest.raises(ValueError):
        k8s.marshalFleet([], bad)
