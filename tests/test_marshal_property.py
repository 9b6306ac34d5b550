"""CPU, property-based: the native JSON marshal (csrc/marshal.cpp) and the Python marshal (k8s.marshalFleet over
json.loads) must produce the same columns for arbitrary Kubernetes-shaped junk — weird resource values, wrong types in
every position, inherited-name phases, duplicate node names, unicode keys."""
import json

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import k8s, native

PREFIX = "gpu.intel.com/"
res_key = st.one_of(
    st.sampled_from([PREFIX + "i915", PREFIX + "xe", PREFIX + "millicores", PREFIX + "memory.max", PREFIX + "tiles", PREFIX + "späße", PREFIX,
                     "cpu", "memory", "gpu.intel.com", "nvidia.com/gpu"]),
    st.text(alphabet="abcxyz/._-é\"\\", min_size=0, max_size=6).map(lambda s: PREFIX + s),
)
weird_value = st.one_of(
    st.sampled_from(["1", "2", "8", "0", "-3", " 7", "7 ", "+4", "1e3", "500m", "1Gi", "0x10", "", "abc", "١٢", "12abc34", "-0", "00012",
                     "9007199254740993", "1.9", " 5", "﻿8", None, True, False, 0, 1, 3, -2, 2.5, 1e21, 1e-7, -0.0, 12345678901,
                     [], {}, [1], {"a": 1}]),
    st.integers(min_value=-(10**6), max_value=10**6),
    st.integers(min_value=0, max_value=10**4).map(str),
)
res_map = st.one_of(st.none(), st.sampled_from([[], "x", 3]), st.dictionaries(res_key, weird_value, max_size=4))
resources = st.one_of(st.none(), st.just("bad"), st.fixed_dictionaries({}, optional={"requests": res_map, "limits": res_map, "other": st.just(1)}))
container = st.one_of(st.none(), st.just(5), st.fixed_dictionaries({"name": st.just("c")}, optional={"resources": resources}))
containers = st.one_of(st.none(), st.just({}), st.lists(container, max_size=3))
node_names = st.sampled_from(["n0", "n1", "n2", "n3", "missing", "", "né", 'q"uote'])
phase = st.sampled_from(["Running", "Pending", "Succeeded", "Failed", "Other", "Unknown", "toString", "constructor", "__proto__", "", None, 3])
pod = st.one_of(
    st.sampled_from([None, 1, "pod", []]),
    st.fixed_dictionaries({}, optional={
        "metadata": st.fixed_dictionaries({"name": st.text(max_size=4)}),
        "spec": st.one_of(st.none(), st.just([]), st.fixed_dictionaries({}, optional={"nodeName": st.one_of(node_names, st.none(), st.just(7)),
                                                                                      "containers": containers, "initContainers": containers})),
        "status": st.one_of(st.none(), st.fixed_dictionaries({}, optional={"phase": phase})),
    }),
)
label_val = st.sampled_from(["true", "True", "false", "", "1", None, True])
node = st.fixed_dictionaries({}, optional={
    "metadata": st.fixed_dictionaries({}, optional={"name": st.one_of(node_names, st.none(), st.just(3)),
                                                    "labels": st.one_of(st.none(), st.dictionaries(st.sampled_from(
                                                        ["intel.feature.node.kubernetes.io/gpu", "node-role.kubernetes.io/gpu",
                                                         "node-role.kubernetes.io/igpu", "x"]), label_val, max_size=3))}),
    "status": st.one_of(st.none(), st.fixed_dictionaries({}, optional={
        "capacity": res_map, "allocatable": res_map,
        "conditions": st.one_of(st.none(), st.lists(st.fixed_dictionaries({}, optional={"type": st.sampled_from(["Ready", "Other"]),
                                                                                      "status": st.sampled_from(["True", "False", True])}), max_size=2))})),
})


def columns_or_error(fn):
    try:
        return fn(), None
    except (ValueError, native.IgpuError) as e:
        return None, type(e).__name__


@settings(max_examples=250, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(nodes=st.lists(node, max_size=5), pods=st.lists(pod, max_size=8), spaced=st.booleans())
def test_native_marshal_equals_python_marshal(nodes, pods, spaced):
    nj = json.dumps(nodes, indent=1 if spaced else None, ensure_ascii=not spaced)
    pj = json.dumps(pods, indent=2 if spaced else None, ensure_ascii=not spaced)
    want, werr = columns_or_error(lambda: k8s.marshalFleet(json.loads(nj), json.loads(pj)))
    got, gerr = columns_or_error(lambda: k8s.marshalFleetJson(nj, pj, threads=2))
    assert (want is None) == (got is None), (werr, gerr, nj, pj)
    if want is None:
        return
    assert want.resource_names == got.resource_names, (nj, pj)
    for f in ("node_flags", "node_cap_i915", "node_cap_xe", "node_alloc_i915", "node_alloc_xe", "node_alloc_own", "pod_node", "pod_phase",
              "pod_entry_off", "entries"):
        assert np.array_equal(getattr(want, f), getattr(got, f)), (f, nj, pj)
