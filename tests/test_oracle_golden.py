"""Pins the object-level oracle to the reference's own known-answer tests (src/api/k8s.test.ts,
transcribed in tests/golden/k8s_test_vectors.json) and cross-checks the two independent
implementations of the JS number built-ins (oracle/js_semantics.py vs the product's jsnum.py)."""
import json
import math
import os
import random

import pytest

from headlamp_intel_gpu_plugin_b200 import jsnum
from oracle import js_semantics as js
from oracle import ref_objects as ro

HERE = os.path.dirname(os.path.abspath(__file__))
VECTORS = json.load(open(os.path.join(HERE, "golden", "k8s_test_vectors.json")))["vectors"]

FN = {
    "isIntelGpuNode": ro.is_intel_gpu_node,
    "filterIntelGpuNodes": lambda items: [n["metadata"]["name"] for n in ro.filter_intel_gpu_nodes(items)],
    "getNodeGpuType": ro.get_node_gpu_type,
    "getNodeGpuCount": ro.get_node_gpu_count,
    "isNodeReady": ro.is_node_ready,
    "isGpuRequestingPod": ro.is_gpu_requesting_pod,
    "filterGpuRequestingPods": lambda items: [p["metadata"]["name"] for p in ro.filter_gpu_requesting_pods(items)],
    "getPodGpuRequests": ro.get_pod_gpu_requests,
}


@pytest.mark.parametrize("v", VECTORS, ids=[f'{v["fn"]}@{v["ref"].split(":")[-1]}' for v in VECTORS])
def test_reference_known_answers(v):
    assert FN[v["fn"]](v["input"]) == v["expect"], v["ref"]


def test_all_hot_path_helpers_have_vectors():
    assert {v["fn"] for v in VECTORS} == set(FN)
    assert len(VECTORS) >= 29


def same(a, b):
    return a == b or (a != a and b != b)


PARSE_CASES = ["500m", "1Gi", "1e3", "0x10", "", " 42", "-0", "+7x", "abc", "9007199254740993", "--1", "+-1", "1 2", ".5", "5.",
               "1e", "-.5e-1x", "  3.25abc", "e5", ".", "Infinity", "-Infinity", "+Inf", "NaN", "1.5e+02", "1e400", "1.e3",
               None, 3, 2.5, True, False, " 7", "﻿8", " 9", "١٢", "-", "+", "0", "-0.0", "00012", "12abc34"]


def test_parse_int_semantics():
    # the JS rules the restatement must reproduce exactly
    assert js.js_parse_int10("500m") == 500 and js.js_parse_int10("1Gi") == 1 and js.js_parse_int10("1e3") == 1
    assert js.js_parse_int10("0x10") == 0 and math.isnan(js.js_parse_int10("")) and js.js_parse_int10(" 42") == 42
    assert math.copysign(1, js.js_parse_int10("-0")) == -1 and js.js_or_zero(js.js_parse_int10("-0")) == 0
    assert math.isnan(js.js_parse_float("NaN")) and math.isnan(js.js_parse_float("+Inf")) and js.js_parse_float("Infinity") == math.inf
    for c in PARSE_CASES:
        assert same(js.js_parse_int10(c), jsnum.parse_int(c)), c
        assert same(js.js_parse_float(c), jsnum.parse_float(c)), c
        assert js.js_truthy(c) == jsnum.truthy(c), c


def test_round_tostring_tofixed():
    assert js.js_math_round(2.5) == 3 and js.js_math_round(-2.5) == -2 and js.js_math_round(0.49999999999999994) == 0
    assert js.js_to_fixed1(45.25) == "45.3" and js.js_to_fixed1(0.05) == "0.1" and js.js_to_fixed1(-1.25) == "-1.3"
    assert js.js_number_to_string(1e21) == "1e+21" and js.js_number_to_string(123456789012345680000.0) == "123456789012345680000"
    assert js.js_number_to_string(1e-7) == "1e-7" and js.js_number_to_string(0.000001) == "0.000001"
    rng = random.Random(3)
    for _ in range(5000):
        x = rng.choice([rng.uniform(-1e3, 1e3), rng.uniform(-1, 1) * 10 ** rng.randint(-10, 25), rng.randint(-10**6, 10**6) / 20])
        assert js.js_number_to_string(x) == jsnum.number_to_string(x)
        assert js.js_to_fixed1(x) == jsnum.to_fixed_1(x)
        assert js.js_math_round(x) == jsnum.math_round(x)


def test_formatters_follow_metrics_ts():
    # src/components/MetricsPage.test.tsx:203-213 expects "45.3 W"; formatters: metrics.ts:161-168
    assert ro.format_watts(45.3) == "45.3 W" and ro.format_watts(120) == "120.0 W"
    assert ro.format_percent(45.3, 120) == "38%" and ro.format_percent(1, 0) == "—" and ro.format_percent(1, -5) == "—"


POD_DETAIL = json.load(open(os.path.join(HERE, "golden", "pod_detail_vectors.json")))["vectors"]


@pytest.mark.parametrize("v", POD_DETAIL, ids=[v["ref"].split(":")[-1] for v in POD_DETAIL])
def test_pod_detail_section_fixtures(v):
    """PodDetailSection.test.tsx: the empty render, "GPU Containers" = 1, the limits-only pod; and the same containers
    through PodsPage's GpuContainerList (PodsPage.tsx:49-88)."""
    got = ro.pod_detail_section(v["pod"])
    assert (got is None) == v["rendersNothing"], v["ref"]
    assert (0 if got is None else got["gpuContainers"]) == v["gpuContainers"], v["ref"]
    if got is not None:
        assert [list(r) for r in got["rows"]] == v["rows"]
    lst = ro.gpu_container_list(v["pod"])
    assert (lst is None) == v["rendersNothing"] and (lst is None or len(lst) == v["gpuContainers"])


def test_gpu_container_list_forms():
    """PodsPage.tsx:64-73: `key: v` when both values are truthy and strictly equal, `req=.. lim=..` when either is truthy,
    nothing when neither; Set order = request keys, then limit keys not seen yet; init containers never listed."""
    i915, xe, mc = "gpu.intel.com/i915", "gpu.intel.com/xe", "gpu.intel.com/millicores"
    pod = {"spec": {"containers": [
        {"name": "a", "resources": {"requests": {xe: "2", i915: "1", "cpu": "1"}, "limits": {mc: "500", i915: "1", xe: "02"}}},
        {"name": "plain", "resources": {"requests": {"cpu": "1"}}},
        {"name": "b", "resources": {"requests": {i915: ""}, "limits": {i915: 0, xe: None}}}],
        "initContainers": [{"name": "init", "resources": {"limits": {i915: "1"}}}]}}
    assert ro.gpu_container_list(pod) == [("a", [(xe, "split", "2", "02"), (i915, "collapsed", "1"), (mc, "split", None, "500")]), ("b", [])]
    d = ro.pod_detail_section(pod)
    assert d["gpuContainers"] == 2
    assert d["rows"] == [("a", xe, "request", "2"), ("a", xe, "limit", "02"), ("a", i915, "request", "1"), ("a", mc, "request", None),
                         ("a", mc, "limit", "500"), ("b", i915, "request", ""), ("b", xe, "request", None)]
    assert ro.gpu_container_list({"spec": {"initContainers": [{"name": "i", "resources": {"limits": {i915: "1"}}}]}}) is None
    assert js.js_strict_equal("1", "1") and not js.js_strict_equal("1", 1) and js.js_strict_equal(1, 1.0) and not js.js_strict_equal([], [])
