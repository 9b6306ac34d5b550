"""Shared comparison helpers: columnar outputs (from the C++ oracle or from the CUDA library) versus
the object-level restatement of the reference (oracle/ref_objects.py)."""
import math

import numpy as np

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import k8s
from oracle import ref_objects as ro
from oracle.js_semantics import js_number_to_string


def nan_eq(a, b):
    if isinstance(a, dict) and isinstance(b, dict):
        return a.keys() == b.keys() and all(nan_eq(a[k], b[k]) for k in a)
    if isinstance(a, (list, tuple)) and isinstance(b, (list, tuple)):
        return len(a) == len(b) and all(nan_eq(x, y) for x, y in zip(a, b))
    if isinstance(a, float) and isinstance(b, float) and math.isnan(a) and math.isnan(b):
        return True
    return a == b


def assert_outputs_equal(a: K.FleetOutputs, b: K.FleetOutputs, what=""):
    for f in K.FleetOutputs.FIELDS:
        x, y = getattr(a, f), getattr(b, f)
        if x is None or y is None:
            assert x is None and y is None, f
            continue
        assert np.array_equal(x, y), f"{what}: column {f} differs at {np.argwhere(x != y)[:5].tolist()}"


def check_against_objects(nodes, pods, cols: K.FleetColumns, out: K.FleetOutputs, order=None, pos=None, container_rows=None):
    """Every reference-visible quantity, recomputed per object by the literal restatement.  With `container_rows`
    (igpu_fleet_container_rows, or the columnar oracle's twin) the per-container views are checked too."""
    r = k8s.FleetRollup(cols, out, order=order, pos=pos, container_rows=container_rows)
    gpu_nodes = ro.filter_intel_gpu_nodes(nodes)
    gpu_pods = ro.filter_gpu_requesting_pods(pods)
    # k8s.ts level
    assert [nodes[i] for i in r.gpu_node_rows()] == gpu_nodes
    assert [pods[i] for i in r.gpu_pod_rows()] == gpu_pods
    for i in r.gpu_node_rows():
        n = nodes[i]
        assert r.node_type(i) == ro.get_node_gpu_type(n)
        assert float(out.node_gpu_count[i]) == ro.get_node_gpu_count(n)
        assert bool(out.node_out_flags[i] & K.NODEOUT_READY) == ro.is_node_ready(n)
    for p, pod in enumerate(pods):
        if not isinstance(pod, dict):
            continue
        assert r.pod_requests(p) == ro.get_pod_gpu_requests(pod), (p, pod)
        if out.pod_lim is not None:
            want = {k: js_number_to_string(v) for k, v in ro.get_pod_gpu_limits_num(pod).items()}
            assert r.pod_limits(p) == want
        # PodDetailSection.tsx:48-52,106-108 "GPU Containers"; PodsPage.tsx:49-88; PodDetailSection.tsx:57-83
        detail = ro.pod_detail_section(pod)
        assert r.pod_gpu_container_count(p) == (0 if detail is None else detail["gpuContainers"]), (p, pod)
        if container_rows is not None:
            assert r.gpu_container_list(p, pod) == ro.gpu_container_list(pod), (p, pod)
            assert r.pod_detail_section(p, pod) == detail, (p, pod)
    # page level
    assert nan_eq(r.overview(), ro.overview_rollup(gpu_nodes, gpu_pods))
    assert nan_eq(r.nodes_page_rows(), ro.nodes_page_rollup(gpu_nodes, gpu_pods))
    for i in r.gpu_node_rows():
        assert nan_eq(r.node_detail(i), ro.node_detail_rollup(nodes[i], gpu_pods)), i
