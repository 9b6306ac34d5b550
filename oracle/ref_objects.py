"""Object-level CPU oracle: the reference's rollup helpers and page loops, restated in Python.

TEST INFRASTRUCTURE ONLY — not shipped, not measured as the product.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this.

Every function works on plain K8s-shaped dict objects (what JSON.parse gives the reference) and
carries numbers as Python floats, i.e. IEEE doubles exactly like JS `number`.  Each function cites
the reference lines it follows.  Pinned against the reference's own known-answer tests
(src/api/k8s.test.ts, transcribed to tests/golden/k8s_test_vectors.json) by
tests/test_oracle_golden.py.  Everything the reference has no numeric test for (page loops,
metrics.ts) follows the cited source literally: parity for those rows is "unpinned by the
reference" (DESIGN.md §oracle).
"""
from __future__ import annotations

import math
from typing import Any

from .js_semantics import (
    NaN,
    js_math_round,
    js_number_to_string,
    js_or_zero,
    js_parse_float,
    js_parse_int10,
    js_strict_equal,
    js_to_fixed1,
    js_truthy,
)

# src/api/k8s.ts:17-28
INTEL_GPU_RESOURCE = "gpu.intel.com/i915"
INTEL_GPU_XE_RESOURCE = "gpu.intel.com/xe"
INTEL_GPU_MILLICORES_RESOURCE = "gpu.intel.com/millicores"
INTEL_GPU_MEMORY_RESOURCE = "gpu.intel.com/memory.max"
INTEL_GPU_RESOURCE_PREFIX = "gpu.intel.com/"
INTEL_GPU_NODE_LABEL = "intel.feature.node.kubernetes.io/gpu"
INTEL_DISCRETE_GPU_NODE_ROLE = "node-role.kubernetes.io/gpu"
INTEL_INTEGRATED_GPU_NODE_ROLE = "node-role.kubernetes.io/igpu"

_COUNTED = (INTEL_GPU_RESOURCE, INTEL_GPU_XE_RESOURCE)

# own + inherited property names of a plain object literal (ECMA-262 Object.prototype + Annex B)
_OBJECT_PROTOTYPE_KEYS = frozenset(
    (
        "constructor", "hasOwnProperty", "isPrototypeOf", "propertyIsEnumerable", "toLocaleString", "toString",
        "valueOf", "__proto__", "__defineGetter__", "__defineSetter__", "__lookupGetter__", "__lookupSetter__",
    )
)


def _get(obj: Any, *path: str) -> Any:
    """Optional chaining a?.b?.c over dicts; anything that is not a dict yields None (undefined)."""
    cur = obj
    for p in path:
        if not isinstance(cur, dict):
            return None
        cur = cur.get(p)
    return cur


def _obj(v: Any) -> dict:
    """`v ?? {}` followed by Object.keys/entries: non-dicts have no own enumerable string keys here."""
    return v if isinstance(v, dict) else {}


def _arr(v: Any) -> list:
    return v if isinstance(v, list) else []


# --------------------------------------------------------------------------- nodes


def is_intel_gpu_node(node: Any) -> bool:
    """src/api/k8s.ts:125-152."""
    if not isinstance(node, dict):
        return False
    labels = _get(node, "metadata", "labels")
    capacity = _get(node, "status", "capacity")
    if isinstance(labels, dict) and labels:
        if (
            labels.get(INTEL_GPU_NODE_LABEL) == "true"
            or labels.get(INTEL_DISCRETE_GPU_NODE_ROLE) == "true"
            or labels.get(INTEL_INTEGRATED_GPU_NODE_ROLE) == "true"
        ):
            return True
    if isinstance(capacity, dict):
        for key in capacity:
            if key.startswith(INTEL_GPU_RESOURCE_PREFIX):
                return True
    return False


def filter_intel_gpu_nodes(items: list) -> list:
    """src/api/k8s.ts:154-156."""
    return [n for n in items if is_intel_gpu_node(n)]


def get_gpu_resources(resources: Any) -> dict:
    """src/api/k8s.ts:159-168 (a key whose value is `undefined` is dropped; JSON null is kept)."""
    if not isinstance(resources, dict):
        return {}
    return {k: v for k, v in resources.items() if k.startswith(INTEL_GPU_RESOURCE_PREFIX)}


def get_node_gpu_count(node: dict) -> float:
    """src/api/k8s.ts:171-180."""
    capacity = _obj(_get(node, "status", "capacity"))
    count = 0.0
    for key, value in capacity.items():
        if key in _COUNTED and js_truthy(value):
            count += js_or_zero(js_parse_int10(value))
    return count


def get_node_gpu_type(node: dict) -> str:
    """src/api/k8s.ts:185-192 — discrete wins over integrated; the NFD label alone is 'unknown'."""
    labels = _obj(_get(node, "metadata", "labels"))
    if labels.get(INTEL_DISCRETE_GPU_NODE_ROLE) == "true":
        return "discrete"
    if labels.get(INTEL_INTEGRATED_GPU_NODE_ROLE) == "true":
        return "integrated"
    return "unknown"


def is_node_ready(node: dict) -> bool:
    """src/api/k8s.ts:329-331."""
    for c in _arr(_get(node, "status", "conditions")):
        if isinstance(c, dict) and c.get("type") == "Ready" and c.get("status") == "True":
            return True
    return False


# --------------------------------------------------------------------------- pods


def _all_containers(pod: Any) -> list:
    return _arr(_get(pod, "spec", "containers")) + _arr(_get(pod, "spec", "initContainers"))


def is_gpu_requesting_pod(pod: Any) -> bool:
    """src/api/k8s.ts:250-264 — any prefixed key in requests or limits, value ignored."""
    if not isinstance(pod, dict):
        return False
    for c in _all_containers(pod):
        requests = _obj(_get(c, "resources", "requests"))
        limits = _obj(_get(c, "resources", "limits"))
        for k in list(requests) + list(limits):
            if k.startswith(INTEL_GPU_RESOURCE_PREFIX):
                return True
    return False


def filter_gpu_requesting_pods(items: list) -> list:
    """src/api/k8s.ts:266-268."""
    return [p for p in items if is_gpu_requesting_pod(p)]


def get_pod_gpu_requests_num(pod: dict) -> dict:
    """src/api/k8s.ts:289-299 — the numeric `totals` map before it is stringified."""
    totals: dict = {}
    for c in _all_containers(pod):
        requests = _obj(_get(c, "resources", "requests"))
        for key, value in requests.items():
            if key.startswith(INTEL_GPU_RESOURCE_PREFIX) and js_truthy(value):
                totals[key] = totals.get(key, 0.0) + js_or_zero(js_parse_int10(value))
    return totals


def get_pod_gpu_requests(pod: dict) -> dict:
    """src/api/k8s.ts:289-301 — totals as decimal strings (String(number))."""
    return {k: js_number_to_string(v) for k, v in get_pod_gpu_requests_num(pod).items()}


def get_pod_gpu_limits_num(pod: dict) -> dict:
    """North-star extension (no reference counterpart; k8s.ts:293 sums requests only): the same
    reduction as get_pod_gpu_requests_num applied to `limits`.  Parity unpinned by construction."""
    totals: dict = {}
    for c in _all_containers(pod):
        limits = _obj(_get(c, "resources", "limits"))
        for key, value in limits.items():
            if key.startswith(INTEL_GPU_RESOURCE_PREFIX) and js_truthy(value):
                totals[key] = totals.get(key, 0.0) + js_or_zero(js_parse_int10(value))
    return totals


# --------------------------------------------------------------------------- per-container views (A14)


def _gpu_containers(pod: Any) -> list:
    """src/components/PodsPage.tsx:50-54 == PodDetailSection.tsx:48-52 — MAIN containers whose merged
    `{...requests, ...limits}` has a key with the prefix (value ignored)."""
    out = []
    for c in _arr(_get(pod, "spec", "containers")):
        resources = {**_obj(_get(c, "resources", "requests")), **_obj(_get(c, "resources", "limits"))}
        if any(k.startswith(INTEL_GPU_RESOURCE_PREFIX) for k in resources):
            out.append(c)
    return out


def _gpu_key_set(requests: dict, limits: dict) -> list:
    """`new Set([...requestKeys.filter(prefix), ...limitKeys.filter(prefix)])` in iteration order
    (PodsPage.tsx:59-62, PodDetailSection.tsx:62-65)."""
    keys = [k for k in requests if k.startswith(INTEL_GPU_RESOURCE_PREFIX)]
    keys += [k for k in limits if k.startswith(INTEL_GPU_RESOURCE_PREFIX) and k not in requests]
    return keys


def gpu_container_list(pod: Any):
    """src/components/PodsPage.tsx:49-88 GpuContainerList, as data instead of JSX: None where the view prints the dash
    (:56), else one (container name, parts) per GPU container; a part is (key, 'collapsed', req) for `key: req`
    (:69-70) or (key, 'split', req ?? None, lim ?? None) for `key: req=.. lim=..` (:71-72); a key whose two values are
    both falsy yields no part.  (The display strings around them — formatGpuResourceName, the dash — are out of scope.)"""
    gpu_containers = _gpu_containers(pod)
    if len(gpu_containers) == 0:
        return None
    out = []
    for c in gpu_containers:
        requests = _obj(_get(c, "resources", "requests"))  # c.resources?.requests ?? {}
        limits = _obj(_get(c, "resources", "limits"))
        parts = []
        for key in _gpu_key_set(requests, limits):
            req, lim = requests.get(key), limits.get(key)
            if js_truthy(req) and js_truthy(lim) and js_strict_equal(req, lim):
                parts.append((key, "collapsed", req))
            elif js_truthy(req) or js_truthy(lim):
                parts.append((key, "split", req, lim))
        out.append((c.get("name") if isinstance(c, dict) else None, parts))
    return out


def pod_detail_section(pod: Any):
    """src/components/PodDetailSection.tsx:48-83,106-108 as data: None when the section renders nothing (:54), else
    {'gpuContainers': gpuContainers.length, 'rows': [...]} with one (container name, key, 'request', req ?? None) per key
    (:71-74) and a (container name, key, 'limit', lim) after it when `lim && lim !== req` (:75-80)."""
    gpu_containers = _gpu_containers(pod)
    if len(gpu_containers) == 0:
        return None
    rows = []
    for c in gpu_containers:
        requests = _obj(_get(c, "resources", "requests"))
        limits = _obj(_get(c, "resources", "limits"))
        name = c.get("name") if isinstance(c, dict) else None
        for key in _gpu_key_set(requests, limits):
            req, lim = requests.get(key), limits.get(key)
            rows.append((name, key, "request", req))
            if js_truthy(lim) and not js_strict_equal(lim, req):
                rows.append((name, key, "limit", lim))
    return {"gpuContainers": len(gpu_containers), "rows": rows}


# --------------------------------------------------------------------------- page loops


def _opt_str(d: dict, key: str, default: str):
    """`d[key] ?? default` — only undefined/null take the default."""
    v = d.get(key)
    return default if v is None else v


def overview_rollup(gpu_nodes: list, gpu_pods: list) -> dict:
    """src/components/OverviewPage.tsx:71-130 (inputs are the already-filtered context lists)."""
    discrete = integrated = unknown = 0
    total_gpu_count = 0.0
    ready = 0
    for node in gpu_nodes:  # :78-86
        t = get_node_gpu_type(node)
        if t == "discrete":
            discrete += 1
        elif t == "integrated":
            integrated += 1
        else:
            unknown += 1
        total_gpu_count += get_node_gpu_count(node)
        if is_node_ready(node):
            ready += 1

    total_capacity = 0.0
    total_allocatable = 0.0
    total_allocated = 0.0
    for node in gpu_nodes:  # :93-102 — note: no `|| 0`, NaN poisons the sums
        capacity = _obj(_get(node, "status", "capacity"))
        allocatable = _obj(_get(node, "status", "allocatable"))
        for key in capacity:
            if key in _COUNTED:
                total_capacity += js_parse_int10(_opt_str(capacity, key, "0"))
                total_allocatable += js_parse_int10(_opt_str(allocatable, key, "0"))

    for pod in gpu_pods:  # :105-113
        if _get(pod, "status", "phase") != "Running":
            continue
        for key, value in get_pod_gpu_requests(pod).items():
            if key in _COUNTED:
                total_allocated += js_or_zero(js_parse_int10(value))

    pct = js_math_round((total_allocated / total_capacity) * 100) if total_capacity > 0 else 0.0  # :115-116

    phases = {"Running": 0, "Pending": 0, "Succeeded": 0, "Failed": 0, "Other": 0}
    for pod in gpu_pods:  # :122-130
        phase = _get(pod, "status", "phase")
        if phase is None:
            phase = "Other"
        if isinstance(phase, str) and phase in phases:
            phases[phase] += 1
        elif isinstance(phase, str) and phase in _OBJECT_PROTOTYPE_KEYS:
            # JS `in` also sees inherited properties: `"toString" in podPhaseCounts` is true, so the
            # reference increments podPhaseCounts.toString (-> NaN) and none of the five buckets.
            pass
        else:
            phases["Other"] += 1

    return {
        "discreteCount": discrete,
        "integratedCount": integrated,
        "unknownCount": unknown,
        "totalGpuCount": total_gpu_count,
        "readyNodeCount": ready,
        "totalGpuNodes": len(gpu_nodes),
        "totalCapacityGpus": total_capacity,
        "totalAllocatableGpus": total_allocatable,
        "totalAllocatedGpus": total_allocated,
        "gpuUtilizationPct": pct,
        "podPhaseCounts": phases,
    }


def allocation_bar(used: float, allocatable: float):
    """src/components/NodesPage.tsx:35-38 — returns None for the '—' case, else (pct, level)."""
    if allocatable == 0:
        return None
    pct = min(100.0, js_math_round((used / allocatable) * 100))
    if pct != pct:  # Math.min(100, NaN) is NaN; comparisons below are then all false
        return (pct, 0)
    level = 2 if pct >= 90 else (1 if pct >= 70 else 0)
    return (pct, level)


def nodes_page_rollup(gpu_nodes: list, gpu_pods: list) -> list:
    """src/components/NodesPage.tsx:152-189 and the bar inputs at :270-273."""
    pods_by_node: dict = {}
    for pod in gpu_pods:  # :153-159
        name = _get(pod, "spec", "nodeName")
        if not js_truthy(name):
            continue
        pods_by_node.setdefault(name, []).append(_get(pod, "metadata", "name"))

    rows = []
    for node in gpu_nodes:  # :162-189
        capacity = _obj(_get(node, "status", "capacity"))
        allocatable = _obj(_get(node, "status", "allocatable"))
        total_cap = 0.0
        total_alloc = 0.0
        for key in capacity:
            if key in _COUNTED:
                total_cap += js_parse_int10(_opt_str(capacity, key, "0"))
                total_alloc += js_parse_int10(_opt_str(allocatable, key, "0"))
        gpu_count = get_node_gpu_count(node)
        pods_on_node = pods_by_node.get(_get(node, "metadata", "name"), [])
        bar_alloc = total_alloc if js_truthy(total_alloc) else gpu_count  # `totalAllocatable || gpuCount`
        rows.append(
            {
                "gpuType": get_node_gpu_type(node),
                "gpuCount": gpu_count,
                "ready": is_node_ready(node),
                "totalCapacity": total_cap,
                "totalAllocatable": total_alloc,
                "podCount": len(pods_on_node),
                "bar": allocation_bar(float(len(pods_on_node)), bar_alloc),
            }
        )
    return rows


def node_detail_rollup(node: dict, gpu_pods: list) -> dict:
    """src/components/NodeDetailSection.tsx:61-91 for one node."""
    name = _get(node, "metadata", "name")
    allocatable = get_gpu_resources(_get(node, "status", "allocatable"))
    pods_on_node = [p for p in gpu_pods if _get(p, "spec", "nodeName") == name and name is not None]  # :62
    gpu_in_use = 0.0
    gpu_allocatable = 0.0
    for key, val in allocatable.items():  # :72-76
        if key in _COUNTED:
            gpu_allocatable += js_or_zero(js_parse_int10(val))
    for pod in pods_on_node:  # :77-87 — Running pods, main containers only
        if _get(pod, "status", "phase") != "Running":
            continue
        for c in _arr(_get(pod, "spec", "containers")):
            for k, val in _obj(_get(c, "resources", "requests")).items():
                if k in _COUNTED:
                    gpu_in_use += js_or_zero(js_parse_int10(val))
    pct = js_math_round((gpu_in_use / gpu_allocatable) * 100) if gpu_allocatable > 0 else 0.0  # :89
    level = 2 if pct >= 90 else (1 if pct >= 70 else 0)  # :90-91 error / warning / success
    return {"gpuInUse": gpu_in_use, "gpuAllocatable": gpu_allocatable, "utilizationPct": pct, "level": level}


# --------------------------------------------------------------------------- metrics.ts


def join_gpu_metrics(chip_results: list, energy_results: list, power_max_results: list, uname_results: list) -> list:
    """src/api/metrics.ts:118-149 — three maps then one row per chip-identification result.

    The chip maps are keyed by the `chip` label only (not chip+instance): last write wins.
    """
    instance_to_node: dict = {}
    for r in uname_results:  # :120-124
        m = _obj(r.get("metric"))
        inst = m.get("instance")
        nodename = m.get("nodename")
        if nodename is None:
            nodename = m.get("node")
        if nodename is None:
            nodename = inst
        if js_truthy(inst):
            instance_to_node[inst] = nodename
    chip_to_power: dict = {}
    for r in energy_results:  # :128-131
        chip = _obj(r.get("metric")).get("chip")
        if js_truthy(chip):
            chip_to_power[chip] = js_parse_float(r["value"][1])
    chip_to_max: dict = {}
    for r in power_max_results:  # :135-138
        chip = _obj(r.get("metric")).get("chip")
        if js_truthy(chip):
            chip_to_max[chip] = js_parse_float(r["value"][1])
    chips = []
    for r in chip_results:  # :141-149
        m = _obj(r.get("metric"))
        chip = m.get("chip")
        chip = "" if chip is None else chip
        instance = m.get("instance")
        instance = "" if instance is None else instance
        node_name = instance_to_node.get(instance)
        if node_name is None:
            node_name = instance
        chips.append(
            {
                "nodeName": node_name,
                "chip": chip,
                "instance": instance,
                "powerWatts": chip_to_power.get(chip) if chip in chip_to_power else None,
                "powerMaxWatts": chip_to_max.get(chip) if chip in chip_to_max else None,
            }
        )
    return chips


def fleet_power(chips: list) -> dict:
    """src/components/MetricsPage.tsx:330-332 with PowerBar :50-53."""
    total = 0.0
    max_total = 0.0
    for c in chips:  # left-to-right reduce; `?? 0` only replaces null
        total += 0.0 if c["powerWatts"] is None else c["powerWatts"]
    for c in chips:
        max_total += 0.0 if c["powerMaxWatts"] is None else c["powerMaxWatts"]
    max_watts = max_total if max_total > 0 else None
    pct = None
    if max_watts is not None and max_watts > 0:
        pct = min(100.0, js_math_round((total / max_watts) * 100))
    return {"totalWatts": total, "maxTotalWatts": max_total, "pct": pct}


def format_watts(w: float) -> str:
    """src/api/metrics.ts:161-163."""
    return f"{js_to_fixed1(w)} W"


def format_percent(used: float, maximum: float) -> str:
    """src/api/metrics.ts:165-168."""
    if maximum <= 0:
        return "—"
    return f"{js_number_to_string(js_math_round((used / maximum) * 100))}%"


__all__ = [n for n in dir() if not n.startswith("_")]
_ = (NaN, math)
