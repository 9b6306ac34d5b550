"""Host side of the allocation rollup: the exported surface of the reference's src/api/k8s.ts, kept
name for name, with the arithmetic moved behind the C ABI (include/igpu_rollup.h).

Two layers:

* marshal — `marshalFleet(nodes, pods)` turns K8s JSON objects into the columnar typed arrays of
  columns.py: strings are dictionary-encoded to integer ids here, and parseInt / truthiness /
  `=== 'true'` are applied here, exactly as the reference applies them per object.
* frozen signatures — `isIntelGpuNode`, `getNodeGpuType`, `getNodeGpuCount`, `isNodeReady`,
  `isGpuRequestingPod`, `getPodGpuRequests`, the two filters, plus the added batch entry point
  `rollupFleet(nodes, pods)` that the data context would call once per refresh.  Every one of them
  runs on the GPU through libigpu_rollup.so; none has a CPU fallback (native.NativeUnavailable).

The reference's presentation helpers (formatGpuType, formatGpuResourceName, formatAge, pluginStatus*; k8s.ts:194-203,
337-386) are string formatting for the pages: out of scope, they stay TypeScript and have no twin here.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

from . import columns as K
from . import jsnum
from . import native

# -- constants, src/api/k8s.ts:13-31 ---------------------------------------------------------------
INTEL_DEVICE_PLUGIN_API_GROUP = "deviceplugin.intel.com"
INTEL_DEVICE_PLUGIN_API_VERSION = "v1"
INTEL_GPU_RESOURCE = "gpu.intel.com/i915"
INTEL_GPU_XE_RESOURCE = "gpu.intel.com/xe"
INTEL_GPU_MILLICORES_RESOURCE = "gpu.intel.com/millicores"
INTEL_GPU_MEMORY_RESOURCE = "gpu.intel.com/memory.max"
INTEL_GPU_RESOURCE_PREFIX = "gpu.intel.com/"
INTEL_GPU_NODE_LABEL = "intel.feature.node.kubernetes.io/gpu"
INTEL_DISCRETE_GPU_NODE_ROLE = "node-role.kubernetes.io/gpu"
INTEL_INTEGRATED_GPU_NODE_ROLE = "node-role.kubernetes.io/igpu"
INTEL_GPU_PLUGIN_LABEL_SELECTOR = "app=intel-gpu-plugin"

# resource dictionary seed: ids 0..3 are fixed by the C ABI; 'tiles' is the fifth name the
# reference knows (formatGpuResourceName, k8s.ts:356-362)
DEFAULT_RESOURCE_NAMES = (
    INTEL_GPU_RESOURCE,
    INTEL_GPU_XE_RESOURCE,
    INTEL_GPU_MILLICORES_RESOURCE,
    INTEL_GPU_MEMORY_RESOURCE,
    "gpu.intel.com/tiles",
)

# `phase in podPhaseCounts` (OverviewPage.tsx:125) is also true for inherited property names
_OBJECT_PROTOTYPE_KEYS = frozenset(
    "constructor __defineGetter__ __defineSetter__ hasOwnProperty __lookupGetter__ __lookupSetter__ "
    "isPrototypeOf propertyIsEnumerable toString valueOf __proto__ toLocaleString".split()
)
_PHASE_CODE = {n: i for i, n in enumerate(K.PHASE_NAMES)}


def _dict(v) -> dict:
    return v if isinstance(v, dict) else {}


def _list(v) -> list:
    return v if isinstance(v, list) else []


def _int_or_zero(v) -> int:
    """parseInt(v, 10) || 0 as a Python int."""
    x = jsnum.parse_int(v)
    if x != x or x == 0:
        return 0
    if abs(x) >= jsnum.EXACT_INT_LIMIT:
        raise ValueError(f"quantity {v!r} is outside the exact-integer domain (|x| < 2^53) of the GPU path")
    return int(x)


def _fits_i32(x: int, what: str) -> int:
    if not (-(1 << 31) <= x < (1 << 31)):
        raise ValueError(f"{what} = {x} does not fit the int32 node columns")
    return x


# --------------------------------------------------------------------------------------- marshal


def marshalNodes(nodes: Sequence[Any]) -> Dict[str, np.ndarray]:
    """Node objects -> node columns.  Field semantics: include/igpu_rollup.h (igpu_fleet_cols)."""
    n = len(nodes)
    flags = np.zeros(n, np.uint8)
    cols = {k: np.zeros(n, np.int32) for k in ("cap_i915", "cap_xe", "alloc_i915", "alloc_xe", "alloc_own")}
    for i, node in enumerate(nodes):
        if not isinstance(node, dict):
            continue
        labels = _dict(_dict(node.get("metadata")).get("labels"))
        status = _dict(node.get("status"))
        capacity = _dict(status.get("capacity"))
        allocatable = _dict(status.get("allocatable"))
        f = 0
        if labels.get(INTEL_DISCRETE_GPU_NODE_ROLE) == "true":
            f |= K.NODE_DISCRETE_ROLE
        if labels.get(INTEL_INTEGRATED_GPU_NODE_ROLE) == "true":
            f |= K.NODE_IGPU_ROLE
        if labels.get(INTEL_GPU_NODE_LABEL) == "true":
            f |= K.NODE_NFD_LABEL
        if any(isinstance(k, str) and k.startswith(INTEL_GPU_RESOURCE_PREFIX) for k in capacity):
            f |= K.NODE_HAS_GPU_CAPACITY
        for c in _list(status.get("conditions")):
            if isinstance(c, dict) and c.get("type") == "Ready" and c.get("status") == "True":
                f |= K.NODE_READY
                break
        for key, short in ((INTEL_GPU_RESOURCE, "i915"), (INTEL_GPU_XE_RESOURCE, "xe")):
            if key in capacity:
                v = capacity[key]
                # getNodeGpuCount (k8s.ts:175-176): `value && (parseInt(value) || 0)`
                cols["cap_" + short][i] = _fits_i32(_int_or_zero(v) if jsnum.truthy(v) else 0, key)
                # Overview/Nodes sums (OverviewPage.tsx:98-99): parseInt(x ?? '0') with no `|| 0`
                if math.isnan(jsnum.parse_int("0" if v is None else v)):
                    f |= K.NODE_CAP_NAN
                a = allocatable.get(key)
                pa = jsnum.parse_int("0" if a is None else a)
                if math.isnan(pa):
                    f |= K.NODE_ALLOC_NAN
                else:
                    cols["alloc_" + short][i] = _fits_i32(_int_or_zero(a if a is not None else "0"), key)
            if key in allocatable:  # NodeDetailSection.tsx:72-76 over getGpuResources(allocatable)
                cols["alloc_own"][i] = _fits_i32(int(cols["alloc_own"][i]) + _int_or_zero(allocatable[key]), key)
        flags[i] = f
    return {
        "node_flags": flags,
        "node_cap_i915": cols["cap_i915"],
        "node_cap_xe": cols["cap_xe"],
        "node_alloc_i915": cols["alloc_i915"],
        "node_alloc_xe": cols["alloc_xe"],
        "node_alloc_own": cols["alloc_own"],
    }


def _phase_code(pod: dict) -> int:
    phase = _dict(pod.get("status")).get("phase")
    if phase is None:
        return K.PHASE_OTHER
    if isinstance(phase, str):
        if phase in _PHASE_CODE:
            return _PHASE_CODE[phase]
        if phase in _OBJECT_PROTOTYPE_KEYS:
            return K.PHASE_PROTO
    return K.PHASE_OTHER


def marshalFleet(
    nodes: Sequence[Any],
    pods: Sequence[Any],
    resource_names: Optional[Sequence[str]] = None,
    options: int = K.OPT_POD_LIMITS,
) -> K.FleetColumns:
    """K8s Node and Pod objects (as JSON.parse / json.loads delivers them) -> FleetColumns.

    All nodes and all pods of the cluster may be passed: the kernels apply the isIntelGpuNode /
    isGpuRequestingPod filters themselves (a pod without any 'gpu.intel.com/' key simply owns no
    entries).  Pods keep their input order; entries are emitted pod by pod.
    """
    res_ids: Dict[str, int] = {n: i for i, n in enumerate(resource_names or DEFAULT_RESOURCE_NAMES)}
    node_cols = marshalNodes(nodes)
    node_names: List[Optional[str]] = []
    node_index: Dict[str, int] = {}
    for i, node in enumerate(nodes):
        name = _dict(_dict(node).get("metadata")).get("name") if isinstance(node, dict) else None
        node_names.append(name if isinstance(name, str) else None)
        if isinstance(name, str) and name and name not in node_index:
            node_index[name] = i

    p = len(pods)
    pod_node = np.full(p, K.NODE_NONE, np.uint32)
    pod_phase = np.full(p, K.PHASE_OTHER, np.uint8)
    pod_off = np.zeros(p + 1, np.uint32)
    e_pod: List[int] = []
    e_res: List[int] = []
    e_flags: List[int] = []
    e_cont: List[int] = []
    e_val: List[int] = []
    pod_names: List[Optional[str]] = []
    for i, pod in enumerate(pods):
        if isinstance(pod, dict):
            spec = _dict(pod.get("spec"))
            pod_names.append(_dict(pod.get("metadata")).get("name"))
            pod_phase[i] = _phase_code(pod)
            node_name = spec.get("nodeName")
            if isinstance(node_name, str) and node_name:
                pod_node[i] = node_index.get(node_name, K.NODE_NONE)
            for init, clist in ((0, _list(spec.get("containers"))), (K.ENTRY_INIT, _list(spec.get("initContainers")))):
                for ci, c in enumerate(clist):
                    resources = _dict(_dict(c).get("resources"))
                    req_map, lim_map = _dict(resources.get("requests")), _dict(resources.get("limits"))
                    for lim, rmap, other in ((0, req_map, lim_map), (K.ENTRY_LIMIT, lim_map, req_map)):
                        for key, value in rmap.items():
                            if not (isinstance(key, str) and key.startswith(INTEL_GPU_RESOURCE_PREFIX)):
                                continue
                            if ci >= K.MAX_CONTAINERS:
                                raise ValueError(f"a pod lists more than {K.MAX_CONTAINERS} containers")
                            rid = res_ids.get(key)
                            if rid is None:
                                rid = res_ids[key] = len(res_ids)
                                if rid >= K.MAX_RESOURCES:
                                    raise ValueError(f"more than {K.MAX_RESOURCES} distinct gpu.intel.com/ resource names")
                            same = K.ENTRY_SAME if key in other and jsnum.strict_equal(value, other[key]) else 0
                            e_pod.append(i)
                            e_res.append(rid)
                            e_cont.append(ci)
                            e_flags.append(init | lim | same | (K.ENTRY_TRUTHY if jsnum.truthy(value) else 0))
                            e_val.append(_int_or_zero(value))
        else:
            pod_names.append(None)
        pod_off[i + 1] = len(e_pod)
    entries = np.zeros(len(e_pod), K.ENTRY_DTYPE)
    entries["pod"] = e_pod
    entries["res"] = e_res
    entries["flags"] = e_flags
    entries["container"] = e_cont
    entries["value"] = e_val
    names = [None] * len(res_ids)
    for k, v in res_ids.items():
        names[v] = k
    cols = K.FleetColumns(
        pod_node=pod_node,
        pod_phase=pod_phase,
        pod_entry_off=pod_off,
        entries=entries,
        n_resources=max(4, len(names)),
        options=options,
        resource_names=names,
        node_names=node_names,
        pod_names=pod_names,
        **node_cols,
    )
    cols.validate()
    return cols


def marshalFleetJson(nodes_json, pods_json, options: int = K.OPT_POD_LIMITS, threads: int = 0) -> K.FleetColumns:
    """The same columns as marshalFleet, straight from the JSON text of the two List responses, parsed by
    the native multithreaded extractor (csrc/marshal.cpp) instead of json.loads + Python loops.
    (`.grouped_by_node()` on the result gives the device-friendly pod order, see rollupFleet.)"""
    nj = nodes_json.encode() if isinstance(nodes_json, str) else bytes(nodes_json)
    pj = pods_json.encode() if isinstance(pods_json, str) else bytes(pods_json)
    return native.marshal_fleet_json(nj, pj, options, threads)


# ---------------------------------------------------------------------------------- result views


@dataclass
class FleetRollup:
    """Decoded outputs of one rollup, shaped like the values the reference's pages compute."""

    cols: K.FleetColumns
    out: K.FleetOutputs
    # Set when the columns were regrouped by node for the device (FleetColumns.grouped_by_node): `order[i]` is the
    # caller's index of the pod in column position i, `pos[p]` the column position of the caller's pod p.  Pod indices
    # taken and returned by the views below are always the caller's.
    order: Optional[np.ndarray] = None
    pos: Optional[np.ndarray] = None
    # igpu_fleet_container_rows, when asked for (container_rows=True): one record per GPU main container, (pod, container) order
    container_rows: Optional[np.ndarray] = None

    def _pos(self, p: int) -> int:
        return p if self.pos is None else int(self.pos[p])

    # -- k8s.ts level
    def gpu_node_rows(self) -> np.ndarray:
        return np.nonzero(self.out.node_out_flags & K.NODEOUT_IS_GPU)[0]

    def gpu_pod_rows(self) -> np.ndarray:
        rows = np.nonzero(self.out.pod_is_gpu)[0]
        return rows if self.order is None else np.sort(self.order[rows])

    def node_type(self, i: int) -> str:
        return K.TYPE_NAMES[int(self.out.node_type[i])]

    def pod_requests(self, p: int) -> Dict[str, str]:
        """getPodGpuRequests(pod): key -> decimal string (k8s.ts:300)."""
        p = self._pos(p)
        mask = int(self.out.pod_req_mask[p])
        names = self.cols.resource_names
        return {names[r]: str(int(self.out.pod_req[r, p])) for r in range(self.cols.n_resources) if mask >> r & 1}

    def pod_limits(self, p: int) -> Dict[str, str]:
        p = self._pos(p)
        mask = int(self.out.pod_lim_mask[p])
        names = self.cols.resource_names
        return {names[r]: str(int(self.out.pod_lim[r, p])) for r in range(self.cols.n_resources) if mask >> r & 1}

    # -- page level
    def overview(self) -> dict:
        """The numbers OverviewPage.tsx:71-130 derives."""
        t = self.out.fleet_totals
        nan = float("nan")
        return {
            "discreteCount": int(t[K.T_DISCRETE]),
            "integratedCount": int(t[K.T_INTEGRATED]),
            "unknownCount": int(t[K.T_UNKNOWN]),
            "totalGpuCount": float(t[K.T_GPU_DEVICES]),
            "readyNodeCount": int(t[K.T_READY_NODES]),
            "totalGpuNodes": int(t[K.T_GPU_NODES]),
            "totalCapacityGpus": nan if t[K.T_CAP_NAN] else float(t[K.T_CAPACITY]),
            "totalAllocatableGpus": nan if t[K.T_ALLOC_NAN] else float(t[K.T_ALLOCATABLE]),
            "totalAllocatedGpus": float(t[K.T_ALLOCATED]),
            "gpuUtilizationPct": float(t[K.T_UTIL_PCT]),
            "podPhaseCounts": {n: int(t[K.T_PHASE0 + i]) for i, n in enumerate(K.PHASE_NAMES)},
        }

    def nodes_page_rows(self) -> List[dict]:
        """One row per GPU node, NodesPage.tsx:162-189 + the allocation bar (:35-38, :270-273)."""
        o, nan = self.out, float("nan")
        rows = []
        for i in self.gpu_node_rows():
            fl = int(o.node_out_flags[i])
            rows.append(
                {
                    "gpuType": K.TYPE_NAMES[int(o.node_type[i])],
                    "gpuCount": float(o.node_gpu_count[i]),
                    "ready": bool(fl & K.NODEOUT_READY),
                    "totalCapacity": nan if fl & K.NODEOUT_CAP_NAN else float(o.node_cap_total[i]),
                    "totalAllocatable": nan if fl & K.NODEOUT_ALLOC_NAN else float(o.node_alloc_total[i]),
                    "podCount": int(o.node_pod_count[i]),
                    "bar": None if fl & K.NODEOUT_BAR_DASH else (float(o.node_bar_pct[i]), int(o.node_bar_level[i])),
                }
            )
        return rows

    # -- per-container views (PodsPage.tsx:49-88, PodDetailSection.tsx:48-83,106-108)
    def pod_gpu_container_count(self, p: int) -> int:
        """gpuContainers.length — what PodDetailSection prints as "GPU Containers" (:106-108)."""
        return int(self.out.pod_gpu_containers[self._pos(p)])

    def _rows_of(self, p: int) -> np.ndarray:
        if self.container_rows is None:
            raise native.IgpuError(-7, "this rollup was made without container rows: rollupFleet(..., container_rows=True)")
        pos = self._pos(p)
        lo, hi = np.searchsorted(self.container_rows["pod"], [pos, pos + 1])
        return self.container_rows[lo:hi]

    def _container_keys(self, p: int, row) -> List[str]:
        """Keys of one container in the order `new Set([...requests keys, ...limits keys])` iterates: the order of the
        container's entries, first occurrence (igpu_rollup.h, igpu_entry)."""
        pos = self._pos(p)
        e = self.cols.entries[int(self.cols.pod_entry_off[pos]):int(self.cols.pod_entry_off[pos + 1])]
        mine = e[((e["flags"] & K.ENTRY_INIT) == 0) & (e["container"] == row["container"])]
        seen, keys = set(), []
        for r in mine["res"]:
            if int(r) not in seen:
                seen.add(int(r))
                keys.append(int(r))
        return keys

    def gpu_container_list(self, p: int, pod: dict):
        """GpuContainerList of PodsPage.tsx:49-88 as data: None where the page prints the dash, else one
        (container name, parts) per GPU container, a part being (key, 'collapsed', req) or (key, 'split', req, lim).
        Which containers, which keys and which form come from the device rows (masks); the raw request / limit values
        and the container name are read from `pod`, the object the caller already holds (strings never cross the ABI)."""
        rows = self._rows_of(p)
        if len(rows) == 0:
            return None
        names, out = self.cols.resource_names, []
        containers = _list(_dict(pod.get("spec")).get("containers"))
        for row in rows:
            c = _dict(containers[int(row["container"])])
            res = _dict(c.get("resources"))
            requests, limits = _dict(res.get("requests")), _dict(res.get("limits"))
            collapsed = int(row["req_truthy"]) & int(row["lim_truthy"]) & int(row["same"])
            shown = (int(row["req_truthy"]) | int(row["lim_truthy"])) & ~collapsed
            parts = []
            for r in self._container_keys(p, row):
                key = names[r]
                if collapsed >> r & 1:
                    parts.append((key, "collapsed", requests.get(key)))
                elif shown >> r & 1:
                    parts.append((key, "split", requests.get(key), limits.get(key)))
            out.append((c.get("name"), parts))
        return out

    def pod_detail_section(self, p: int, pod: dict):
        """The data of PodDetailSection.tsx:48-83,106-108: None when the section renders nothing, else
        {'gpuContainers': n, 'rows': [(container, key, 'request', req), (container, key, 'limit', lim)?, ...]}."""
        rows = self._rows_of(p)
        if len(rows) == 0:
            return None
        names, out = self.cols.resource_names, []
        containers = _list(_dict(pod.get("spec")).get("containers"))
        for row in rows:
            c = _dict(containers[int(row["container"])])
            res = _dict(c.get("resources"))
            requests, limits = _dict(res.get("requests")), _dict(res.get("limits"))
            limit_row = int(row["lim_truthy"]) & ~int(row["same"])
            for r in self._container_keys(p, row):
                key = names[r]
                out.append((c.get("name"), key, "request", requests.get(key)))
                if limit_row >> r & 1:
                    out.append((c.get("name"), key, "limit", limits.get(key)))
        return {"gpuContainers": self.pod_gpu_container_count(p), "rows": out}

    def node_detail(self, i: int) -> dict:
        """NodeDetailSection.tsx:69-91 for node row i."""
        o = self.out
        return {
            "gpuInUse": float(o.node_inuse[i]),
            "gpuAllocatable": float(self.cols.node_alloc_own[i]),
            "utilizationPct": float(o.node_util_pct[i]),
            "level": int(o.node_util_level[i]),
        }


# Below this many pods regrouping costs more on the host than it saves on the device (the call is launch-bound anyway).
GROUP_BY_NODE_MIN_PODS = 50_000


def rollupFleet(nodes: Sequence[Any], pods: Sequence[Any], ctx: Optional[native.Context] = None,
                group_by_node: Optional[bool] = None, container_rows: bool = False) -> FleetRollup:
    """Added batch entry point: one marshal + one GPU rollup for the whole fleet.

    group_by_node: lay the pod columns out with the pods of each node adjacent, so the device reduces a node's pods in
    a warp and issues one set of atomics per run instead of one per pod.  Results are identical; the views of the
    returned FleetRollup keep speaking the caller's pod indices.  Default: only for large fleets."""
    cols = marshalFleet(nodes, pods)
    ctx = ctx or native.default_context()
    if group_by_node is None:
        group_by_node = cols.n_pods >= GROUP_BY_NODE_MIN_PODS
    if group_by_node:
        cols, _ = cols.grouped_by_node()
    return _rollup_columns(cols, ctx, container_rows)


def _rollup_columns(cols: K.FleetColumns, ctx: native.Context, container_rows: bool = False) -> FleetRollup:
    rows = None
    if container_rows:  # the staged form: the rows are read off the device-resident fleet after the rollup
        fleet = ctx.fleet_create(max(cols.n_nodes, 1), max(cols.n_pods, 1), max(cols.n_entries, 1), cols.n_resources, cols.options)
        try:
            fleet.upload(cols)
            fleet.rollup()
            out = fleet.download()
            rows = fleet.container_rows()
        finally:
            fleet.destroy()
    else:
        out = ctx.alloc_rollup(cols)
    if cols.pod_order is None:
        return FleetRollup(cols, out, container_rows=rows)
    order = np.asarray(cols.pod_order, np.int64)
    pos = np.empty_like(order)
    pos[order] = np.arange(len(order))
    return FleetRollup(cols, out, order=order, pos=pos, container_rows=rows)


def rollupFleetJson(nodes_json, pods_json, ctx: Optional[native.Context] = None, group_by_node: Optional[bool] = None,
                    threads: int = 0, container_rows: bool = False) -> FleetRollup:
    """rollupFleet from the TEXT of the two List responses: native marshal (csrc/marshal.cpp, grouping included) + one
    GPU rollup.  No json.loads, no per-object Python."""
    ctx = ctx or native.default_context()
    pj = pods_json.encode() if isinstance(pods_json, str) else bytes(pods_json)
    if group_by_node is None:
        group_by_node = len(pj) >= GROUP_BY_NODE_MIN_PODS * 300  # ~300 bytes of JSON per pod at the least
    opts = K.OPT_POD_LIMITS | (K.OPT_GROUP_PODS_BY_NODE if group_by_node else 0)
    return _rollup_columns(marshalFleetJson(nodes_json, pj, opts, threads), ctx, container_rows)


# ------------------------------------------------------------- frozen per-object signatures


def isIntelGpuNode(node: Any) -> bool:
    """k8s.ts:125-152."""
    if not node or not isinstance(node, dict):
        return False
    return bool(rollupFleet([node], []).out.node_out_flags[0] & K.NODEOUT_IS_GPU)


def filterIntelGpuNodes(items: Sequence[Any]) -> list:
    """k8s.ts:154-156 — one batch for the whole list."""
    if not items:
        return []
    r = rollupFleet(list(items), [])
    return [items[i] for i in r.gpu_node_rows()]


def getGpuResources(resources: Any) -> Dict[str, Any]:
    """k8s.ts:159-168 (pure key filter, no arithmetic: stays on the host)."""
    if not isinstance(resources, dict):
        return {}
    return {k: v for k, v in resources.items() if isinstance(k, str) and k.startswith(INTEL_GPU_RESOURCE_PREFIX)}


def getNodeGpuCount(node: Any) -> int:
    """k8s.ts:171-180."""
    return int(rollupFleet([node], []).out.node_gpu_count[0])


def getNodeGpuType(node: Any) -> str:
    """k8s.ts:185-192."""
    return rollupFleet([node], []).node_type(0)


def isNodeReady(node: Any) -> bool:
    """k8s.ts:329-331."""
    return bool(rollupFleet([node], []).out.node_out_flags[0] & K.NODEOUT_READY)


def isGpuRequestingPod(pod: Any) -> bool:
    """k8s.ts:250-264."""
    if not pod or not isinstance(pod, dict):
        return False
    return bool(rollupFleet([], [pod]).out.pod_is_gpu[0])


def filterGpuRequestingPods(items: Sequence[Any]) -> list:
    """k8s.ts:266-268."""
    if not items:
        return []
    r = rollupFleet([], list(items))
    return [items[i] for i in r.gpu_pod_rows()]


def getPodGpuRequests(pod: Any) -> Dict[str, str]:
    """k8s.ts:289-301."""
    return rollupFleet([], [pod]).pod_requests(0)
