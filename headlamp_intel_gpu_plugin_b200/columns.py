"""Columnar layout shared by the host marshal layer and the CUDA kernels.

Mirrors include/igpu_rollup.h one to one (same names, same bit values); the reference-side meaning
of every column is documented there with file:line citations.  Nothing in here computes: these are
containers, constants and consistency checks.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

# node_flags bits (k8s.ts:136-138,146-148,329-331; OverviewPage.tsx:98-99)
NODE_DISCRETE_ROLE = 0x01
NODE_IGPU_ROLE = 0x02
NODE_NFD_LABEL = 0x04
NODE_HAS_GPU_CAPACITY = 0x08
NODE_READY = 0x10
NODE_CAP_NAN = 0x20
NODE_ALLOC_NAN = 0x40

# pod_phase codes (OverviewPage.tsx:122-130)
PHASE_RUNNING, PHASE_PENDING, PHASE_SUCCEEDED, PHASE_FAILED, PHASE_OTHER, PHASE_PROTO = range(6)
PHASE_NAMES = ("Running", "Pending", "Succeeded", "Failed", "Other")

NODE_NONE = 0xFFFFFFFF

RES_I915, RES_XE, RES_MILLICORES, RES_MEMORY_MAX = 0, 1, 2, 3
MAX_RESOURCES = 32

ENTRY_LIMIT = 0x01
ENTRY_INIT = 0x02
ENTRY_TRUTHY = 0x04
ENTRY_SAME = 0x08  # the container holds the same key in its other map with a strictly equal raw value (req === lim)
MAX_CONTAINERS = 255

OPT_POD_LIMITS = 0x1
OPT_GROUP_PODS_BY_NODE = 0x100  # marshal-only: igpu_marshal_fleet_json lays pods out grouped by node

ENTRY_DTYPE = np.dtype(
    [("pod", "<u4"), ("res", "<u2"), ("flags", "u1"), ("container", "u1"), ("value", "<i8")], align=True
)
assert ENTRY_DTYPE.itemsize == 16

# fleet_totals slots
T_GPU_NODES, T_DISCRETE, T_INTEGRATED, T_UNKNOWN, T_GPU_DEVICES, T_READY_NODES = 0, 1, 2, 3, 4, 5
T_CAPACITY, T_ALLOCATABLE, T_CAP_NAN, T_ALLOC_NAN, T_ALLOCATED, T_GPU_PODS = 6, 7, 8, 9, 10, 11
T_PHASE0, T_SCHEDULED_PODS, T_UTIL_PCT, T_REQ0, T_LIM0, TOTALS_LEN = 12, 17, 18, 32, 64, 96
# slots that are NOT additive across shards (recomputed after the reduction)
T_BAD_INPUT = 19
T_NON_ADDITIVE = (T_UTIL_PCT,)

NODEOUT_IS_GPU, NODEOUT_READY, NODEOUT_CAP_NAN, NODEOUT_ALLOC_NAN, NODEOUT_BAR_DASH = 0x01, 0x02, 0x04, 0x08, 0x10
TYPE_DISCRETE, TYPE_INTEGRATED, TYPE_UNKNOWN = 0, 1, 2
TYPE_NAMES = ("discrete", "integrated", "unknown")

POWER_HAS_WATTS, POWER_HAS_MAX, POWER_HAS_NODE = 0x1, 0x2, 0x4
F_SUM, F_COUNT, F_MIN, F_MAX, F_NODES, FLEET_LEN = 0, 1, 2, 3, 4, 8

# igpu_container_row: one per main container that names a gpu.intel.com/ key (PodsPage.tsx:49-62)
CONTAINER_ROW_DTYPE = np.dtype([("pod", "<u4"), ("container", "<u4"), ("req_keys", "<u4"), ("lim_keys", "<u4"), ("req_truthy", "<u4"),
                                ("lim_truthy", "<u4"), ("same", "<u4"), ("n_entries", "<u4")])
assert CONTAINER_ROW_DTYPE.itemsize == 32

NODE_STATS_DTYPE = np.dtype([("sum", "<f8"), ("count", "<u8"), ("min", "<f4"), ("max", "<f4")], align=True)
assert NODE_STATS_DTYPE.itemsize == 24


def _c(a, dtype, n=None):
    a = np.ascontiguousarray(a, dtype=dtype)
    if n is not None and a.shape != (n,):
        raise ValueError(f"column has shape {a.shape}, expected ({n},)")
    return a


@dataclass
class FleetColumns:
    """Input columns of one fleet or one node-range shard (igpu_fleet_cols)."""

    node_flags: np.ndarray
    node_cap_i915: np.ndarray
    node_cap_xe: np.ndarray
    node_alloc_i915: np.ndarray
    node_alloc_xe: np.ndarray
    node_alloc_own: np.ndarray
    pod_node: np.ndarray
    pod_phase: np.ndarray
    pod_entry_off: np.ndarray
    entries: np.ndarray
    n_resources: int = 5
    options: int = OPT_POD_LIMITS
    # host-side dictionaries (never cross the C ABI)
    resource_names: Optional[List[str]] = None
    node_names: Optional[List[str]] = None
    pod_names: Optional[List[str]] = None
    # set when the pod columns are a regrouping of the caller's pods: pod_order[i] = the caller's index of the pod in
    # column position i (grouped_by_node, or the native marshal with OPT_GROUP_PODS_BY_NODE)
    pod_order: Optional[np.ndarray] = None

    def __post_init__(self):
        n = len(self.node_flags)
        self.node_flags = _c(self.node_flags, np.uint8, n)
        for k in ("node_cap_i915", "node_cap_xe", "node_alloc_i915", "node_alloc_xe", "node_alloc_own"):
            setattr(self, k, _c(getattr(self, k), np.int32, n))
        p = len(self.pod_node)
        self.pod_node = _c(self.pod_node, np.uint32, p)
        self.pod_phase = _c(self.pod_phase, np.uint8, p)
        self.pod_entry_off = _c(self.pod_entry_off, np.uint32, p + 1)
        self.entries = np.ascontiguousarray(self.entries, dtype=ENTRY_DTYPE)

    @property
    def n_nodes(self) -> int:
        return len(self.node_flags)

    @property
    def n_pods(self) -> int:
        return len(self.pod_node)

    @property
    def n_entries(self) -> int:
        return len(self.entries)

    def validate(self) -> None:
        """Checks the invariants the kernels rely on; raises ValueError with the first violation."""
        if not (4 <= self.n_resources <= MAX_RESOURCES):
            raise ValueError(f"n_resources {self.n_resources} outside [4, {MAX_RESOURCES}]")
        off = self.pod_entry_off
        if off[0] != 0 or off[-1] != self.n_entries or np.any(np.diff(off.astype(np.int64)) < 0):
            raise ValueError("pod_entry_off must start at 0, end at n_entries and be non-decreasing")
        if self.n_entries:
            e = self.entries
            if int(e["res"].max()) >= self.n_resources:
                raise ValueError("entry resource id >= n_resources")
            owner = np.repeat(np.arange(self.n_pods, dtype=np.uint32), np.diff(off))
            if not np.array_equal(owner, e["pod"]):
                raise ValueError("entries[i].pod disagrees with pod_entry_off")
            if int(np.abs(e["value"]).astype(object).sum()) >= (1 << 53):
                raise ValueError("sum of |entry.value| reaches 2^53: outside the exact-integer domain of the reference")
            # inside a pod: main containers in order, then initContainers in order (container runs are contiguous)
            key = ((e["flags"] & ENTRY_INIT).astype(np.int64) << 7) | e["container"].astype(np.int64)
            same_pod = e["pod"][1:] == e["pod"][:-1]
            if np.any(same_pod & (key[1:] < key[:-1])):
                raise ValueError("entries of a pod must be grouped by container: spec.containers in order, then initContainers")
            if int(e["container"].max()) >= MAX_CONTAINERS:
                raise ValueError(f"container index >= {MAX_CONTAINERS}")
        sched = self.pod_node[self.pod_node != NODE_NONE]
        if sched.size and int(sched.max()) >= self.n_nodes:
            raise ValueError("pod_node id >= n_nodes")
        if self.pod_phase.size and int(self.pod_phase.max()) > PHASE_PROTO:
            raise ValueError("pod_phase code out of range")

    def nbytes_in(self) -> int:
        """Algorithmic input bytes of the allocation rollup: 16E + 9P + 4 + 21N."""
        return 16 * self.n_entries + 9 * self.n_pods + 4 + 21 * self.n_nodes

    def grouped_by_node(self):
        """The same fleet with the pods of each node adjacent (stable counting sort on pod_node, unscheduled pods
        last) -> (columns, order) where order[i] = index, in THIS object, of the pod now at position i.

        Results do not depend on pod order; grouped pods let the device sum a node's pods with one warp reduction
        and one set of atomics per run instead of one per pod (csrc/alloc_rollup.cu, tiled kernel).  Per-pod outputs
        of a rollup over the grouped columns are in grouped order: `outputs[field][..., inverse]` with
        `inverse = np.argsort(order)` restores this object's order."""
        order = np.argsort(self.pod_node, kind="stable").astype(np.int64)
        off = self.pod_entry_off.astype(np.int64)
        lens = (off[1:] - off[:-1])[order]
        new_off = np.zeros(self.n_pods + 1, np.int64)
        np.cumsum(lens, out=new_off[1:])
        # entry i of the new stream comes from old position start_of_its_pod + offset_inside_the_pod
        src = np.repeat(off[:-1][order] - new_off[:-1], lens) + np.arange(int(new_off[-1]), dtype=np.int64)
        entries = self.entries[src].copy()
        entries["pod"] = np.repeat(np.arange(self.n_pods, dtype=np.uint32), lens)
        cols = FleetColumns(
            node_flags=self.node_flags, node_cap_i915=self.node_cap_i915, node_cap_xe=self.node_cap_xe,
            node_alloc_i915=self.node_alloc_i915, node_alloc_xe=self.node_alloc_xe, node_alloc_own=self.node_alloc_own,
            pod_node=self.pod_node[order], pod_phase=self.pod_phase[order], pod_entry_off=new_off.astype(np.uint32), entries=entries,
            n_resources=self.n_resources, options=self.options, resource_names=self.resource_names, node_names=self.node_names,
            pod_names=[self.pod_names[i] for i in order] if self.pod_names else None,
            pod_order=order if self.pod_order is None else np.asarray(self.pod_order)[order],
        )
        return cols, order


@dataclass
class FleetOutputs:
    """Output columns (igpu_fleet_out) as numpy arrays."""

    pod_is_gpu: np.ndarray
    pod_req_mask: np.ndarray
    pod_req: np.ndarray  # [R, P]
    pod_lim_mask: Optional[np.ndarray]
    pod_lim: Optional[np.ndarray]
    node_out_flags: np.ndarray
    node_type: np.ndarray
    node_gpu_count: np.ndarray
    node_cap_total: np.ndarray
    node_alloc_total: np.ndarray
    node_pod_count: np.ndarray
    node_bar_pct: np.ndarray
    node_bar_level: np.ndarray
    node_inuse: np.ndarray
    node_util_pct: np.ndarray
    node_util_level: np.ndarray
    node_req: np.ndarray  # [R, N]
    node_lim: Optional[np.ndarray]
    fleet_totals: np.ndarray
    pod_gpu_containers: Optional[np.ndarray] = None  # [P] gpuContainers.length (PodDetailSection.tsx:48-52,106-108)

    @staticmethod
    def allocate(n_nodes: int, n_pods: int, n_resources: int, options: int = OPT_POD_LIMITS) -> "FleetOutputs":
        lim = bool(options & OPT_POD_LIMITS)
        z = np.zeros
        return FleetOutputs(
            pod_is_gpu=z(n_pods, np.uint8),
            pod_req_mask=z(n_pods, np.uint32),
            pod_req=z((n_resources, n_pods), np.int64),
            pod_lim_mask=z(n_pods, np.uint32) if lim else None,
            pod_lim=z((n_resources, n_pods), np.int64) if lim else None,
            node_out_flags=z(n_nodes, np.uint8),
            node_type=z(n_nodes, np.uint8),
            node_gpu_count=z(n_nodes, np.int64),
            node_cap_total=z(n_nodes, np.int64),
            node_alloc_total=z(n_nodes, np.int64),
            node_pod_count=z(n_nodes, np.uint32),
            node_bar_pct=z(n_nodes, np.int64),
            node_bar_level=z(n_nodes, np.uint8),
            node_inuse=z(n_nodes, np.int64),
            node_util_pct=z(n_nodes, np.int64),
            node_util_level=z(n_nodes, np.uint8),
            node_req=z((n_resources, n_nodes), np.int64),
            node_lim=z((n_resources, n_nodes), np.int64) if lim else None,
            fleet_totals=z(TOTALS_LEN, np.int64),
            pod_gpu_containers=z(n_pods, np.uint8),
        )

    FIELDS = (
        "pod_is_gpu pod_req_mask pod_req pod_lim_mask pod_lim node_out_flags node_type node_gpu_count "
        "node_cap_total node_alloc_total node_pod_count node_bar_pct node_bar_level node_inuse node_util_pct "
        "node_util_level node_req node_lim fleet_totals pod_gpu_containers"
    ).split()

    def nbytes_out(self) -> int:
        return sum(getattr(self, f).nbytes for f in self.FIELDS if getattr(self, f) is not None)


@dataclass
class PowerColumns:
    """Instant-vector join inputs (igpu_power_cols); ids index host-side string dictionaries."""

    chip_chip: np.ndarray
    chip_instance: np.ndarray
    energy_chip: np.ndarray
    energy_value: np.ndarray
    power_max_chip: np.ndarray
    power_max_value: np.ndarray
    uname_instance: np.ndarray
    uname_node: np.ndarray
    n_chip_ids: int
    n_instance_ids: int
    chip_names: List[str] = field(default_factory=list)
    instance_names: List[str] = field(default_factory=list)
    node_names: List[str] = field(default_factory=list)

    def __post_init__(self):
        for k in ("chip_chip", "chip_instance", "energy_chip", "power_max_chip", "uname_instance", "uname_node"):
            setattr(self, k, np.ascontiguousarray(getattr(self, k), dtype=np.uint32))
        for k in ("energy_value", "power_max_value"):
            setattr(self, k, np.ascontiguousarray(getattr(self, k), dtype=np.float64))
