"""Synthetic fleets of the shapes BASELINE.json names, reproducible everywhere.

Every random draw is counter-based: x = splitmix64(seed ^ (stream << 56) ^ index).  numpy here, the
C++ CPU baseline and the CUDA generator kernel (csrc/series_gen.cu) evaluate the same integer hash
and — for the power series — the same sequence of individually rounded float32 operations (no
libm, no fused multiply-add), so a shard generated on any device is bit-identical to the same rows
generated on the host.  (SURVEY's sketch used sin(); it is replaced by a parabolic day wave because
libm sinf and CUDA sinf differ in the last bits.)

`synth_fleet` emits the columnar form directly (vectorised; 10^6..10^7 pods) and can also emit
the same fleet as K8s-shaped dict objects for the marshal / object-oracle cross checks.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from . import columns as K
from .k8s import DEFAULT_RESOURCE_NAMES, INTEL_DISCRETE_GPU_NODE_ROLE, INTEL_GPU_NODE_LABEL, INTEL_INTEGRATED_GPU_NODE_ROLE

DEFAULT_SEED = 20261106

_U64 = np.uint64
_M1, _M2, _GOLD = _U64(0xBF58476D1CE4E5B9), _U64(0x94D049BB133111EB), _U64(0x9E3779B97F4A7C15)

# stream ids (shifted into the top byte)
S_NODE, S_POD, S_SLOT, S_SERIES_PARAM, S_SAMPLE, S_NAN, S_CHIP = 0x01, 0x02, 0x03, 0x11, 0x12, 0x13, 0x21


def splitmix64(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = (z + _GOLD).astype(_U64)
        z = (z ^ (z >> _U64(30))) * _M1
        z = (z ^ (z >> _U64(27))) * _M2
        return z ^ (z >> _U64(31))


def draw(seed: int, stream: int, index: np.ndarray, salt: int = 0) -> np.ndarray:
    """uint64 hash of (seed, stream, index[, salt]); salt selects independent sub-draws."""
    base = _U64((seed ^ (stream << 56)) & 0xFFFFFFFFFFFFFFFF)
    x = splitmix64(np.asarray(index, dtype=_U64) ^ base)
    if salt:
        x = splitmix64(x ^ _U64(salt))
    return x


def _pct(r: np.ndarray, shift: int = 0) -> np.ndarray:
    """0..9999 from 14+ bits of r."""
    return ((r >> _U64(shift)) % _U64(10000)).astype(np.int64)


# ------------------------------------------------------------------------------------------ fleet


def synth_nodes(n_nodes: int, seed: int = DEFAULT_SEED, first: int = 0, non_gpu_share_bp: int = 0):
    """Node columns. Mix: 60 % discrete role, 25 % igpu role, 10 % NFD label only, 5 % capacity only;
    `non_gpu_share_bp` basis points of plain nodes on top (default none)."""
    idx = np.arange(first, first + n_nodes, dtype=_U64)
    r = draw(seed, S_NODE, idx)
    r2 = draw(seed, S_NODE, idx, salt=1)
    kind_p = _pct(r)
    plain = _pct(r2, 40) < non_gpu_share_bp
    discrete = (kind_p < 6000) & ~plain
    igpu = (kind_p >= 6000) & (kind_p < 8500) & ~plain
    nfd = (kind_p >= 8500) & (kind_p < 9500) & ~plain
    cap_only = (kind_p >= 9500) & ~plain
    pick4 = np.array([1, 2, 4, 8], np.int32)[((r >> _U64(16)) % _U64(4)).astype(np.int64)]
    pick2 = np.array([1, 2], np.int32)[((r >> _U64(20)) % _U64(2)).astype(np.int64)]
    cap_i915 = np.where(discrete, pick4, np.where(igpu, 1, np.where(nfd, pick2, np.where(cap_only, np.minimum(pick4, 4), 0)))).astype(np.int32)
    has_xe = discrete & (_pct(r, 24) < 1000)
    cap_xe = np.where(has_xe, np.array([1, 2], np.int32)[((r >> _U64(44)) % _U64(2)).astype(np.int64)], 0).astype(np.int32)
    ready = _pct(r2) < 9700
    flags = (
        discrete * K.NODE_DISCRETE_ROLE
        + igpu * K.NODE_IGPU_ROLE
        + nfd * K.NODE_NFD_LABEL
        + ((cap_i915 > 0) | (cap_xe > 0)) * K.NODE_HAS_GPU_CAPACITY
        + ready * K.NODE_READY
    ).astype(np.uint8)
    return {
        "node_flags": flags,
        "node_cap_i915": cap_i915,
        "node_cap_xe": cap_xe,
        "node_alloc_i915": cap_i915.copy(),  # allocatable = capacity
        "node_alloc_xe": cap_xe.copy(),
        "node_alloc_own": (cap_i915 + cap_xe).astype(np.int32),
    }


_KEY_CDF = np.array([7000, 8000, 9000, 9500, 10000])  # i915, xe, millicores, memory.max, tiles
_SLOTS = 4  # up to 3 main containers + 1 init container


def _slot_value(key: np.ndarray, r: np.ndarray) -> np.ndarray:
    v_dev = 1 + (r % _U64(8)).astype(np.int64)  # i915 / xe: 1..8 devices
    v_mc = 100 + (r % _U64(901)).astype(np.int64)  # millicores 100..1000
    v_mem = 256 * (1 + (r % _U64(64)).astype(np.int64))  # memory.max in MiB, 256..16384
    v_tile = 1 + (r % _U64(4)).astype(np.int64)
    return np.where(key <= 1, v_dev, np.where(key == 2, v_mc, np.where(key == 3, v_mem, v_tile)))


def synth_pod_tables(n_pods: int, n_nodes_total: int, seed: int = DEFAULT_SEED, first: int = 0):
    """Per-pod and per-container-slot draws for pods [first, first + n_pods) of the fleet."""
    pidx = np.arange(first, first + n_pods, dtype=_U64)
    r = draw(seed, S_POD, pidx)
    r2 = draw(seed, S_POD, pidx, salt=1)
    u = (r2 >> _U64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    node = np.minimum(n_nodes_total - 1, np.floor(n_nodes_total * (u * u))).astype(np.int64)  # Zipf-ish, s = 0.5
    unscheduled = _pct(r) < 200
    ph = _pct(r, 14)
    phase = np.where(ph < 8500, 0, np.where(ph < 9300, 1, np.where(ph < 9700, 2, np.where(ph < 9900, 3, 4)))).astype(np.uint8)
    n_main = 1 + ((r >> _U64(28)) % _U64(3)).astype(np.int64)
    has_init = ((r >> _U64(32)) % _U64(2)).astype(bool)

    sidx = (pidx[:, None] * _U64(_SLOTS) + np.arange(_SLOTS, dtype=_U64)[None, :]).reshape(-1)
    s = draw(seed, S_SLOT, sidx).reshape(n_pods, _SLOTS)
    s2 = draw(seed, S_SLOT, sidx, salt=1).reshape(n_pods, _SLOTS)
    s3 = draw(seed, S_SLOT, sidx, salt=2).reshape(n_pods, _SLOTS)
    slot = np.arange(_SLOTS)[None, :]
    exists = np.where(slot < 3, slot < n_main[:, None], has_init[:, None])
    keyed = exists & (_pct(s) < 7000)
    key1 = np.searchsorted(_KEY_CDF, _pct(s, 14), side="right").astype(np.int64)
    key2 = np.searchsorted(_KEY_CDF, _pct(s, 28), side="right").astype(np.int64)
    second = keyed & (_pct(s, 42) < 3000) & (key2 != key1)
    mode_p = _pct(s2)
    mode = np.where(mode_p < 8000, 0, np.where(mode_p < 8500, 1, 2))  # 0 both (lim == req), 1 limits only, 2 requests only
    val1 = _slot_value(key1, s2 >> _U64(16))
    val2 = _slot_value(key2, s3 >> _U64(16))
    return {
        "node": node,
        "unscheduled": unscheduled,
        "phase": phase,
        "n_main": n_main,
        "has_init": has_init,
        "exists": exists,
        "keyed": keyed,
        "second": second,
        "key1": key1,
        "key2": key2,
        "mode": mode,
        "val1": val1,
        "val2": val2,
    }


def pods_per_node(n_pods: int, n_nodes_total: int, seed: int = DEFAULT_SEED) -> np.ndarray:
    """How many scheduled pods of the fleet sit on each node (only the two draws that decide it: cheap at 10 M pods)."""
    pidx = np.arange(n_pods, dtype=_U64)
    u = (draw(seed, S_POD, pidx, salt=1) >> _U64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    node = np.minimum(n_nodes_total - 1, np.floor(n_nodes_total * (u * u))).astype(np.int64)
    sched = _pct(draw(seed, S_POD, pidx)) >= 200
    return np.bincount(node[sched], minlength=n_nodes_total)


def synth_fleet(
    n_nodes: int,
    n_pods: int,
    seed: int = DEFAULT_SEED,
    node_range: Optional[Tuple[int, int]] = None,
    options: int = K.OPT_POD_LIMITS,
    non_gpu_share_bp: int = 0,
    with_tables: bool = False,
    unscheduled_share: Optional[Tuple[int, int]] = None,
):
    """Columns of the fleet (n_nodes, n_pods).  With `node_range=(lo, hi)` only nodes [lo, hi) and
    the pods scheduled onto them are emitted (node ids rebased to lo) — the multi-GPU shard of
    SURVEY §8(e).  Pods without a node belong to no range: they are dropped from ranged output unless
    `unscheduled_share=(rank, world)` asks for this rank's round-robin share of them (fleet-wide pod
    index % world == rank), which makes the result exactly `shard.shard_fleet(whole fleet, rank, world)`
    without building the whole fleet's entry stream."""
    lo, hi = node_range if node_range else (0, n_nodes)
    nodes = synth_nodes(hi - lo, seed, first=lo, non_gpu_share_bp=non_gpu_share_bp)
    t = synth_pod_tables(n_pods, n_nodes, seed)
    pod_node = np.where(t["unscheduled"], np.int64(K.NODE_NONE), t["node"])
    if node_range:
        keep = (~t["unscheduled"]) & (t["node"] >= lo) & (t["node"] < hi)
        if unscheduled_share is not None:
            rank, world = unscheduled_share
            keep |= t["unscheduled"] & (np.arange(n_pods) % world == rank)
        t = {k: v[keep] for k, v in t.items()}
        pod_node = np.where(t["unscheduled"], np.int64(K.NODE_NONE), t["node"] - lo).astype(np.int64)
    p = len(pod_node)
    # dense candidate table [P, slot, 4]: req k1, req k2, lim k1, lim k2
    keyed, second, mode = t["keyed"], t["second"], t["mode"]
    has_req, has_lim = keyed & (mode != 1), keyed & (mode != 2)
    valid = np.stack([has_req, has_req & second, has_lim, has_lim & second], axis=-1)
    res = np.stack([t["key1"], t["key2"], t["key1"], t["key2"]], axis=-1)
    val = np.stack([t["val1"], t["val2"], t["val1"], t["val2"]], axis=-1)
    is_lim = np.broadcast_to(np.array([0, 0, 1, 1])[None, None, :], valid.shape)
    is_init = np.broadcast_to((np.arange(_SLOTS) == 3)[None, :, None], valid.shape)
    # slots 0..2 are spec.containers[0..2], slot 3 is initContainers[0]; a key in both maps carries the same string
    container = np.broadcast_to(np.where(np.arange(_SLOTS) == 3, 0, np.arange(_SLOTS))[None, :, None], valid.shape)
    is_same = np.broadcast_to((keyed & (mode == 0))[:, :, None], valid.shape)
    per_pod = valid.reshape(p, _SLOTS * 4).sum(axis=1)
    off = np.zeros(p + 1, np.uint32)
    np.cumsum(per_pod, out=off[1:])
    flat = valid.reshape(-1)
    entries = np.zeros(int(off[-1]), K.ENTRY_DTYPE)
    entries["pod"] = np.repeat(np.arange(p, dtype=np.uint32), per_pod)
    entries["res"] = res.reshape(-1)[flat]
    entries["flags"] = (is_lim.reshape(-1)[flat] * K.ENTRY_LIMIT + is_init.reshape(-1)[flat] * K.ENTRY_INIT + K.ENTRY_TRUTHY
                        + is_same.reshape(-1)[flat] * K.ENTRY_SAME).astype(np.uint8)
    entries["container"] = container.reshape(-1)[flat]
    entries["value"] = val.reshape(-1)[flat]
    cols = K.FleetColumns(
        pod_node=pod_node.astype(np.uint32),
        pod_phase=t["phase"],
        pod_entry_off=off,
        entries=entries,
        n_resources=len(DEFAULT_RESOURCE_NAMES),
        options=options,
        resource_names=list(DEFAULT_RESOURCE_NAMES),
        **nodes,
    )
    return (cols, t) if with_tables else cols


def fleet_objects(n_nodes: int, n_pods: int, seed: int = DEFAULT_SEED, non_gpu_share_bp: int = 0):
    """The same fleet as `synth_fleet(n_nodes, n_pods, seed)` but as K8s-shaped dict objects
    (Python loops: small fleets only).  marshalFleet(*fleet_objects(...)) == synth_fleet(...)."""
    cols, t = synth_fleet(n_nodes, n_pods, seed, non_gpu_share_bp=non_gpu_share_bp, with_tables=True)
    names = DEFAULT_RESOURCE_NAMES
    nodes = []
    for i in range(n_nodes):
        f = int(cols.node_flags[i])
        labels = {"kubernetes.io/hostname": f"node-{i:06d}"}
        if f & K.NODE_DISCRETE_ROLE:
            labels[INTEL_DISCRETE_GPU_NODE_ROLE] = "true"
        if f & K.NODE_IGPU_ROLE:
            labels[INTEL_INTEGRATED_GPU_NODE_ROLE] = "true"
        if f & K.NODE_NFD_LABEL:
            labels[INTEL_GPU_NODE_LABEL] = "true"
        cap = {"cpu": "64", "memory": "263856Mi"}
        if cols.node_cap_i915[i]:
            cap[names[0]] = str(int(cols.node_cap_i915[i]))
        if cols.node_cap_xe[i]:
            cap[names[1]] = str(int(cols.node_cap_xe[i]))
        nodes.append(
            {
                "apiVersion": "v1",
                "kind": "Node",
                "metadata": {"name": f"node-{i:06d}", "labels": labels},
                "status": {
                    "capacity": cap,
                    "allocatable": dict(cap),
                    "conditions": [
                        {"type": "MemoryPressure", "status": "False"},
                        {"type": "Ready", "status": "True" if f & K.NODE_READY else "False"},
                    ],
                },
            }
        )
    pods = []
    for p in range(n_pods):
        mains, inits = [], []
        for sl in range(_SLOTS):
            if not t["exists"][p, sl]:
                continue
            c = {"name": f"c{sl}", "image": "registry.example/app:1"}
            req, lim = {"cpu": "500m"}, {}
            if t["keyed"][p, sl]:
                pairs = [(names[int(t["key1"][p, sl])], str(int(t["val1"][p, sl])))]
                if t["second"][p, sl]:
                    pairs.append((names[int(t["key2"][p, sl])], str(int(t["val2"][p, sl]))))
                for k, v in pairs:
                    if t["mode"][p, sl] != 1:
                        req[k] = v
                    if t["mode"][p, sl] != 2:
                        lim[k] = v
            c["resources"] = {"requests": req, **({"limits": lim} if lim else {})}
            (inits if sl == 3 else mains).append(c)
        spec = {"containers": mains}
        if inits:
            spec["initContainers"] = inits
        if not t["unscheduled"][p]:
            spec["nodeName"] = f"node-{int(t['node'][p]):06d}"
        ph = int(t["phase"][p])
        status = {"phase": K.PHASE_NAMES[ph]} if ph < 4 else {"phase": "Unknown"}
        pods.append(
            {"apiVersion": "v1", "kind": "Pod", "metadata": {"name": f"pod-{p:07d}", "namespace": "default"}, "spec": spec, "status": status}
        )
    return nodes, pods


# ----------------------------------------------------------------------------------- power series


def series_params(seed: int, global_series: np.ndarray):
    r = draw(seed, S_SERIES_PARAM, global_series)
    base = (np.float32(40.0) + (r & _U64(0xFFFF)).astype(np.float32) * np.float32(180.0 / 65536.0)).astype(np.float32)
    k = (((r >> _U64(16)) & _U64(0xFF)) % _U64(3)).astype(np.int64)
    tdp = np.array([150.0, 250.0, 300.0], np.float32)[k]
    phase = ((r >> _U64(24)) % _U64(86400)).astype(np.uint32)
    return base, tdp, phase


def synth_series(n_series: int, n_time: int, seed: int = DEFAULT_SEED, global_first: int = 0, nan_per_million: int = 0) -> np.ndarray:
    """[n_series, n_time] float32 power samples, bit-identical to csrc/series_gen.cu and
    oracle_generate_series:  p = clip(base + 0.25·base·wave(t) + 0.05·base·ξ, 0, tdp)."""
    f32 = np.float32
    gs = np.arange(global_first, global_first + n_series, dtype=_U64)
    base, tdp, phase = series_params(seed, gs)
    out = np.empty((n_series, n_time), f32)
    t = np.arange(n_time, dtype=np.uint32)
    a = (f32(0.25) * base).astype(f32)
    b = (f32(0.05) * base).astype(f32)
    for s in range(n_series):
        idx = gs[s] * _U64(n_time) + t.astype(_U64)
        r = draw(seed, S_SAMPLE, idx)
        xi = ((r >> _U64(40)).astype(np.uint32).astype(f32) * f32(1.0 / 8388608.0) - f32(1.0)).astype(f32)
        ph = (t + phase[s]) % np.uint32(86400)
        x = (ph.astype(f32) / f32(86400.0)).astype(f32)
        u = (f32(1.0) - x).astype(f32)
        y = (x * u).astype(f32)
        wave = (f32(8.0) * y - f32(1.0)).astype(f32)
        v = (base[s] + (a[s] * wave).astype(f32)).astype(f32)
        v = (v + (b[s] * xi).astype(f32)).astype(f32)
        v = np.where(v < f32(0.0), f32(0.0), v)
        v = np.where(v > tdp[s], tdp[s], v).astype(f32)
        if nan_per_million:
            rn = draw(seed, S_NAN, idx)
            v = np.where((rn % _U64(1000000)).astype(np.uint32) < np.uint32(nan_per_million), f32(np.nan), v)
        out[s] = v
    return out


def even_node_offsets(n_nodes: int, chips_per_node: int) -> np.ndarray:
    return (np.arange(n_nodes + 1, dtype=np.uint64) * chips_per_node).astype(np.uint32)


# ------------------------------------------------------------------------- instant power vectors


def synth_power_results(n_nodes: int, chips_per_node: int, seed: int = DEFAULT_SEED, missing_bp: int = 300):
    """Prometheus instant-vector results (dict objects, metrics.ts:44-47 shape) for the four queries
    of fetchGpuMetrics.  Chip label strings are fleet-unique; `missing_bp` basis
    points of chips lack an energy-rate or a TDP sample."""
    chips, energy, pmax, uname = [], [], [], []
    for n in range(n_nodes):
        inst = f"10.{(n >> 16) & 255}.{(n >> 8) & 255}.{n & 255}:9100"
        uname.append({"metric": {"__name__": "node_uname_info", "instance": inst, "nodename": f"node-{n:06d}"}, "value": [1793990000.0, "1"]})
        for c in range(chips_per_node):
            g = n * chips_per_node + c
            r = int(draw(seed, S_CHIP, np.array([g], dtype=_U64))[0])
            chip = f"0000:{(g >> 8) & 255:02x}:{g & 255:02x}_0_n{n}"
            chips.append({"metric": {"chip": chip, "chip_name": "i915", "instance": inst}, "value": [1793990000.0, "1"]})
            watts = 40 + (r & 0xFFFF) * 180.0 / 65536.0
            tdp = (150, 250, 300)[(r >> 16) % 3]
            if (r >> 24) % 10000 >= missing_bp:
                energy.append({"metric": {"chip": chip, "chip_name": "i915", "instance": inst}, "value": [1793990000.0, repr(watts)]})
            if (r >> 40) % 10000 >= missing_bp:
                pmax.append({"metric": {"chip": chip, "chip_name": "i915", "instance": inst}, "value": [1793990000.0, str(tdp)]})
    return chips, energy, pmax, uname
