"""ctypes binding of libigpu_rollup.so (include/igpu_rollup.h) — the verified drop-in boundary.

There is deliberately no fallback: if the shared library is missing, or no CUDA device can be
opened, every entry point raises NativeUnavailable / IgpuError.  Nothing in this package computes a
rollup on the CPU.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence, Tuple

import numpy as np

from . import columns as K

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libigpu_rollup.so")

EXPORTS = (
    "igpu_version igpu_strerror igpu_ctx_create igpu_ctx_destroy igpu_last_error igpu_ctx_sync igpu_ctx_reload_env igpu_get_timing "
    "igpu_host_alloc igpu_host_free igpu_comm_unique_id igpu_comm_init igpu_comm_destroy igpu_alloc_rollup "
    "igpu_fleet_create igpu_fleet_upload igpu_fleet_rollup igpu_fleet_download igpu_fleet_destroy igpu_fleet_container_rows "
    "igpu_marshal_fleet_json igpu_marshal_columns igpu_marshal_pod_order igpu_marshal_resource_name igpu_marshal_free "
    "igpu_power_join_sum igpu_power_join_sum_static igpu_series_create igpu_series_destroy igpu_series_upload igpu_series_download "
    "igpu_series_generate igpu_series_window_stats igpu_series_read_window_stats igpu_series_quantile_by_node igpu_series_rollup "
    "igpu_series_quantile_fleet igpu_series_fleet_quantile_reads igpu_series_window_rollup igpu_series_keep_window_counts "
    "igpu_counters_create igpu_counters_destroy igpu_counters_upload igpu_counters_generate igpu_counters_gap_stat igpu_counters_rate_len igpu_counters_rate "
    "igpu_marshal_matrix_json igpu_matrix_n_series igpu_matrix_values igpu_matrix_label igpu_matrix_dropped igpu_matrix_free"
).split()


class NativeUnavailable(RuntimeError):
    """libigpu_rollup.so is not built / not loadable.  There is no CPU fallback."""


class IgpuError(RuntimeError):
    def __init__(self, status: int, text: str):
        super().__init__(f"igpu status {status}: {text}")
        self.status = status


class FleetColsC(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_uint32),
        ("n_pods", C.c_uint32),
        ("n_entries", C.c_uint64),
        ("n_resources", C.c_uint32),
        ("options", C.c_uint32),
        ("node_flags", C.c_void_p),
        ("node_cap_i915", C.c_void_p),
        ("node_cap_xe", C.c_void_p),
        ("node_alloc_i915", C.c_void_p),
        ("node_alloc_xe", C.c_void_p),
        ("node_alloc_own", C.c_void_p),
        ("pod_node", C.c_void_p),
        ("pod_phase", C.c_void_p),
        ("pod_entry_off", C.c_void_p),
        ("entries", C.c_void_p),
    ]


_OUT_FIELDS = K.FleetOutputs.FIELDS


class FleetOutC(C.Structure):
    _fields_ = [(f, C.c_void_p) for f in _OUT_FIELDS]


class PowerColsC(C.Structure):
    _fields_ = [
        ("n_chips", C.c_uint32),
        ("n_energy", C.c_uint32),
        ("n_power_max", C.c_uint32),
        ("n_uname", C.c_uint32),
        ("n_chip_ids", C.c_uint32),
        ("n_instance_ids", C.c_uint32),
        ("chip_chip", C.c_void_p),
        ("chip_instance", C.c_void_p),
        ("energy_chip", C.c_void_p),
        ("energy_value", C.c_void_p),
        ("power_max_chip", C.c_void_p),
        ("power_max_value", C.c_void_p),
        ("uname_instance", C.c_void_p),
        ("uname_node", C.c_void_p),
    ]


class PowerOutC(C.Structure):
    _fields_ = [
        ("chip_flags", C.c_void_p),
        ("chip_watts", C.c_void_p),
        ("chip_max_watts", C.c_void_p),
        ("chip_node", C.c_void_p),
        ("fleet", C.c_void_p),
    ]


class WindowRollupOutC(C.Structure):
    """igpu_window_rollup_out (include/igpu_rollup.h)."""

    _fields_ = [(f, C.c_void_p) for f in "node_min node_mean node_max node_count fleet_min fleet_mean fleet_max fleet_count slab_reads".split()]


class TimingC(C.Structure):
    _fields_ = [
        ("h2d_ms", C.c_float),
        ("kernel_ms", C.c_float),
        ("d2h_ms", C.c_float),
        ("total_ms", C.c_float),
        ("h2d_bytes", C.c_uint64),
        ("d2h_bytes", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("kernel_detail_ms", C.c_float * 12),
    ]


KERNEL_NAMES = ("pod_rollup", "node_rollup", "window_stats", "quantile_sample", "quantile_scan", "quantile_finish",
                "quantile_tail", "power_join", "series_fused", "counter_rate")


class RateParamsC(C.Structure):
    """igpu_rate_params (include/igpu_rollup.h)."""

    _fields_ = [("step_ms", C.c_uint32), ("range_samples", C.c_uint32), ("first_eval", C.c_uint32),
                ("eval_stride", C.c_uint32), ("options", C.c_uint32)]


RATE_CLOSED_LEFT, RATE_INCREASE, RATE_GAUGE = 1, 2, 4


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data


def fleet_cols_struct(cols: K.FleetColumns) -> FleetColsC:
    s = FleetColsC()
    s.n_nodes, s.n_pods, s.n_entries = cols.n_nodes, cols.n_pods, cols.n_entries
    s.n_resources, s.options = cols.n_resources, cols.options
    for name in (
        "node_flags node_cap_i915 node_cap_xe node_alloc_i915 node_alloc_xe node_alloc_own "
        "pod_node pod_phase pod_entry_off entries"
    ).split():
        setattr(s, name, _ptr(getattr(cols, name)))
    return s


def fleet_out_struct(out: K.FleetOutputs) -> FleetOutC:
    s = FleetOutC()
    for f in _OUT_FIELDS:
        setattr(s, f, _ptr(getattr(out, f)))
    return s


def power_cols_struct(p: K.PowerColumns) -> PowerColsC:
    s = PowerColsC()
    s.n_chips, s.n_energy = len(p.chip_chip), len(p.energy_chip)
    s.n_power_max, s.n_uname = len(p.power_max_chip), len(p.uname_instance)
    s.n_chip_ids, s.n_instance_ids = p.n_chip_ids, p.n_instance_ids
    for name in (
        "chip_chip chip_instance energy_chip energy_value power_max_chip power_max_value uname_instance uname_node"
    ).split():
        setattr(s, name, _ptr(getattr(p, name)))
    return s


STATIC_JOIN_MAX_IDS = 1 << 18

_lib = None


def load_library(path: Optional[str] = None) -> C.CDLL:
    """dlopen the product library and declare prototypes.  Raises NativeUnavailable if absent."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("IGPU_ROLLUP_LIB", LIB_PATH)
    if not os.path.exists(p):
        raise NativeUnavailable(
            f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the rollup path)"
        )
    try:
        lib = C.CDLL(p, mode=C.RTLD_GLOBAL)
    except OSError as e:  # e.g. libcudart missing
        raise NativeUnavailable(f"cannot load {p}: {e}") from e
    vp, i32, u32, u64 = C.c_void_p, C.c_int32, C.c_uint32, C.c_uint64
    sig = {
        "igpu_version": (i32, []),
        "igpu_strerror": (C.c_char_p, [i32]),
        "igpu_ctx_create": (i32, [i32, vp, C.POINTER(vp)]),
        "igpu_ctx_destroy": (i32, [vp]),
        "igpu_last_error": (C.c_char_p, [vp]),
        "igpu_ctx_sync": (i32, [vp]),
        "igpu_ctx_reload_env": (i32, [vp]),
        "igpu_get_timing": (i32, [vp, C.POINTER(TimingC)]),
        "igpu_host_alloc": (i32, [u64, C.POINTER(vp)]),
        "igpu_host_free": (i32, [vp]),
        "igpu_comm_unique_id": (i32, [vp]),
        "igpu_comm_init": (i32, [vp, vp, i32, i32]),
        "igpu_comm_destroy": (i32, [vp]),
        "igpu_alloc_rollup": (i32, [vp, C.POINTER(FleetColsC), C.POINTER(FleetOutC)]),
        "igpu_fleet_create": (i32, [vp, u32, u32, u64, u32, u32, C.POINTER(vp)]),
        "igpu_fleet_upload": (i32, [vp, vp, C.POINTER(FleetColsC)]),
        "igpu_fleet_rollup": (i32, [vp, vp]),
        "igpu_fleet_download": (i32, [vp, vp, C.POINTER(FleetOutC)]),
        "igpu_fleet_container_rows": (i32, [vp, vp, vp, u64, C.POINTER(u64)]),
        "igpu_fleet_destroy": (i32, [vp, vp]),
        "igpu_marshal_fleet_json": (i32, [C.c_char_p, u64, C.c_char_p, u64, u32, i32, C.POINTER(vp)]),
        "igpu_marshal_columns": (i32, [vp, C.POINTER(FleetColsC)]),
        "igpu_marshal_pod_order": (vp, [vp]),
        "igpu_marshal_resource_name": (C.c_char_p, [vp, u32]),
        "igpu_marshal_free": (i32, [vp]),
        "igpu_power_join_sum": (i32, [vp, C.POINTER(PowerColsC), C.POINTER(PowerOutC)]),
        "igpu_power_join_sum_static": (i32, [i32, vp, C.POINTER(PowerColsC), C.POINTER(PowerOutC)]),
        "igpu_series_create": (i32, [vp, u32, u32, u32, vp, C.POINTER(vp)]),
        "igpu_series_destroy": (i32, [vp, vp]),
        "igpu_series_upload": (i32, [vp, vp, u32, u32, vp]),
        "igpu_series_download": (i32, [vp, vp, u32, u32, vp]),
        "igpu_series_generate": (i32, [vp, vp, u64, u64, u32]),
        "igpu_series_window_stats": (i32, [vp, vp, u32, vp, vp, vp]),
        "igpu_series_read_window_stats": (i32, [vp, vp, u32, u32, vp, vp, vp]),
        "igpu_series_quantile_by_node": (i32, [vp, vp, C.c_double, i32, vp, vp, vp]),
        "igpu_series_rollup": (i32, [vp, vp, u32, C.c_double, i32, vp, vp, vp, vp, vp, vp]),
        "igpu_series_quantile_fleet": (i32, [vp, vp, C.c_double, i32, C.POINTER(C.c_double), C.POINTER(u64)]),
        "igpu_series_window_rollup": (i32, [vp, vp, u32, C.POINTER(WindowRollupOutC)]),
        "igpu_series_keep_window_counts": (i32, [vp, vp, i32]),
        "igpu_series_fleet_quantile_reads": (u32, [vp]),
        "igpu_counters_create": (i32, [vp, u32, u32, C.POINTER(vp)]),
        "igpu_counters_destroy": (i32, [vp, vp]),
        "igpu_counters_upload": (i32, [vp, vp, u32, u32, vp]),
        "igpu_counters_generate": (i32, [vp, vp, u64, u64, u32, u32, u32]),
        "igpu_counters_gap_stat": (i32, [vp, vp, C.POINTER(u64), C.POINTER(u64)]),
        "igpu_counters_rate_len": (u32, [u32, C.POINTER(RateParamsC)]),
        "igpu_counters_rate": (i32, [vp, vp, C.POINTER(RateParamsC), vp, vp]),
        "igpu_marshal_matrix_json": (i32, [C.c_char_p, u64, C.c_int64, u32, u32, C.POINTER(C.c_char_p), u32, i32, C.POINTER(vp)]),
        "igpu_matrix_n_series": (u32, [vp]),
        "igpu_matrix_values": (vp, [vp]),
        "igpu_matrix_label": (C.c_char_p, [vp, u32, u32]),
        "igpu_matrix_dropped": (u64, [vp]),
        "igpu_matrix_free": (i32, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError here = header and library disagree
        fn.restype, fn.argtypes = res, args
    if path is None:
        _lib = lib
    return lib


def marshal_fleet_json(nodes_json: bytes, pods_json: bytes, options: int = K.OPT_POD_LIMITS, threads: int = 0) -> K.FleetColumns:
    """igpu_marshal_fleet_json: K8s JSON text -> FleetColumns on the host cores (no GPU needed).
    The columns are copied out of the native object, which is freed before returning.
    With K.OPT_GROUP_PODS_BY_NODE in `options` the pods come out grouped by node and `cols.pod_order` holds the
    permutation (position in the JSON of the pod in each column position)."""
    lib = load_library()
    h = C.c_void_p()
    st = lib.igpu_marshal_fleet_json(nodes_json, len(nodes_json), pods_json, len(pods_json), options, threads, C.byref(h))
    if st != 0:
        raise IgpuError(st, (lib.igpu_last_error(None) or b"").decode() or lib.igpu_strerror(st).decode())
    try:
        cs = FleetColsC()
        lib.igpu_marshal_columns(h, C.byref(cs))

        def arr(ptr, dtype, n):
            if not n:
                return np.zeros(0, dtype)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n * np.dtype(dtype).itemsize,)).view(dtype).copy()

        names, i = [], 0
        while True:
            nm = lib.igpu_marshal_resource_name(h, i)
            if nm is None:
                break
            names.append(nm.decode("utf-8", errors="replace"))  # invalid UTF-8 in a key: U+FFFD, as a text decoder would
            i += 1
        n, p, e = cs.n_nodes, cs.n_pods, cs.n_entries
        return K.FleetColumns(
            node_flags=arr(cs.node_flags, np.uint8, n), node_cap_i915=arr(cs.node_cap_i915, np.int32, n),
            node_cap_xe=arr(cs.node_cap_xe, np.int32, n), node_alloc_i915=arr(cs.node_alloc_i915, np.int32, n),
            node_alloc_xe=arr(cs.node_alloc_xe, np.int32, n), node_alloc_own=arr(cs.node_alloc_own, np.int32, n),
            pod_node=arr(cs.pod_node, np.uint32, p), pod_phase=arr(cs.pod_phase, np.uint8, p),
            pod_entry_off=arr(cs.pod_entry_off, np.uint32, p + 1), entries=arr(cs.entries, K.ENTRY_DTYPE, e),
            n_resources=cs.n_resources, options=cs.options, resource_names=names,
            pod_order=arr(lib.igpu_marshal_pod_order(h), np.uint32, p).astype(np.int64) if lib.igpu_marshal_pod_order(h) else None)
    finally:
        lib.igpu_marshal_free(h)


def marshal_matrix_json(text: bytes, start_ms: int, step_ms: int, n_time: int, labels: Sequence[str] = ("chip", "instance"),
                        threads: int = 0):
    """Prometheus query_range response text -> (float64 [series][n_time] with NaN gaps, per-series label tuples, dropped count)."""
    lib = load_library()
    names = (C.c_char_p * max(len(labels), 1))(*[l.encode() for l in labels])
    h = C.c_void_p()
    st = lib.igpu_marshal_matrix_json(text, len(text), start_ms, step_ms, n_time, names, len(labels), threads, C.byref(h))
    if st != 0:
        raise IgpuError(st, (lib.igpu_last_error(None) or b"").decode(errors="replace"))
    try:
        n = int(lib.igpu_matrix_n_series(h))
        src = (C.c_double * (n * n_time)).from_address(lib.igpu_matrix_values(h)) if n else None
        values = np.frombuffer(src, np.float64).reshape(n, n_time).copy() if n else np.empty((0, n_time))
        labs = []
        for s_ in range(n):
            row = []
            for l in range(len(labels)):
                v = lib.igpu_matrix_label(h, s_, l)
                row.append(None if v is None else v.decode("utf-8", errors="replace"))
            labs.append(tuple(row))
        return values, labs, int(lib.igpu_matrix_dropped(h))
    finally:
        lib.igpu_matrix_free(h)


class PinnedBuffer:
    """Page-locked host memory from igpu_host_alloc, viewable as numpy arrays."""

    def __init__(self, nbytes: int):
        self._lib = load_library()
        p = C.c_void_p()
        st = self._lib.igpu_host_alloc(max(1, nbytes), C.byref(p))
        if st != 0:
            raise IgpuError(st, self._lib.igpu_strerror(st).decode())
        self.ptr, self.nbytes = p.value, nbytes
        self._raw = (C.c_uint8 * max(1, nbytes)).from_address(self.ptr)

    def view(self, dtype, shape, offset: int = 0) -> np.ndarray:
        dt = np.dtype(dtype)
        n = int(np.prod(shape)) if not isinstance(shape, int) else shape
        a = np.frombuffer(self._raw, dtype=dt, count=n, offset=offset)
        return a.reshape(shape)

    def free(self):
        if self.ptr:
            self._raw = None
            self._lib.igpu_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def pinned_copy_of_columns(cols: K.FleetColumns):
    """Copies every column into one pinned slab (so igpu_fleet_upload runs at full PCIe rate)."""
    names = (
        "node_flags node_cap_i915 node_cap_xe node_alloc_i915 node_alloc_xe node_alloc_own "
        "pod_node pod_phase pod_entry_off entries"
    ).split()
    offs, total = {}, 0
    for n in names:
        total = (total + 255) & ~255
        offs[n] = total
        total += getattr(cols, n).nbytes
    buf = PinnedBuffer(total)
    kw = {}
    for n in names:
        a = getattr(cols, n)
        v = buf.view(a.dtype, a.shape, offs[n])
        v[...] = a
        kw[n] = v
    out = K.FleetColumns(n_resources=cols.n_resources, options=cols.options, resource_names=cols.resource_names,
                         node_names=cols.node_names, pod_names=cols.pod_names, **kw)
    out._pinned = buf  # keep alive
    return out


def pinned_outputs(n_nodes: int, n_pods: int, n_resources: int, options: int) -> K.FleetOutputs:
    proto = K.FleetOutputs.allocate(0, 0, n_resources, options)
    shapes = K.FleetOutputs.allocate(1, 1, n_resources, options)
    del proto
    spec, total = {}, 0
    for f in K.FleetOutputs.FIELDS:
        a = getattr(shapes, f)
        if a is None:
            spec[f] = None
            continue
        if f.startswith("pod_"):
            shape = (n_resources, n_pods) if a.ndim == 2 else (n_pods,)
        elif f.startswith("node_"):
            shape = (n_resources, n_nodes) if a.ndim == 2 else (n_nodes,)
        else:
            shape = (K.TOTALS_LEN,)
        total = (total + 255) & ~255
        spec[f] = (a.dtype, shape, total)
        total += int(np.prod(shape)) * a.dtype.itemsize
    buf = PinnedBuffer(total)
    kw = {f: (None if s is None else buf.view(*s)) for f, s in spec.items()}
    out = K.FleetOutputs(**kw)
    out._pinned = buf
    return out


class Context:
    """One CUDA device + stream (igpu_ctx).  Not thread-safe; one call at a time."""

    def __init__(self, device: int = 0, stream: Optional[int] = None):
        self.lib = load_library()
        h = C.c_void_p()
        st = self.lib.igpu_ctx_create(device, C.c_void_p(stream) if stream else None, C.byref(h))
        if st != 0:
            raise IgpuError(st, self.lib.igpu_strerror(st).decode())
        self.h = h
        self.device = device

    # -- plumbing
    def _ck(self, st: int):
        if st != 0:
            raise IgpuError(st, (self.lib.igpu_last_error(self.h) or b"").decode() or self.lib.igpu_strerror(st).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.igpu_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        self._ck(self.lib.igpu_ctx_sync(self.h))

    def reload_env(self):
        """The IGPU_* switches are read when the context is created; this reads them again (tests, sweeps)."""
        self._ck(self.lib.igpu_ctx_reload_env(self.h))

    def timing(self) -> dict:
        t = TimingC()
        self._ck(self.lib.igpu_get_timing(self.h, C.byref(t)))
        d = {f: getattr(t, f) for f, _ in TimingC._fields_ if f != "kernel_detail_ms"}
        d["kernels"] = {n: float(t.kernel_detail_ms[i]) for i, n in enumerate(KERNEL_NAMES)}
        return d

    # -- communicator
    def comm_unique_id(self) -> bytes:
        b = C.create_string_buffer(128)
        st = self.lib.igpu_comm_unique_id(b)
        if st != 0:
            raise IgpuError(st, self.lib.igpu_strerror(st).decode())
        return b.raw

    def comm_init(self, uid: bytes, rank: int, nranks: int):
        b = C.create_string_buffer(uid, 128)
        self._ck(self.lib.igpu_comm_init(self.h, b, rank, nranks))

    # -- K1
    def alloc_rollup(self, cols: K.FleetColumns, out: Optional[K.FleetOutputs] = None) -> K.FleetOutputs:
        """igpu_alloc_rollup: host columns in, host columns out (H2D + kernels + D2H + sync)."""
        if out is None:
            out = K.FleetOutputs.allocate(cols.n_nodes, cols.n_pods, cols.n_resources, cols.options)
        cs, os_ = fleet_cols_struct(cols), fleet_out_struct(out)
        self._ck(self.lib.igpu_alloc_rollup(self.h, C.byref(cs), C.byref(os_)))
        return out

    def fleet_create(self, max_nodes, max_pods, max_entries, n_resources, options=K.OPT_POD_LIMITS) -> "Fleet":
        h = C.c_void_p()
        self._ck(self.lib.igpu_fleet_create(self.h, max_nodes, max_pods, max_entries, n_resources, options, C.byref(h)))
        return Fleet(self, h)

    # -- K2
    def power_join_sum(self, p: K.PowerColumns) -> dict:
        n = len(p.chip_chip)
        out = {
            "chip_flags": np.zeros(n, np.uint8),
            "chip_watts": np.zeros(n, np.float64),
            "chip_max_watts": np.zeros(n, np.float64),
            "chip_node": np.zeros(n, np.uint32),
            "fleet": np.zeros(4, np.float64),
        }
        cs = power_cols_struct(p)
        os_ = PowerOutC(*[_ptr(out[k]) for k in ("chip_flags", "chip_watts", "chip_max_watts", "chip_node", "fleet")])
        self._ck(self.lib.igpu_power_join_sum(self.h, C.byref(cs), C.byref(os_)))
        return out

    # -- K3 / K4
    def series_create(self, n_series: int, n_time: int, node_series_off: Sequence[int]) -> "Series":
        off = np.ascontiguousarray(node_series_off, dtype=np.uint32)
        h = C.c_void_p()
        self._ck(self.lib.igpu_series_create(self.h, n_series, n_time, len(off) - 1, off.ctypes.data, C.byref(h)))
        return Series(self, h, n_series, n_time, len(off) - 1)

    # -- N3
    def counters_create(self, n_series: int, n_time: int) -> "Counters":
        h = C.c_void_p()
        self._ck(self.lib.igpu_counters_create(self.h, n_series, n_time, C.byref(h)))
        return Counters(self, h, n_series, n_time)


class Fleet:
    """Device-resident fleet columns (igpu_fleet)."""

    def __init__(self, ctx: Context, h):
        self.ctx, self.h = ctx, h

    def upload(self, cols: K.FleetColumns):
        cs = fleet_cols_struct(cols)
        self.ctx._ck(self.ctx.lib.igpu_fleet_upload(self.ctx.h, self.h, C.byref(cs)))
        self._shape = (cols.n_nodes, cols.n_pods, cols.n_resources, cols.options)

    def rollup(self):
        self.ctx._ck(self.ctx.lib.igpu_fleet_rollup(self.ctx.h, self.h))

    def download(self, out: Optional[K.FleetOutputs] = None) -> K.FleetOutputs:
        if out is None:
            out = K.FleetOutputs.allocate(*self._shape)
        os_ = fleet_out_struct(out)
        self.ctx._ck(self.ctx.lib.igpu_fleet_download(self.ctx.h, self.h, C.byref(os_)))
        return out

    def container_rows(self) -> np.ndarray:
        """igpu_fleet_container_rows: one record per main container that names a gpu.intel.com/ key, in (pod, container)
        order (the filter and key Set of PodsPage.tsx:49-62 / PodDetailSection.tsx:48-65 for the whole fleet)."""
        n = C.c_uint64(0)
        self.ctx._ck(self.ctx.lib.igpu_fleet_container_rows(self.ctx.h, self.h, None, 0, C.byref(n)))
        rows = np.zeros(n.value, K.CONTAINER_ROW_DTYPE)
        if n.value:
            self.ctx._ck(self.ctx.lib.igpu_fleet_container_rows(self.ctx.h, self.h, rows.ctypes.data, n.value, C.byref(n)))
        return rows

    def destroy(self):
        if self.h:
            self.ctx.lib.igpu_fleet_destroy(self.ctx.h, self.h)
            self.h = None


class Series:
    """Device-resident [series][time] float32 slab (igpu_series)."""

    def __init__(self, ctx: Context, h, n_series, n_time, n_nodes):
        self.ctx, self.h = ctx, h
        self.n_series, self.n_time, self.n_nodes = n_series, n_time, n_nodes

    def upload(self, rows: np.ndarray, first: int = 0):
        rows = np.ascontiguousarray(rows, dtype=np.float32)
        assert rows.ndim == 2 and rows.shape[1] == self.n_time
        self.ctx._ck(self.ctx.lib.igpu_series_upload(self.ctx.h, self.h, first, rows.shape[0], rows.ctypes.data))
        self.ctx.sync()  # the source array may be freed by the caller right after

    def upload_async(self, rows: np.ndarray, first: int = 0):
        assert rows.dtype == np.float32 and rows.flags.c_contiguous
        self.ctx._ck(self.ctx.lib.igpu_series_upload(self.ctx.h, self.h, first, rows.shape[0], rows.ctypes.data))

    def download(self, first: int, count: int) -> np.ndarray:
        out = np.empty((count, self.n_time), np.float32)
        self.ctx._ck(self.ctx.lib.igpu_series_download(self.ctx.h, self.h, first, count, out.ctypes.data))
        return out

    def generate(self, seed: int, global_first: int = 0, nan_per_million: int = 0):
        self.ctx._ck(self.ctx.lib.igpu_series_generate(self.ctx.h, self.h, seed, global_first, nan_per_million))

    def n_windows(self, window: int) -> int:
        return (self.n_time + window - 1) // window

    def window_stats(self, window: int, fetch: bool = True):
        """min/mean/max per window; fetch=False leaves the result on the device."""
        if not fetch:
            self.ctx._ck(self.ctx.lib.igpu_series_window_stats(self.ctx.h, self.h, window, None, None, None))
            return None
        nw = self.n_windows(window)
        mn, me, mx = (np.empty((self.n_series, nw), np.float32) for _ in range(3))
        self.ctx._ck(
            self.ctx.lib.igpu_series_window_stats(self.ctx.h, self.h, window, mn.ctypes.data, me.ctypes.data, mx.ctypes.data)
        )
        return mn, me, mx

    def read_window_stats(self, window: int, first: int, count: int):
        nw = self.n_windows(window)
        mn, me, mx = (np.empty((count, nw), np.float32) for _ in range(3))
        self.ctx._ck(
            self.ctx.lib.igpu_series_read_window_stats(
                self.ctx.h, self.h, first, count, mn.ctypes.data, me.ctypes.data, mx.ctypes.data
            )
        )
        return mn, me, mx

    def quantile_by_node(self, q: float = 0.95, mode: int = 0):
        qv = np.empty(self.n_nodes, np.float64)
        st = np.zeros(self.n_nodes, K.NODE_STATS_DTYPE)
        fl = np.zeros(K.FLEET_LEN, np.float64)
        self.ctx._ck(
            self.ctx.lib.igpu_series_quantile_by_node(
                self.ctx.h, self.h, q, mode, qv.ctypes.data, st.ctypes.data, fl.ctypes.data
            )
        )
        return qv, st, fl

    def rollup(self, window: int = 60, q: float = 0.95, mode: int = 0, fetch_windows: bool = True):
        """Window min/mean/max AND per-node quantile in one read (igpu_series_rollup).
        Returns ((min, mean, max) or None, quantile[n_nodes], stats, fleet)."""
        nw = self.n_windows(window)
        w = tuple(np.empty((self.n_series, nw), np.float32) for _ in range(3)) if fetch_windows else None
        qv = np.empty(self.n_nodes, np.float64)
        st = np.zeros(self.n_nodes, K.NODE_STATS_DTYPE)
        fl = np.zeros(K.FLEET_LEN, np.float64)
        wp = [a.ctypes.data for a in w] if w else [None, None, None]
        self.ctx._ck(
            self.ctx.lib.igpu_series_rollup(
                self.ctx.h, self.h, window, q, mode, wp[0], wp[1], wp[2], qv.ctypes.data, st.ctypes.data, fl.ctypes.data
            )
        )
        return w, qv, st, fl

    def quantile_fleet(self, q: float = 0.95, mode: int = 0):
        """igpu_series_quantile_fleet: the quantile over every valid sample of the fleet (of all ranks, when the context
        has a communicator) -> (value, number of valid samples)."""
        out, cnt = C.c_double(0), C.c_uint64(0)
        self.ctx._ck(self.ctx.lib.igpu_series_quantile_fleet(self.ctx.h, self.h, q, mode, C.byref(out), C.byref(cnt)))
        return out.value, cnt.value

    def fleet_quantile_reads(self) -> int:
        """Full passes over the slab the last quantile_fleet() made: 1 (bracketed by a sample), 3 (exact passes), 4 (bracket missed)."""
        return int(self.ctx.lib.igpu_series_fleet_quantile_reads(self.h))

    def window_rollup(self, window: int = 60, nodes: bool = True):
        """igpu_series_window_rollup: window min / mean / max / count per node (this rank's) and per fleet (all ranks').
        Returns (node dict of [n_nodes][n_win] arrays or None, fleet dict of [n_win] arrays)."""
        nw = self.n_windows(window)
        node = None
        if nodes:
            node = {"min": np.empty((self.n_nodes, nw), np.float32), "mean": np.empty((self.n_nodes, nw), np.float32),
                    "max": np.empty((self.n_nodes, nw), np.float32), "count": np.empty((self.n_nodes, nw), np.uint32)}
        fleet = {"min": np.empty(nw, np.float32), "mean": np.empty(nw, np.float32), "max": np.empty(nw, np.float32),
                 "count": np.empty(nw, np.uint64)}
        o = WindowRollupOutC()
        for k in ("min", "mean", "max", "count"):
            setattr(o, "node_" + k, node[k].ctypes.data if node is not None and node[k].size else None)
            setattr(o, "fleet_" + k, fleet[k].ctypes.data if nw else None)
        reads = C.c_uint32(1)
        o.slab_reads = C.addressof(reads)
        self.ctx._ck(self.ctx.lib.igpu_series_window_rollup(self.ctx.h, self.h, window, C.byref(o)))
        self.last_window_rollup_slab_reads = int(reads.value)  # 0: folded from the planes a rollup of this window kept
        return node, fleet

    def keep_window_counts(self, on: bool = True):
        """igpu_series_keep_window_counts: rollup() also keeps valid-sample counts per (series, window), so that a
        window_rollup() of the same window needs no second read of the slab."""
        self.ctx._ck(self.ctx.lib.igpu_series_keep_window_counts(self.ctx.h, self.h, 1 if on else 0))

    def rollup_into(self, window: int, q: float, mode: int, wmin, wmean, wmax, qv: np.ndarray, st: np.ndarray, fl: np.ndarray):
        """igpu_series_rollup writing into caller-owned host arrays (pinned ones make the copies asynchronous DMA);
        wmin / wmean / wmax may be None to leave the window planes on the device."""
        for a_ in (wmin, wmean, wmax):
            assert a_ is None or (a_.dtype == np.float32 and a_.flags.c_contiguous and a_.size == self.n_series * self.n_windows(window))
        assert qv.dtype == np.float64 and qv.size == self.n_nodes and st.dtype == K.NODE_STATS_DTYPE and st.size == self.n_nodes
        assert fl.dtype == np.float64 and fl.size == K.FLEET_LEN
        self.ctx._ck(
            self.ctx.lib.igpu_series_rollup(self.ctx.h, self.h, window, q, mode, _ptr(wmin), _ptr(wmean), _ptr(wmax), qv.ctypes.data,
                                            st.ctypes.data, fl.ctypes.data)
        )

    def destroy(self):
        if self.h:
            self.ctx.lib.igpu_series_destroy(self.ctx.h, self.h)
            self.h = None


class Counters:
    """Device-resident [series][time] float64 slab of cumulative counters (igpu_counters)."""

    def __init__(self, ctx: Context, h, n_series, n_time):
        self.ctx, self.h = ctx, h
        self.n_series, self.n_time = n_series, n_time

    def upload(self, rows: np.ndarray, first: int = 0):
        rows = np.ascontiguousarray(rows, dtype=np.float64)
        assert rows.ndim == 2 and rows.shape[1] == self.n_time
        self.ctx._ck(self.ctx.lib.igpu_counters_upload(self.ctx.h, self.h, first, rows.shape[0], rows.ctypes.data))
        self.ctx.sync()

    def generate(self, seed: int, global_first: int = 0, step_ms: int = 15000, reset_period: int = 0, missing_per_million: int = 0):
        self.ctx._ck(self.ctx.lib.igpu_counters_generate(self.ctx.h, self.h, seed, global_first, step_ms, reset_period, missing_per_million))

    def gap_stat(self) -> Tuple[int, int]:
        """(missing, looked at) of the sparse sample upload / generate took of the slab; rate() picks its kernel by it."""
        miss, seen = C.c_uint64(0), C.c_uint64(0)
        self.ctx._ck(self.ctx.lib.igpu_counters_gap_stat(self.ctx.h, self.h, C.byref(miss), C.byref(seen)))
        return int(miss.value), int(seen.value)

    @staticmethod
    def params(step_ms: int, range_samples: int, first_eval: Optional[int] = None, eval_stride: int = 1, options: int = 0) -> RateParamsC:
        if first_eval is None:  # first evaluation whose range lies wholly inside the slab
            first_eval = range_samples if options & RATE_CLOSED_LEFT else range_samples - 1
        return RateParamsC(step_ms, range_samples, first_eval, eval_stride, options)

    def rate_len(self, prm: RateParamsC) -> int:
        return int(self.ctx.lib.igpu_counters_rate_len(self.n_time, C.byref(prm)))

    def rate(self, prm: RateParamsC, dst: "Series", want_f64: bool = False) -> Optional[np.ndarray]:
        """rate()/increase()/delta() of every counter into `dst` (device to device); optionally the float64 results."""
        r64 = np.empty((self.n_series, self.rate_len(prm)), np.float64) if want_f64 else None
        self.ctx._ck(self.ctx.lib.igpu_counters_rate(self.ctx.h, self.h, C.byref(prm), dst.h, _ptr(r64)))
        return r64

    def destroy(self):
        if self.h:
            self.ctx.lib.igpu_counters_destroy(self.ctx.h, self.h)
            self.h = None


_default: Optional[Context] = None


def default_context() -> Context:
    """Lazily created process-wide context on device $IGPU_DEVICE (default 0)."""
    global _default
    if _default is None:
        _default = Context(int(os.environ.get("IGPU_DEVICE", os.environ.get("LOCAL_RANK", "0"))))
    return _default
