"""ctypes loader for oracle/liborollup_oracle.so (rollup_oracle.cpp) — TEST INFRASTRUCTURE ONLY.

The struct layouts are the product's own (include/igpu_rollup.h), so the ctypes Structure classes
are borrowed from the package; the arithmetic is entirely the oracle's.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200.native import (
    FleetColsC,
    FleetOutC,
    PowerColsC,
    PowerOutC,
    RateParamsC,
    fleet_cols_struct,
    fleet_out_struct,
    power_cols_struct,
)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "liborollup_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "rollup_oracle.cpp")
    hdr = os.path.join(_HERE, "..", "include", "igpu_rollup.h")
    stale = (not os.path.exists(LIB)) or any(os.path.getmtime(f) > os.path.getmtime(LIB) for f in (src, hdr))
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "liborollup_oracle.so"], check=True, capture_output=True)
    return LIB


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = C.CDLL(LIB)
        vp, i32, u32, u64 = C.c_void_p, C.c_int32, C.c_uint32, C.c_uint64
        L.oracle_alloc_rollup.restype = i32
        L.oracle_alloc_rollup.argtypes = [C.POINTER(FleetColsC), C.POINTER(FleetOutC), i32]
        L.oracle_container_rows.restype = i32
        L.oracle_container_rows.argtypes = [C.POINTER(FleetColsC), vp, u64, C.POINTER(u64)]
        L.oracle_power_join_sum.restype = i32
        L.oracle_power_join_sum.argtypes = [C.POINTER(PowerColsC), C.POINTER(PowerOutC)]
        L.oracle_window_stats.restype = i32
        L.oracle_window_stats.argtypes = [vp, u32, u32, u32, vp, vp, vp, i32]
        L.oracle_quantile_by_node.restype = i32
        L.oracle_quantile_by_node.argtypes = [vp, u32, u32, u32, vp, C.c_double, i32, vp, vp, vp, i32]
        L.oracle_quantile_fleet.restype = i32
        L.oracle_quantile_fleet.argtypes = [vp, u64, C.c_double, i32, C.POINTER(C.c_double), C.POINTER(u64)]
        L.oracle_window_rollup.restype = i32
        L.oracle_window_rollup.argtypes = [vp, u32, u32, u32, vp, u32] + [vp] * 8
        L.oracle_generate_series.restype = i32
        L.oracle_generate_series.argtypes = [vp, u32, u32, u64, u64, u32, i32]
        L.oracle_counters_rate_len.restype = u32
        L.oracle_counters_rate_len.argtypes = [u32, C.POINTER(RateParamsC)]
        L.oracle_counters_rate.restype = i32
        L.oracle_counters_rate.argtypes = [vp, u32, u32, C.POINTER(RateParamsC), vp, vp, i32]
        L.oracle_counters_generate.restype = i32
        L.oracle_counters_generate.argtypes = [vp, u32, u32, u64, u64, u32, u32, u32, i32]
        L.oracle_num_threads.restype = i32
        _lib = L
    return _lib


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def alloc_rollup(cols: K.FleetColumns, threads: int = 1) -> K.FleetOutputs:
    out = K.FleetOutputs.allocate(cols.n_nodes, cols.n_pods, cols.n_resources, cols.options)
    cs, os_ = fleet_cols_struct(cols), fleet_out_struct(out)
    st = lib().oracle_alloc_rollup(C.byref(cs), C.byref(os_), threads)
    assert st == 0, st
    return out


def container_rows(cols: K.FleetColumns) -> np.ndarray:
    """One igpu_container_row per main container that owns a gpu.intel.com/ key (PodsPage.tsx:49-62)."""
    cs = fleet_cols_struct(cols)
    n = C.c_uint64(0)
    rows = np.zeros(max(1, cols.n_entries), K.CONTAINER_ROW_DTYPE)
    st = lib().oracle_container_rows(C.byref(cs), rows.ctypes.data, len(rows), C.byref(n))
    assert st == 0, st
    return rows[: n.value].copy()


def power_join_sum(p: K.PowerColumns) -> dict:
    n = len(p.chip_chip)
    out = {
        "chip_flags": np.zeros(n, np.uint8),
        "chip_watts": np.zeros(n, np.float64),
        "chip_max_watts": np.zeros(n, np.float64),
        "chip_node": np.zeros(n, np.uint32),
        "fleet": np.zeros(4, np.float64),
    }
    cs = power_cols_struct(p)
    os_ = PowerOutC(*[out[k].ctypes.data for k in ("chip_flags", "chip_watts", "chip_max_watts", "chip_node", "fleet")])
    st = lib().oracle_power_join_sum(C.byref(cs), C.byref(os_))
    assert st == 0, st
    return out


def window_stats(x: np.ndarray, window: int, threads: int = 1, out=None):
    x = np.ascontiguousarray(x, np.float32)
    s, t = x.shape
    nw = (t + window - 1) // window
    mn, me, mx = out if out is not None else (np.empty((s, nw), np.float32) for _ in range(3))
    assert all(a.dtype == np.float32 and a.shape == (s, nw) and a.flags.c_contiguous for a in (mn, me, mx))
    st = lib().oracle_window_stats(x.ctypes.data, s, t, window, mn.ctypes.data, me.ctypes.data, mx.ctypes.data, threads)
    assert st == 0, st
    return mn, me, mx


def quantile_by_node(x: np.ndarray, node_series_off, q: float = 0.95, mode: int = 0, threads: int = 1):
    x = np.ascontiguousarray(x, np.float32)
    off = np.ascontiguousarray(node_series_off, np.uint32)
    n = len(off) - 1
    qv = np.empty(n, np.float64)
    stats = np.zeros(n, K.NODE_STATS_DTYPE)
    fleet = np.zeros(K.FLEET_LEN, np.float64)
    st = lib().oracle_quantile_by_node(
        x.ctypes.data, x.shape[0], x.shape[1], n, off.ctypes.data, q, mode, qv.ctypes.data, stats.ctypes.data, fleet.ctypes.data, threads
    )
    assert st == 0, st
    return qv, stats, fleet


def quantile_fleet(x: np.ndarray, q: float = 0.95, mode: int = 0):
    """Quantile over every valid sample of the slab -> (value, number of valid samples)."""
    x = np.ascontiguousarray(x, np.float32)
    out, cnt = C.c_double(0), C.c_uint64(0)
    st = lib().oracle_quantile_fleet(x.ctypes.data, x.size, q, mode, C.byref(out), C.byref(cnt))
    assert st == 0, st
    return out.value, cnt.value


def window_rollup(x: np.ndarray, node_series_off, window: int):
    """Window min / mean / max / count per node and per fleet -> (node dict of [N][nw] arrays, fleet dict of [nw] arrays)."""
    x = np.ascontiguousarray(x, np.float32)
    off = np.ascontiguousarray(node_series_off, np.uint32)
    n, nw = len(off) - 1, (x.shape[1] + window - 1) // window
    node = {"min": np.empty((n, nw), np.float32), "mean": np.empty((n, nw), np.float32), "max": np.empty((n, nw), np.float32),
            "count": np.empty((n, nw), np.uint32)}
    fleet = {"min": np.empty(nw, np.float32), "mean": np.empty(nw, np.float32), "max": np.empty(nw, np.float32), "count": np.empty(nw, np.uint64)}
    st = lib().oracle_window_rollup(x.ctypes.data, x.shape[0], x.shape[1], n, off.ctypes.data, window,
                                    *[node[k].ctypes.data for k in ("min", "mean", "max", "count")],
                                    *[fleet[k].ctypes.data for k in ("min", "mean", "max", "count")])
    assert st == 0, st
    return node, fleet


def generate_series(n_series: int, n_time: int, seed: int, global_first: int = 0, nan_per_million: int = 0, threads: int = 1):
    out = np.empty((n_series, n_time), np.float32)
    st = lib().oracle_generate_series(out.ctypes.data, n_series, n_time, seed, global_first, nan_per_million, threads)
    assert st == 0, st
    return out


def counters_rate(v: np.ndarray, prm: RateParamsC, threads: int = 1):
    """Prometheus extrapolatedRate per evaluation point -> (float32 [S][n_out], float64 [S][n_out])."""
    v = np.ascontiguousarray(v, dtype=np.float64)
    n_out = int(lib().oracle_counters_rate_len(v.shape[1], C.byref(prm)))
    o32 = np.empty((v.shape[0], n_out), np.float32)
    o64 = np.empty((v.shape[0], n_out), np.float64)
    st = lib().oracle_counters_rate(v.ctypes.data, v.shape[0], v.shape[1], C.byref(prm), o32.ctypes.data, o64.ctypes.data, threads)
    assert st == 0, st
    return o32, o64


def counters_generate(n_series: int, n_time: int, seed: int, global_first: int = 0, step_ms: int = 15000, reset_period: int = 0,
                      missing_per_million: int = 0, threads: int = 1):
    out = np.empty((n_series, n_time), np.float64)
    st = lib().oracle_counters_generate(out.ctypes.data, n_series, n_time, seed, global_first, step_ms, reset_period, missing_per_million, threads)
    assert st == 0, st
    return out
