"""The N-API addon (bindings/node/igpu_addon.c) EXECUTED without Node.js: its C code runs against a stand-in
for the Node-API calls it uses (tests/napi_shim/napi_shim.c, a toy value model driven through ctypes).  What this proves: the
addon's marshalling of typed arrays into igpu_fleet_cols / igpu_power_cols, its argument validation, its output typed arrays
and the module-load behaviour (no CPU fallback) — against the ctypes path on the same inputs.  What it does not prove: anything
about V8 or a real Node.js, where the addon remains unverified."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import k8s, metrics, native, synth

from .dirty import dirty_fleet

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "napi_shim", "_build")
SO = os.path.join(BUILD, "igpu_addon_shim.so")
# napi_typedarray_type
U8, I32, U32, F32, F64, BI64 = 1, 5, 6, 7, 8, 9
NP_OF = {0: np.int8, 1: np.uint8, 2: np.uint8, 3: np.int16, 4: np.uint16, 5: np.int32, 6: np.uint32, 7: np.float32, 8: np.float64,
         9: np.int64, 10: np.uint64}
V_STRING, V_OBJECT, V_TYPEDARRAY = 2, 3, 5


def build_shim() -> str:
    srcs = [os.path.join(ROOT, "tests", "napi_shim", "napi_shim.c"), os.path.join(ROOT, "bindings", "node", "igpu_addon.c")]
    deps = srcs + [os.path.join(ROOT, "bindings", "node", "node_api_min.h"), os.path.join(ROOT, "include", "igpu_rollup.h"), native.LIB_PATH]
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        os.makedirs(BUILD, exist_ok=True)
        csrc = os.path.dirname(native.LIB_PATH)
        subprocess.run(["gcc", "-shared", "-fPIC", "-O1", "-g", "-Wall", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "bindings", "node"),
                        *srcs, "-L" + csrc, "-ligpu_rollup", "-Wl,-rpath," + csrc, "-o", SO], check=True, capture_output=True)
    return SO


class Shim:
    def __init__(self):
        native.load_library()  # the addon links against the product library
        L = self.L = C.CDLL(build_shim())
        vp = C.c_void_p
        for name, res, args in [("shim_env_new", vp, []), ("shim_env_free", None, [vp]), ("shim_load", vp, [vp]), ("shim_exception", C.c_char_p, [vp]),
                                ("shim_exception_code", C.c_char_p, [vp]), ("shim_clear_exception", None, [vp]), ("shim_object", vp, [vp]),
                                ("shim_number", vp, [vp, C.c_double]), ("shim_typedarray", vp, [vp, C.c_int, vp, C.c_size_t]),
                                ("shim_set", C.c_int, [vp, vp, C.c_char_p, vp]), ("shim_get", vp, [vp, vp, C.c_char_p]), ("shim_kind", C.c_int, [vp]),
                                ("shim_string", C.c_char_p, [vp]), ("shim_typedarray_info", C.c_int, [vp, C.POINTER(C.c_int), C.POINTER(C.c_size_t), C.POINTER(vp)]),
                                ("shim_key", C.c_char_p, [vp, C.c_size_t]), ("shim_call", vp, [vp, C.c_char_p, C.POINTER(vp), C.c_size_t])]:
            getattr(L, name).restype, getattr(L, name).argtypes = res, args
        self.env = L.shim_env_new()
        self.keep = []

    def load(self):
        return self.L.shim_load(self.env)

    def exception(self):
        e = self.L.shim_exception(self.env)
        return None if e is None else (self.L.shim_exception_code(self.env).decode(), e.decode())

    def obj(self, **props):
        o = self.L.shim_object(self.env)
        for k, v in props.items():
            self.L.shim_set(self.env, o, k.encode(), v)
        return o

    def ta(self, kind, a):
        a = np.ascontiguousarray(a)
        self.keep.append(a)
        return self.L.shim_typedarray(self.env, kind, a.ctypes.data if a.size else None, a.size)

    def num(self, x):
        return self.L.shim_number(self.env, float(x))

    def call(self, name, *args):
        argv = (C.c_void_p * max(1, len(args)))(*args)
        self.L.shim_clear_exception(self.env)
        return self.L.shim_call(self.env, name.encode(), argv, len(args))

    def to_python(self, v):
        kind = self.L.shim_kind(v)
        if kind == V_TYPEDARRAY:
            t, n, d = C.c_int(), C.c_size_t(), C.c_void_p()
            assert self.L.shim_typedarray_info(v, C.byref(t), C.byref(n), C.byref(d)) == 0
            dt = np.dtype(NP_OF[t.value])
            return np.frombuffer((C.c_char * (n.value * dt.itemsize)).from_address(d.value), dt).copy() if n.value else np.zeros(0, dt)
        if kind == V_STRING:
            return self.L.shim_string(v).decode()
        if kind == V_OBJECT:
            out, i = {}, 0
            while (k := self.L.shim_key(v, i)) is not None:
                out[k.decode()] = self.to_python(self.L.shim_get(self.env, v, k))
                i += 1
            return out
        raise AssertionError(f"unexpected value kind {kind}")

    def close(self):
        self.L.shim_env_free(self.env)


def fleet_arg(sh: Shim, cols: K.FleetColumns, **override):
    props = dict(nodeFlags=sh.ta(U8, cols.node_flags), capI915=sh.ta(I32, cols.node_cap_i915), capXe=sh.ta(I32, cols.node_cap_xe),
                 allocI915=sh.ta(I32, cols.node_alloc_i915), allocXe=sh.ta(I32, cols.node_alloc_xe), allocOwn=sh.ta(I32, cols.node_alloc_own),
                 podNode=sh.ta(U32, cols.pod_node), podPhase=sh.ta(U8, cols.pod_phase), podEntryOff=sh.ta(U32, cols.pod_entry_off),
                 entries=sh.ta(U8, cols.entries.view(np.uint8)), nResources=sh.num(cols.n_resources), options=sh.num(cols.options))
    props.update(override)
    return sh.obj(**{k: v for k, v in props.items() if v is not None})


ADDON_FIELD = {"pod_is_gpu": "podIsGpu", "pod_gpu_containers": "podGpuContainers", "pod_req_mask": "podReqMask", "pod_req": "podReq",
               "pod_lim_mask": "podLimMask", "pod_lim": "podLim", "node_out_flags": "nodeOutFlags", "node_type": "nodeType",
               "node_gpu_count": "nodeGpuCount", "node_cap_total": "nodeCapTotal", "node_alloc_total": "nodeAllocTotal",
               "node_pod_count": "nodePodCount", "node_bar_pct": "nodeBarPct", "node_bar_level": "nodeBarLevel", "node_inuse": "nodeInUse",
               "node_util_pct": "nodeUtilPct", "node_util_level": "nodeUtilLevel", "node_req": "nodeReq", "node_lim": "nodeLim",
               "fleet_totals": "fleetTotals"}


def assert_same_as_ctypes(got: dict, want: K.FleetOutputs):
    assert set(ADDON_FIELD.values()) <= set(got), sorted(set(ADDON_FIELD.values()) - set(got))
    for f, name in ADDON_FIELD.items():
        assert np.array_equal(got[name], getattr(want, f).reshape(-1)), f


def test_shim_builds_and_module_load_has_no_cpu_fallback():
    """Without a CUDA device `require()` of the addon throws (the TypeScript layer then stays on the reference path)."""
    import torch

    sh = Shim()
    exports = sh.load()
    if torch.cuda.is_available():
        assert exports and sh.exception() is None
    else:
        assert not exports
        code, msg = sh.exception()
        assert "CUDA" in msg or "device" in msg, msg
    sh.close()


@pytest.mark.gpu
def test_rollup_fleet_through_the_addon_equals_the_ctypes_path(ctx):
    sh = Shim()
    assert sh.load(), sh.exception()
    for cols in (synth.synth_fleet(50, 2000), k8s.marshalFleet(*dirty_fleet(40, 400, 8)), k8s.marshalFleet([], [])):
        out = sh.call("rollupFleet", fleet_arg(sh, cols))
        assert out, sh.exception()
        assert_same_as_ctypes(sh.to_python(out), ctx.alloc_rollup(cols))
    sh.close()


@pytest.mark.gpu
def test_rollup_fleet_json_through_the_addon(ctx):
    nodes, pods = dirty_fleet(30, 500, 3)
    nj, pj = json.dumps(nodes).encode(), json.dumps(pods).encode()
    sh = Shim()
    assert sh.load(), sh.exception()
    out = sh.call("rollupFleetJson", sh.ta(U8, np.frombuffer(nj, np.uint8)), sh.ta(U8, np.frombuffer(pj, np.uint8)))
    assert out, sh.exception()
    got = sh.to_python(out)
    r = k8s.rollupFleetJson(nj, pj, ctx, group_by_node=False)
    assert_same_as_ctypes(got, r.out)
    assert [got["resourceNames"][str(i)] for i in range(len(got["resourceNames"]))] == r.cols.resource_names
    assert not sh.call("rollupFleetJson", sh.ta(U8, np.frombuffer(b"[", np.uint8)), sh.ta(U8, np.frombuffer(b"[]", np.uint8)))
    assert "JSON" in sh.exception()[1]  # the library's message travels as the exception text
    sh.close()


@pytest.mark.gpu
def test_power_join_through_the_addon(ctx):
    ch, en, mx, un = synth.synth_power_results(60, 4)
    p = metrics.marshalMetrics(ch, en, mx, un)
    sh = Shim()
    assert sh.load(), sh.exception()
    arg = sh.obj(chipChip=sh.ta(U32, p.chip_chip), chipInstance=sh.ta(U32, p.chip_instance), energyChip=sh.ta(U32, p.energy_chip),
                 energyValue=sh.ta(F64, p.energy_value), powerMaxChip=sh.ta(U32, p.power_max_chip), powerMaxValue=sh.ta(F64, p.power_max_value),
                 unameInstance=sh.ta(U32, p.uname_instance), unameNode=sh.ta(U32, p.uname_node), nChipIds=sh.num(p.n_chip_ids),
                 nInstanceIds=sh.num(p.n_instance_ids))
    out = sh.call("powerJoinSum", arg)
    assert out, sh.exception()
    got, want = sh.to_python(out), ctx.power_join_sum(p)
    for a, b in (("chipFlags", "chip_flags"), ("chipWatts", "chip_watts"), ("chipMaxWatts", "chip_max_watts"), ("chipNode", "chip_node"), ("fleet", "fleet")):
        assert np.array_equal(got[a], want[b], equal_nan=True), a
    short = sh.obj(chipChip=sh.ta(U32, p.chip_chip), chipInstance=sh.ta(U32, p.chip_instance[:-1]))
    assert not sh.call("powerJoinSum", short) and "chipInstance" in sh.exception()[1]
    sh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("fleet_views", [0, 1])
def test_range_view_through_the_addon(ctx, oracle_lib, fleet_views):
    """rangeView: upload, single-read rollup, and with fleetViews the fleet quantile and the curve — against the ctypes path
    on the same slab (bit for bit: same library, same kernels) and against the oracle."""
    nodes, chips, T, W = 6, 4, 2880, 60
    x = synth.synth_series(nodes * chips, T, seed=4, nan_per_million=3000)
    x[chips:2 * chips, :2 * W] = np.nan
    off = synth.even_node_offsets(nodes, chips)
    sh = Shim()
    assert sh.load(), sh.exception()
    arg = sh.obj(x=sh.ta(F32, x.reshape(-1)), nodeOff=sh.ta(U32, off), nTime=sh.num(T), window=sh.num(W), q=sh.num(0.95), fleetViews=sh.num(fleet_views))
    out = sh.call("rangeView", arg)
    assert out, sh.exception()
    got = sh.to_python(out)
    s = ctx.series_create(nodes * chips, T, off)
    s.upload(x)
    _, qv, st, fl = s.rollup(W, 0.95, 0, fetch_windows=False)
    assert np.array_equal(got["nodeQuantile"], qv) and np.array_equal(got["fleet"], fl)
    assert np.array_equal(got["nodeSum"], st["sum"]) and np.array_equal(got["nodeCount"], st["count"].astype(np.float64))
    assert np.array_equal(got["nodeMin"], st["min"]) and np.array_equal(got["nodeMax"], st["max"])
    oq, ost, ofl = oracle_lib.quantile_by_node(x, off, 0.95)
    assert np.allclose(got["nodeQuantile"], oq, rtol=1e-5) and np.array_equal(got["nodeCount"], ost["count"].astype(np.float64))
    if fleet_views:
        assert got["fleetQuantile"][0] == s.quantile_fleet(0.95)[0] == oracle_lib.quantile_fleet(x, 0.95)[0]
        _, ocurve = oracle_lib.window_rollup(x, off, W)
        assert np.array_equal(got["curveCount"], ocurve["count"].astype(np.float64))
        assert np.array_equal(got["curveMax"], ocurve["max"], equal_nan=True) and np.array_equal(got["curveMin"], ocurve["min"], equal_nan=True)
        assert np.allclose(got["curveMean"], ocurve["mean"], rtol=1e-5, equal_nan=True)
    else:
        assert "fleetQuantile" not in got and "curveMean" not in got
    s.destroy()
    # a slab that does not match its offsets never reaches the library
    bad = sh.obj(x=sh.ta(F32, x.reshape(-1)[:-5]), nodeOff=sh.ta(U32, off), nTime=sh.num(T), window=sh.num(W), q=sh.num(0.95))
    assert not sh.call("rangeView", bad) and "nodeOff" in sh.exception()[1]
    wrong = sh.obj(x=sh.ta(F64, x.astype(np.float64).reshape(-1)), nodeOff=sh.ta(U32, off), nTime=sh.num(T))
    assert not sh.call("rangeView", wrong) and "'x'" in sh.exception()[1]
    sh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("what", ["short_column", "wrong_type", "missing", "entries_not_records", "offsets_disagree"])
def test_addon_rejects_columns_the_library_would_overrun(ctx, what):
    """A shorter typed array must never reach igpu_alloc_rollup (it would be read past its end): the call throws."""
    cols = synth.synth_fleet(20, 300)
    sh = Shim()
    assert sh.load(), sh.exception()
    if what == "short_column":
        arg, needle = fleet_arg(sh, cols, capXe=sh.ta(I32, cols.node_cap_xe[:-1])), "capXe"
    elif what == "wrong_type":
        arg, needle = fleet_arg(sh, cols, podPhase=sh.ta(U32, cols.pod_phase.astype(np.uint32))), "podPhase"
    elif what == "missing":
        arg, needle = fleet_arg(sh, cols, allocOwn=None), "allocOwn"
    elif what == "entries_not_records":
        arg, needle = fleet_arg(sh, cols, entries=sh.ta(U8, cols.entries.view(np.uint8)[:-3])), "entries"
    else:
        arg, needle = fleet_arg(sh, cols, podEntryOff=sh.ta(U32, cols.pod_entry_off + 1)), "entries"
    assert not sh.call("rollupFleet", arg)
    code, msg = sh.exception()
    assert needle in msg and code.startswith("ERR_INVALID_ARG"), (code, msg)
    assert sh.call("rollupFleet", fleet_arg(sh, cols)), sh.exception()  # and the addon keeps working
    sh.close()
