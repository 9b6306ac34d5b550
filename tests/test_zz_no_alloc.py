"""GPU test of igpu_power_join_sum_static, the entry point for callers that must not allocate
(a server whose own pool owns all device memory, a scrape loop that cannot take the device-wide stall
of cudaMalloc / cudaFree, a program where only its own allocator may touch the GPU).

The caller — here the test, through torch — owns every input and output buffer on the device and
passes raw device pointers through the C ABI; the library's only working memory is its built-in
array.  The result must equal igpu_power_join_sum's on the same input, on two calls in a row: the
second call is a different, smaller fleet, so anything the first call left in the shared array would
show up as a wrong join.  Because both entry points share their kernels, the result is also held
against the CPU oracle and the per-object restatement of metrics.ts:118-149 directly.

This file sorts last on purpose (pytest collects test files alphabetically), so a run that stops at
the first failure has already run every other test.  It works alone in a fresh process as well."""
import ctypes as C

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import metrics, native, synth
from oracle import columnar as oracle
from oracle import ref_objects as ro

from .parity import nan_eq

pytestmark = pytest.mark.gpu
RTOL = 1e-5  # the float tolerance of the other power tests (north star: 1e-5 relative)


def _on_device(p):
    """Copies the marshalled columns to device tensors owned by the caller; returns (struct, keepalive)."""
    import torch

    keep = {}
    cs = native.PowerColsC()
    cs.n_chips, cs.n_energy = len(p.chip_chip), len(p.energy_chip)
    cs.n_power_max, cs.n_uname = len(p.power_max_chip), len(p.uname_instance)
    cs.n_chip_ids, cs.n_instance_ids = p.n_chip_ids, p.n_instance_ids
    for name in "chip_chip chip_instance energy_chip energy_value power_max_chip power_max_value uname_instance uname_node".split():
        a = getattr(p, name)
        # torch has no uint32 arithmetic, but it can hold the bytes: view ids as int32
        t = torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a).cuda() if a.size else None
        keep[name] = t
        setattr(cs, name, t.data_ptr() if t is not None else None)
    return cs, keep


def _device_outputs(n):
    import torch

    t = {
        "chip_flags": torch.full((max(n, 1),), 0x7F, dtype=torch.uint8, device="cuda"),
        "chip_watts": torch.full((max(n, 1),), -1.0, dtype=torch.float64, device="cuda"),
        "chip_max_watts": torch.full((max(n, 1),), -1.0, dtype=torch.float64, device="cuda"),
        "chip_node": torch.full((max(n, 1),), -7, dtype=torch.int32, device="cuda"),
        "fleet": torch.full((4,), -1.0, dtype=torch.float64, device="cuda"),
    }
    os_ = native.PowerOutC(*[t[k].data_ptr() for k in ("chip_flags", "chip_watts", "chip_max_watts", "chip_node", "fleet")])
    return os_, t


def _close(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return bool(np.all((np.isnan(a) & np.isnan(b)) | (np.abs(a - b) <= RTOL * np.maximum(np.abs(b), 1e-300))))


def _dirty(ch, en, mx, un):
    en.append({"metric": {"chip": ch[0]["metric"]["chip"], "instance": "other"}, "value": [0, "7.5"]})  # last write wins
    mx.append({"metric": {"chip": ch[2]["metric"]["chip"]}, "value": [0, "NaN"]})                        # poisons the TDP sum
    un.append({"metric": {"instance": ch[5]["metric"]["instance"], "node": "renamed"}, "value": [0, "1"]})
    ch.append({"metric": {"chip": "lonely"}, "value": [0, "1"]})


def _host(out, n):
    got = {k: v.cpu().numpy()[:n] for k, v in out.items() if k != "fleet"}
    got["chip_node"] = got["chip_node"].view(np.uint32)
    got["fleet"] = out["fleet"].cpu().numpy()
    return got


def _check_against_oracles(p, out, results=None):
    """Static join vs the columnar CPU oracle and (given the Prometheus result objects) vs the literal
    restatement of fetchGpuMetrics' join and MetricsPage's sums."""
    n = len(p.chip_chip)
    got, want = _host(out, n), oracle.power_join_sum(p)
    assert np.array_equal(got["chip_flags"], want["chip_flags"])
    has_w, has_m = (got["chip_flags"] & K.POWER_HAS_WATTS) != 0, (got["chip_flags"] & K.POWER_HAS_MAX) != 0
    has_n = (got["chip_flags"] & K.POWER_HAS_NODE) != 0
    assert np.array_equal(got["chip_node"][has_n], want["chip_node"][has_n])
    assert _close(got["chip_watts"][has_w], want["chip_watts"][has_w])
    assert _close(got["chip_max_watts"][has_m], want["chip_max_watts"][has_m])
    assert _close(got["fleet"], want["fleet"]), (got["fleet"], want["fleet"])
    if results is not None:
        rows = metrics._decode_chips(p, got)
        ref = ro.join_gpu_metrics(*results)
        assert nan_eq(rows, ref)
        fp = ro.fleet_power(ref)
        assert _close(got["fleet"][0], fp["totalWatts"]) and _close(got["fleet"][1], fp["maxTotalWatts"])
        assert (fp["pct"] is None and np.isnan(got["fleet"][2])) or got["fleet"][2] == fp["pct"]
        assert got["fleet"][3] == len(ref)


def _run_static(p, device=0, stream=None):
    import torch

    lib = native.load_library()
    cs, keep = _on_device(p)
    os_, out = _device_outputs(len(p.chip_chip))
    st = lib.igpu_power_join_sum_static(device, C.c_void_p(stream) if stream else None, C.byref(cs), C.byref(os_))
    torch.cuda.synchronize()
    return st, out, keep


def test_static_join_equals_allocating_join_on_two_calls_in_a_row(ctx):
    import torch

    lib = native.load_library()
    stream = torch.cuda.current_stream().cuda_stream
    for n_nodes, chips, dirty in ((300, 8, True), (40, 2, False)):  # big first, small second: stale tables would bite
        ch, en, mx, un = synth.synth_power_results(n_nodes, chips, seed=n_nodes)
        if dirty:
            _dirty(ch, en, mx, un)
        p = metrics.marshalMetrics(ch, en, mx, un)
        want = ctx.power_join_sum(p)  # the existing, allocating entry point
        cs, keep = _on_device(p)
        os_, out = _device_outputs(len(p.chip_chip))
        st = lib.igpu_power_join_sum_static(0, C.c_void_p(stream) if stream else None, C.byref(cs), C.byref(os_))
        assert st == 0, lib.igpu_last_error(None)
        torch.cuda.synchronize()
        n = len(p.chip_chip)
        assert np.array_equal(out["chip_flags"].cpu().numpy()[:n], want["chip_flags"])
        assert np.array_equal(out["chip_node"].cpu().numpy()[:n].view(np.uint32), want["chip_node"])
        assert _close(out["chip_watts"].cpu().numpy()[:n], want["chip_watts"])
        assert _close(out["chip_max_watts"].cpu().numpy()[:n], want["chip_max_watts"])
        assert _close(out["fleet"].cpu().numpy(), want["fleet"]), (out["fleet"].cpu().numpy(), want["fleet"])
        _check_against_oracles(p, out, (ch, en, mx, un))
        del keep


def test_static_join_refuses_what_does_not_fit(ctx):
    import torch

    lib = native.load_library()
    ch, en, mx, un = synth.synth_power_results(3, 2)
    p = metrics.marshalMetrics(ch, en, mx, un)
    cs, keep = _on_device(p)
    os_, out = _device_outputs(len(p.chip_chip))
    cs.n_chip_ids = native.STATIC_JOIN_MAX_IDS + 1  # one id more than the built-in array can index
    assert lib.igpu_power_join_sum_static(0, None, C.byref(cs), C.byref(os_)) == -6  # IGPU_ERR_RANGE
    msg = lib.igpu_last_error(None).decode()
    assert "scratch limit" in msg and str(native.STATIC_JOIN_MAX_IDS) in msg
    torch.cuda.synchronize()
    assert float(out["fleet"][0]) == -1.0  # nothing was launched
    os_.fleet = None
    cs.n_chip_ids = p.n_chip_ids
    assert lib.igpu_power_join_sum_static(0, None, C.byref(cs), C.byref(os_)) == -1  # IGPU_ERR_ARG: outputs are required


@pytest.mark.parametrize("watts,tdp,pct", [(["200"], ["100"], 100.0),         # min(100, 200)
                                           (["1"], ["8"], 13.0),              # 12.5: Math.round goes up
                                           (["45.3", "0"], ["120", "0"], 38.0),
                                           (["5"], ["0"], None),              # max <= 0: no bar
                                           (["0.5", "0.5"], ["40", "160"], 1.0)])  # 0.5 % rounds to 1
def test_static_join_power_bar_percentage(ctx, watts, tdp, pct):
    """The percentage is computed on the device in this entry point (PowerBar, MetricsPage.tsx:50-53)."""
    ch = [{"metric": {"chip": f"c{i}", "instance": "a:9100"}, "value": [0, "1"]} for i in range(len(watts))]
    en = [{"metric": {"chip": f"c{i}"}, "value": [0, w]} for i, w in enumerate(watts)]
    mx = [{"metric": {"chip": f"c{i}"}, "value": [0, t]} for i, t in enumerate(tdp)]
    un = [{"metric": {"instance": "a:9100", "nodename": "na"}, "value": [0, "1"]}]
    p = metrics.marshalMetrics(ch, en, mx, un)
    st, out, _keep = _run_static(p)
    assert st == 0
    f = out["fleet"].cpu().numpy()
    assert (pct is None and np.isnan(f[2])) or f[2] == pct, f
    _check_against_oracles(p, out, (ch, en, mx, un))


def test_static_join_at_exactly_the_limit(ctx):
    """n_chip_ids == n_instance_ids == IGPU_STATIC_JOIN_MAX_IDS: the last id sits in the last word of each table."""
    m = native.STATIC_JOIN_MAX_IDS
    rng = np.random.default_rng(3)
    n = 5000
    chip = np.concatenate([[0, m - 1], rng.integers(0, m, n - 2)]).astype(np.uint32)
    inst = np.concatenate([[m - 1, 0], rng.integers(0, m, n - 2)]).astype(np.uint32)
    p = K.PowerColumns(chip_chip=chip, chip_instance=inst, energy_chip=chip[::2].copy(), energy_value=rng.uniform(1, 200, (n + 1) // 2),
                       power_max_chip=chip[::-1].copy(), power_max_value=rng.uniform(150, 300, n),
                       uname_instance=inst[1::2].copy(), uname_node=rng.integers(0, 50, n // 2).astype(np.uint32),
                       n_chip_ids=m, n_instance_ids=m)
    for _ in range(2):
        st, out, _keep = _run_static(p)
        assert st == 0, native.load_library().igpu_last_error(None)
        _check_against_oracles(p, out)


def test_static_join_unknown_device_touches_nothing(ctx):
    import torch

    lib = native.load_library()
    ch, en, mx, un = synth.synth_power_results(3, 2)
    p = metrics.marshalMetrics(ch, en, mx, un)
    for bad in (torch.cuda.device_count(), 63, -1, 64):
        st, out, _keep = _run_static(p, device=bad)
        assert st == -1, (bad, st)  # IGPU_ERR_ARG
        assert float(out["fleet"][0]) == -1.0 and int(out["chip_flags"][0]) == 0x7F  # nothing ran on the current device
    assert "device" in lib.igpu_last_error(None).decode()
    st, out, _keep = _run_static(p)  # and the entry point still works afterwards
    assert st == 0
    _check_against_oracles(p, out, (ch, en, mx, un))


def test_static_join_second_device_default_stream(ctx):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ch, en, mx, un = synth.synth_power_results(20, 4)
    p = metrics.marshalMetrics(ch, en, mx, un)
    with torch.cuda.device(1):
        st, out, _keep = _run_static(p, device=1, stream=None)
        torch.cuda.synchronize()
        assert st == 0
        _check_against_oracles(p, out, (ch, en, mx, un))
