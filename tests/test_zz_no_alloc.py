"""GPU test of igpu_power_join_sum_static, the entry point for callers that must not allocate
(a server whose own pool owns all device memory, a scrape loop that cannot take the device-wide stall
of cudaMalloc / cudaFree, a program where only its own allocator may touch the GPU).

The caller — here the test, through torch — owns every input and output buffer on the device and
passes raw device pointers through the C ABI; the library's only working memory is its built-in
array.  The result must equal igpu_power_join_sum's on the same input, on two calls in a row: the
second call is a different, smaller fleet, so anything the first call left in the shared array would
show up as a wrong join.

This file sorts last on purpose (pytest collects test files alphabetically), so a run that stops at
the first failure has already run every other test.  It works alone in a fresh process as well."""
import ctypes as C

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import metrics, native, synth

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
