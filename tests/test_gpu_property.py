"""GPU, property-based: arbitrary Kubernetes-shaped junk -> JSON text -> native marshal -> CUDA rollup (+ per-container rows)
-> views, against the literal per-object restatement of the reference (oracle/ref_objects.py).  The CPU twins of this test
(test_oracle_property.py, test_marshal_property.py) found two real divergences in round 1; this one puts the device path
under the same pressure."""
import json

import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import k8s, native

from .parity import check_against_objects
from .test_marshal_property import node, pod

pytestmark = pytest.mark.gpu


@settings(max_examples=200, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large, HealthCheck.function_scoped_fixture])
@given(nodes=st.lists(node, max_size=5), pods=st.lists(pod, max_size=10), grouped=st.booleans())
def test_json_text_to_cuda_equals_the_object_restatement(ctx, nodes, pods, grouped):
    nodes, pods = json.loads(json.dumps(nodes)), json.loads(json.dumps(pods))  # what a List response can carry
    try:
        r = k8s.rollupFleetJson(json.dumps(nodes), json.dumps(pods), ctx, group_by_node=grouped, container_rows=True)
    except native.IgpuError as e:
        assert e.status == -6, e  # outside a documented limit (resource names, containers, 2^53): rejected by design
        return
    check_against_objects(nodes, pods, r.cols, r.out, r.order, r.pos, container_rows=r.container_rows)


# ---- rate() / increase() / delta(): every kernel of csrc/rate.cu against the serial restatement --------------------------
import numpy as np

from headlamp_intel_gpu_plugin_b200 import synth
from headlamp_intel_gpu_plugin_b200.native import RATE_CLOSED_LEFT, RATE_GAUGE, RATE_INCREASE, Counters


@st.composite
def counter_case(draw):
    S = draw(st.integers(1, 4))
    T = draw(st.integers(1, 700))
    K = draw(st.sampled_from([1, 2, 3, 19, 20, 31, 32, 33, 34, 40, 60, 64, 65, 128, 255, 256, 257, 300]))
    opts = draw(st.sampled_from([0, RATE_CLOSED_LEFT, RATE_INCREASE, RATE_GAUGE, RATE_GAUGE | RATE_CLOSED_LEFT]))
    stride = draw(st.sampled_from([1, 1, 2, 7]))
    first = draw(st.one_of(st.none(), st.integers(0, max(0, T - 1))))
    seed = draw(st.integers(0, 2**31 - 1))
    gap = draw(st.sampled_from([0.0, 0.0, 0.001, 0.02, 0.3, 0.9]))
    resets = draw(st.integers(0, 4))
    tier = draw(st.sampled_from(["0", "1", "2"]))
    return S, T, K, opts, stride, first, seed, gap, resets, tier


@settings(max_examples=150, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])
@given(case=counter_case())
def test_rate_any_shape_is_bit_exact(ctx, oracle_lib, switch, case):
    S, T, K, opts, stride, first, seed, gap, resets, tier = case
    rng = np.random.default_rng(seed)
    v = np.cumsum(rng.integers(0, 4000, (S, T)).astype(np.float64) * 0.125, axis=1) + rng.integers(0, 1000, (S, 1))
    for s in range(S):
        for at in rng.integers(0, T, resets):
            v[s, at:] -= v[s, at] - float(rng.integers(0, 2))   # a counter reset (sometimes to exactly zero: the clamp's edge)
        v[s, rng.random(T) < gap] = np.nan
    prm = Counters.params(15_000, K, first_eval=first, eval_stride=stride, options=opts)
    switch("IGPU_RATE_TIER", tier)
    c = ctx.counters_create(S, T)
    c.upload(v)
    n_out = c.rate_len(prm)
    want32, want64 = oracle_lib.counters_rate(v, prm)
    if n_out == 0:
        assert want64.shape[1] == 0
        c.destroy()
        return
    dst = ctx.series_create(S, n_out, synth.even_node_offsets(1, S))
    got64 = c.rate(prm, dst, want_f64=True)
    got32 = dst.download(0, S)
    dst.destroy()
    c.destroy()
    same = lambda a, b: np.array_equal(a.view(np.uint64 if a.dtype == np.float64 else np.uint32),
                                       b.view(np.uint64 if b.dtype == np.float64 else np.uint32)) or np.all((a == b) | (np.isnan(a) & np.isnan(b)))
    assert same(got64, want64), (case, np.argwhere(~((got64 == want64) | (np.isnan(got64) & np.isnan(want64))))[:5])
    assert same(got32, want32)


# ---- series rollup: every window / row shape through whichever kernel takes it, against the oracle -------------------------
@st.composite
def series_case(draw):
    S = draw(st.integers(1, 12))
    T = draw(st.one_of(st.integers(1, 400), st.integers(2000, 6100)))
    W = draw(st.sampled_from([1, 2, 4, 6, 7, 10, 15, 20, 30, 60, 64, 90, 100, 120, 150, 192, 256, 300, 404, 1024, 1440, 3000]))
    q = draw(st.sampled_from([0.0, 0.01, 0.5, 0.95, 0.999, 1.0]))
    mode = draw(st.sampled_from([0, 0, 1]))
    nan_ppm = draw(st.sampled_from([0, 0, 1000, 50000, 600000]))
    seed = draw(st.integers(0, 2**31 - 1))
    n_cuts = draw(st.integers(0, min(S, 4)))
    cuts = sorted(draw(st.lists(st.integers(0, S), min_size=n_cuts, max_size=n_cuts)))
    return S, T, W, q, mode, nan_ppm, seed, cuts


@settings(max_examples=120, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])
@given(case=series_case())
def test_series_rollup_any_shape_matches_the_oracle(ctx, oracle_lib, switch, case):
    S, T, W, q, mode, nan_ppm, seed, cuts = case
    x = synth.synth_series(S, T, seed=seed % 1000, nan_per_million=nan_ppm)
    if S > 1 and seed % 3 == 0:
        x[1] = np.nan  # a chip that never reported
    off = np.array([0] + cuts + [S], np.uint32)  # nodes of any size, empty ones included
    s = ctx.series_create(S, T, off)
    s.upload(x)
    s.keep_window_counts(True)
    (mn, me, mx), qv, st_, fl = s.rollup(W, q, mode)
    omn, ome, omx = oracle_lib.window_stats(x, W)
    oq, ost, ofl = oracle_lib.quantile_by_node(x, off, q, mode)
    eq = lambda a, b: np.all((a == b) | (np.isnan(a) & np.isnan(b)))
    close = lambda a, b, r=1e-5: np.all((np.isnan(a) & np.isnan(b)) | (np.abs(np.asarray(a, np.float64) - b) <= r * np.maximum(np.abs(b), 1e-30)))
    assert eq(mn, omn) and eq(mx, omx) and close(me, ome), case
    assert close(qv, oq, 1e-12) and np.array_equal(st_["count"], ost["count"]) and eq(st_["min"], ost["min"]) and eq(st_["max"], ost["max"]), case
    node, fleet = s.window_rollup(W)   # from the kept planes where the single-read kernel ran, from the slab otherwise
    onode, ofleet = oracle_lib.window_rollup(x, off, W)
    for got, want in ((node, onode), (fleet, ofleet)):
        assert np.array_equal(got["count"], want["count"]) and eq(got["min"], want["min"]) and eq(got["max"], want["max"]), case
        assert close(got["mean"], want["mean"]), case
    switch("IGPU_FLEET_QUANTILE_MIN_SAMPLE", "1")
    (gq, gc), (wq, wc) = s.quantile_fleet(q, mode), oracle_lib.quantile_fleet(x, q, mode)
    assert gc == wc and (gq == wq or (np.isnan(gq) and np.isnan(wq))), case  # (no valid sample at all: NaN)
    s.destroy()
