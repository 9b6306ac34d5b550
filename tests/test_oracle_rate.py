"""CPU: the N3 oracle (oracle_counters_rate = Prometheus extrapolatedRate on a regular grid).

prometheus/prometheus is a server the reference only queries (metrics.ts:106-109); it is not vendored under
/root/reference, so these known answers are restated from promql/testdata/functions.test as recalled, and each was
re-derived by hand from the published algorithm below.  Parity for N3 is therefore "unpinned by the reference".
A second, independently written pure-Python evaluation guards the C++ restatement against transcription slips.
"""
import math

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200.native import RATE_CLOSED_LEFT, RATE_GAUGE, RATE_INCREASE, Counters

MIN5 = 300_000


def py_rate(row, e, step_ms, K, options):
    """One evaluation, written from the prose of the algorithm (not from the C++)."""
    closed, is_rate, is_counter = options & RATE_CLOSED_LEFT, not options & RATE_INCREASE, not options & RATE_GAUGE
    end, start = e * step_ms, e * step_ms - K * step_ms
    pts = []
    for i in range(max(0, e - K), min(len(row) - 1, e) + 1):
        t = i * step_ms
        inside = (start <= t <= end) if closed else (start < t <= end)
        if inside and not math.isnan(row[i]):
            pts.append((t, float(row[i])))
    if len(pts) < 2:
        return math.nan
    (t0, v0), (t1, v1) = pts[0], pts[-1]
    result = v1 - v0
    if is_counter:
        for (_, a), (_, b) in zip(pts, pts[1:]):
            if b < a:
                result += a
    to_start, to_end, sampled = (t0 - start) / 1000, (end - t1) / 1000, (t1 - t0) / 1000
    avg = sampled / (len(pts) - 1)
    limit = avg * 1.1
    if to_start >= limit:
        to_start = avg / 2
    if is_counter and result > 0 and v0 >= 0:
        to_zero = sampled * (v0 / result)
        if to_zero < to_start:
            to_start = to_zero
    if to_end >= limit:
        to_end = avg / 2
    factor = (sampled + to_start + to_end) / sampled
    if is_rate:
        factor /= (K * step_ms) / 1000
    return result * factor


def grid(*rows):
    n = max(len(r) for r in rows)
    v = np.full((len(rows), n), np.nan)
    for i, r in enumerate(rows):
        v[i, : len(r)] = r
    return v


def test_promql_known_answers_v2_range(oracle_lib):
    # load 5m: http_requests{path="/foo"} 0+10x10 ; {path="/bar"} 0+10x5 0+10x5 ; testcounter_reset_middle 0+10x4 0+10x5
    foo = list(range(0, 110, 10))
    bar = [0, 10, 20, 30, 40, 50, 0, 10, 20, 30, 40, 50]
    mid = [0, 10, 20, 30, 40, 0, 10, 20, 30, 40, 50]
    v = grid(foo, bar, mid)
    # eval instant at 50m rate(...[50m]) -> 0.03333..., 0.03, 0.03 ; increase -> 100, 90, 90   (closed-left range of 2.x)
    prm = Counters.params(MIN5, 10, first_eval=10, options=RATE_CLOSED_LEFT)
    _, r = oracle_lib.counters_rate(v, prm)
    assert r[:, 0] == pytest.approx([100 / 3000, 0.03, 0.03], rel=1e-15)
    prm = Counters.params(MIN5, 10, first_eval=10, options=RATE_CLOSED_LEFT | RATE_INCREASE)
    _, r = oracle_lib.counters_rate(v, prm)
    assert r[:, 0] == pytest.approx([100, 90, 90], rel=1e-15)


def test_left_open_range_of_v3(oracle_lib):
    # (0, 50m] drops the sample at t=0: /foo sees 10..100 over 45 m and extrapolates one step to the range start
    v = grid(list(range(0, 110, 10)))
    _, r = oracle_lib.counters_rate(v, Counters.params(MIN5, 10, first_eval=10))
    assert r[0, 0] == pytest.approx(90 * (3000 / 2700) / 3000, rel=1e-15)


def test_zero_point_clamp(oracle_lib):
    # a counter that started just before the range must not be extrapolated below zero:
    # samples 1, 11, 21 at steps 5..7 of a 10-step range -> duration_to_zero = 20 s * (1 / 20) = 1 s
    row = [math.nan] * 5 + [1, 11, 21] + [math.nan] * 3
    _, r = oracle_lib.counters_rate(grid(row), Counters.params(10_000, 10, first_eval=10, options=RATE_INCREASE))
    sampled, avg = 20.0, 10.0
    assert r[0, 0] == pytest.approx(20 * (sampled + 1.0 + avg / 2) / sampled, rel=1e-15)


def test_fewer_than_two_samples_is_nan(oracle_lib):
    row = [math.nan, 5.0, math.nan, math.nan, math.nan, math.nan, 7.0]
    _, r = oracle_lib.counters_rate(grid(row), Counters.params(1000, 3, first_eval=0))
    assert np.isnan(r[0, [0, 1, 2, 3, 4, 5, 6]]).all()
    _, r = oracle_lib.counters_rate(grid(row), Counters.params(1000, 6, first_eval=6))
    assert r[0, 0] == pytest.approx(py_rate(row, 6, 1000, 6, 0), rel=1e-15) and not math.isnan(r[0, 0])


@pytest.mark.parametrize("options", [0, RATE_CLOSED_LEFT, RATE_INCREASE, RATE_GAUGE, RATE_GAUGE | RATE_INCREASE | RATE_CLOSED_LEFT])
@pytest.mark.parametrize("K,stride,first", [(1, 1, 0), (2, 1, 0), (20, 1, 19), (20, 7, 3), (33, 1, 0), (47, 5, 11)])
def test_cpp_restatement_equals_python_restatement(oracle_lib, options, K, stride, first):
    rng = np.random.default_rng(K * 1000 + stride * 10 + options)
    S, T = 6, 260
    v = np.cumsum(rng.integers(0, 50, (S, T)).astype(np.float64), axis=1)
    for s in range(S):  # counter resets, gaps, a dead stretch, a late start
        for at in rng.integers(1, T, 4):
            v[s, at:] -= v[s, at] - float(rng.integers(0, 3))
        v[s, rng.random(T) < 0.1 * s] = np.nan
    v[2, 100:170] = np.nan
    v[3, :40] = np.nan
    prm = Counters.params(15_000, K, first_eval=first, eval_stride=stride, options=options)
    o32, o64 = oracle_lib.counters_rate(v, prm, threads=3)
    n_out = (T - 1 - first) // stride + 1
    assert o64.shape == (S, n_out)
    for s in range(S):
        for j in range(n_out):
            want = py_rate(v[s], first + j * stride, 15_000, K, options)
            got = o64[s, j]
            assert (math.isnan(want) and math.isnan(got)) or got == want, (s, j, got, want)
    assert np.array_equal(o32.view(np.uint32), o64.astype(np.float32).view(np.uint32))


def test_generator_is_a_counter(oracle_lib):
    v = oracle_lib.counters_generate(8, 3000, seed=11, global_first=5, step_ms=15_000, reset_period=700, missing_per_million=20_000)
    w = oracle_lib.counters_generate(4, 3000, seed=11, global_first=7, step_ms=15_000, reset_period=700, missing_per_million=20_000)
    assert np.array_equal(v[2:6].view(np.uint64), w.view(np.uint64))  # rows depend on the fleet-wide index only
    assert 0.015 < np.isnan(v).mean() < 0.025
    clean = oracle_lib.counters_generate(8, 3000, seed=11, global_first=5, step_ms=15_000, reset_period=700)
    d = np.diff(clean, axis=1)
    assert ((d < 0).sum(axis=1) >= 4).all() and ((d < 0).sum(axis=1) <= 5).all()  # one drop per reset period
    assert (clean >= 0).all()
    # between resets the rate is the series' power: 50..300 W
    _, r = oracle_lib.counters_rate(clean, Counters.params(15_000, 20))
    assert (r > 45).all() and (r < 305).all()


def test_rate_len(oracle_lib):
    import ctypes as C

    L = oracle_lib.lib()
    for T, first, stride, want in [(10, 0, 1, 10), (10, 9, 1, 1), (10, 10, 1, 0), (10, 3, 4, 2), (1, 0, 5, 1)]:
        prm = Counters.params(1000, 2, first_eval=first, eval_stride=stride)
        assert L.oracle_counters_rate_len(T, C.byref(prm)) == want
