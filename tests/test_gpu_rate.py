"""GPU parity (through the C ABI): N3 — rate()/increase()/delta() over energy counters (csrc/rate.cu) against
oracle_counters_rate.  The kernel adds the counter-reset terms in time order with explicitly rounded float64
operations, so the bar here is bit-exact float64 (far inside the north star's 1e-5 relative), and the float32
slab it writes is the correctly rounded float64 result."""
import os

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import native, synth
from headlamp_intel_gpu_plugin_b200.native import RATE_CLOSED_LEFT, RATE_GAUGE, RATE_INCREASE, Counters

pytestmark = pytest.mark.gpu
STEP = 15_000


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint64 if a.dtype == np.float64 else np.uint32),
                          np.ascontiguousarray(b).view(np.uint64 if b.dtype == np.float64 else np.uint32)) or \
        np.all((a == b) | (np.isnan(a) & np.isnan(b)))


def run_rate(ctx, v, prm):
    S, T = v.shape
    c = ctx.counters_create(S, T)
    c.upload(v)
    n_out = c.rate_len(prm)
    dst = ctx.series_create(S, max(n_out, 1), synth.even_node_offsets(1, S)) if n_out else None
    if dst is None:
        c.destroy()
        return np.empty((S, 0), np.float32), np.empty((S, 0))
    r64 = c.rate(prm, dst, want_f64=True)
    r32 = dst.download(0, S)
    dst.destroy()
    c.destroy()
    return r32, r64


def dirty_counters(S, T, seed):
    rng = np.random.default_rng(seed)
    v = np.cumsum(rng.integers(0, 5000, (S, T)).astype(np.float64) * 0.25, axis=1) + rng.integers(0, 10**6, (S, 1))
    for s in range(S):
        for at in rng.integers(1, max(T, 2), 1 + s % 5):
            if at < T:
                v[s, at:] -= v[s, at] - float(rng.integers(0, 3))
        v[s, rng.random(T) < 0.02 * (s % 6)] = np.nan
    if S > 3:
        v[2, T // 3: T // 3 + 90] = np.nan   # a dead stretch longer than two blocks
        v[3, : T // 4] = np.nan              # a series that starts late (zero-point clamp region follows)
        v[3, T // 4:] -= np.nanmin(v[3, T // 4:])
    return v


# The warp path has two kernels (rate.cu: streaming, and one that takes every block from the bit masks); which one runs is
# decided on the device from a gap statistic of the slab.  Every shape below runs three ways: as decided, and each forced.
TIERS = pytest.mark.parametrize("tier", ["auto", "1", "2"])


def pick_tier(switch, tier):
    if tier != "auto":
        switch("IGPU_RATE_TIER", tier)


@TIERS
@pytest.mark.parametrize("options", [0, RATE_CLOSED_LEFT, RATE_INCREASE, RATE_GAUGE, RATE_GAUGE | RATE_INCREASE | RATE_CLOSED_LEFT])
@pytest.mark.parametrize("K", [1, 2, 5, 20, 32, 33])
def test_rate_bit_exact_vs_oracle(ctx, oracle_lib, switch, options, K, tier):
    if K == 33 and options & RATE_CLOSED_LEFT:
        pytest.skip("34 samples per range: covered by the generic-kernel test")
    pick_tier(switch, tier)
    v = dirty_counters(13, 1000 + K, seed=K * 8 + options)
    prm = Counters.params(STEP, K, options=options)
    want32, want64 = oracle_lib.counters_rate(v, prm, threads=4)
    got32, got64 = run_rate(ctx, v, prm)
    assert bits_equal(got64, want64)
    assert bits_equal(got32, want32)


@pytest.mark.parametrize("first,stride", [(0, 1), (5, 1), (19, 3), (31, 32), (32, 33), (100, 7), (999, 1), (1000, 1)])
@TIERS
def test_first_eval_and_stride(ctx, oracle_lib, switch, first, stride, tier):
    pick_tier(switch, tier)
    v = dirty_counters(7, 1000, seed=first + stride)
    prm = Counters.params(STEP, 20, first_eval=first, eval_stride=stride)
    want32, want64 = oracle_lib.counters_rate(v, prm)
    got32, got64 = run_rate(ctx, v, prm)
    assert got64.shape == want64.shape
    assert bits_equal(got64, want64) and bits_equal(got32, want32)


@pytest.mark.parametrize("T", [1, 2, 31, 32, 33, 63, 64, 65, 2048 + 19, 2 * 2048 + 40, 4133])
@TIERS
def test_ragged_lengths_and_segment_seams(ctx, oracle_lib, switch, T, tier):
    pick_tier(switch, tier)
    # 2048 grid samples per work item: lengths around the seams exercise the warm-up blocks of the second item
    v = dirty_counters(5, T, seed=T)
    prm = Counters.params(STEP, 20, first_eval=0)
    want32, want64 = oracle_lib.counters_rate(v, prm)
    got32, got64 = run_rate(ctx, v, prm)
    assert bits_equal(got64, want64) and bits_equal(got32, want32)


@pytest.mark.parametrize("K,options", [(34, 0), (33, RATE_CLOSED_LEFT), (80, 0), (240, RATE_INCREASE), (256, 0), (255, RATE_CLOSED_LEFT),
                                       (60, RATE_GAUGE), (257, 0), (256, RATE_CLOSED_LEFT), (400, RATE_INCREASE), (1000, 0), (2890, 0)])
@pytest.mark.parametrize("stride", [1, 3])
def test_long_ranges(ctx, oracle_lib, switch, K, options, stride):
    """Ranges of more than 33 samples take the long-range streaming kernel (the first sample of a clean range read from its
    shared-memory ring up to 256 samples back, from global memory beyond; dirty blocks marked and evaluated by a second kernel,
    cooperatively up to 256 samples, lane by lane beyond).  Same bits as the oracle, and as the thread-per-evaluation kernel."""
    v = dirty_counters(6, 2900, seed=K)
    v[4] = np.cumsum(np.full(2900, 7.25))        # a clean series: every block on the fast path
    v[5, 1500:] -= v[5, 1500] - 3.0              # one reset in an otherwise clean series (and a clamp region after it)
    v[5, :1500] = np.cumsum(np.full(1500, 2.5))
    v[5, 1500:] = 3.0 + np.cumsum(np.full(1400, 2.5))
    prm = Counters.params(STEP, K, options=options, eval_stride=stride)
    want32, want64 = oracle_lib.counters_rate(v, prm)
    got32, got64 = run_rate(ctx, v, prm)
    assert bits_equal(got64, want64) and bits_equal(got32, want32)
    switch("IGPU_RATE_FORCE_GENERIC")
    gen32, gen64 = run_rate(ctx, v, prm)
    assert bits_equal(gen64, got64) and bits_equal(gen32, got32)


@pytest.mark.parametrize("first,T", [(0, 700), (5, 700), (300, 1301), (59, 64), (10, 40), (0, 4133)])
def test_long_ranges_first_eval_and_seams(ctx, oracle_lib, first, T):
    """Evaluations whose range is cut by the start of the series, rows shorter than a chunk, and work-item seams."""
    v = dirty_counters(5, T, seed=first + T)
    v[4] = np.cumsum(np.full(T, 1.5))
    prm = Counters.params(STEP, 60, first_eval=first)
    want32, want64 = oracle_lib.counters_rate(v, prm)
    got32, got64 = run_rate(ctx, v, prm)
    assert bits_equal(got64, want64) and bits_equal(got32, want32)


def test_generic_and_warp_kernels_agree(ctx, oracle_lib, switch):
    v = dirty_counters(9, 3000, seed=77)
    prm = Counters.params(STEP, 20)
    a32, a64 = run_rate(ctx, v, prm)
    switch("IGPU_RATE_FORCE_GENERIC")
    b32, b64 = run_rate(ctx, v, prm)
    assert bits_equal(a64, b64) and bits_equal(a32, b32)


def test_promql_known_answers_on_device(ctx):
    foo = list(range(0, 110, 10)) + [np.nan]
    bar = [0, 10, 20, 30, 40, 50, 0, 10, 20, 30, 40, 50]
    mid = [0, 10, 20, 30, 40, 0, 10, 20, 30, 40, 50, np.nan]
    v = np.array([foo, bar, mid], np.float64)
    _, r = run_rate(ctx, v, Counters.params(300_000, 10, first_eval=10, options=RATE_CLOSED_LEFT))
    assert r[:, 0] == pytest.approx([100 / 3000, 0.03, 0.03], rel=1e-15)
    _, r = run_rate(ctx, v, Counters.params(300_000, 10, first_eval=10, options=RATE_CLOSED_LEFT | RATE_INCREASE))
    assert r[:, 0] == pytest.approx([100, 90, 90], rel=1e-15)


def test_device_generator_matches_oracle(ctx, oracle_lib):
    c = ctx.counters_create(33, 2500)
    c.generate(synth.DEFAULT_SEED, global_first=1234, step_ms=STEP, reset_period=900, missing_per_million=3000)
    prm = Counters.params(STEP, 20)
    dst = ctx.series_create(33, c.rate_len(prm), synth.even_node_offsets(3, 11))
    got64 = c.rate(prm, dst, want_f64=True)
    host = oracle_lib.counters_generate(33, 2500, synth.DEFAULT_SEED, 1234, STEP, 900, 3000)
    _, want64 = oracle_lib.counters_rate(host, prm, threads=4)
    assert bits_equal(got64, want64)
    assert np.isnan(want64).sum() < want64.size // 100 and (want64[~np.isnan(want64)] > 0).all()
    dst.destroy()
    c.destroy()


def test_counters_to_rollup_pipeline(ctx, oracle_lib):
    """counters -> rate (device) -> window stats + per-node p95 (device): the range view fed by raw energy counters."""
    S, T, nodes = 48, 5779, 12   # 5760 rate points after a 20-step range
    c = ctx.counters_create(S, T)
    c.generate(7, step_ms=STEP, reset_period=2000, missing_per_million=500)
    prm = Counters.params(STEP, 20)
    off = synth.even_node_offsets(nodes, S // nodes)
    dst = ctx.series_create(S, c.rate_len(prm), off)
    c.rate(prm, dst)
    (wmin, wmean, wmax), q, st, fl = dst.rollup(60, 0.95)
    host = oracle_lib.counters_generate(S, T, 7, 0, STEP, 2000, 500)
    power, _ = oracle_lib.counters_rate(host, prm, threads=4)
    omin, omean, omax = oracle_lib.window_stats(power, 60, threads=4)
    oq, ost, ofl = oracle_lib.quantile_by_node(power, off, 0.95, threads=4)
    same = lambda a, b: np.all((a == b) | (np.isnan(a) & np.isnan(b)))
    assert same(wmin, omin) and same(wmax, omax)
    assert np.allclose(wmean, omean, rtol=1e-5, equal_nan=True)
    assert np.allclose(q, oq, rtol=1e-5, equal_nan=True)
    assert np.array_equal(st["count"], ost["count"])
    dst.destroy()
    c.destroy()


def test_host_api_energy_counters_to_power_series(ctx, oracle_lib):
    from headlamp_intel_gpu_plugin_b200 import metrics

    v = dirty_counters(16, 2000, seed=3)
    ec = metrics.EnergyCounters.from_array(v, step_ms=STEP, ctx=ctx)
    off = synth.even_node_offsets(4, 4)
    ps = ec.power(off, range_ms=300_000, eval_stride=4)
    want32, _ = oracle_lib.counters_rate(v, Counters.params(STEP, 20, eval_stride=4))
    assert ps.n_time == want32.shape[1]
    assert bits_equal(ps.h.download(0, 16), want32)
    with pytest.raises(ValueError):
        ec.power(off, range_ms=100_000)
    ps.close()
    ec.close()


def test_prometheus_text_to_p95_end_to_end(ctx, oracle_lib):
    """N3 whole chain: query_range response TEXT -> native ingest -> counters on the device -> rate() -> fused
    window stats + per-node p95, against the oracle chain fed with the same numbers."""
    import json

    from headlamp_intel_gpu_plugin_b200 import metrics

    S, T, nodes = 12, 500, 4
    start = 1_700_000_000_000
    v = oracle_lib.counters_generate(S, T, 99, 0, STEP, 150, 20_000)
    result = []
    for s_ in range(S):
        vals = [[(start + i * STEP) / 1000, repr(float(v[s_, i]))] for i in range(T) if not np.isnan(v[s_, i])]
        result.append({"metric": {"__name__": "node_hwmon_energy_joule_total", "chip": f"chip{s_}", "instance": f"10.0.0.{s_ // 3}:9100"}, "values": vals})
    text = json.dumps({"status": "success", "data": {"resultType": "matrix", "result": result}})
    slab, labels, dropped = metrics.marshalMatrixJson(text, start, STEP, T)
    assert dropped == 0 and labels[5] == ("chip5", "10.0.0.1:9100")
    assert np.all((slab == v) | (np.isnan(slab) & np.isnan(v)))  # repr() round-trips float64 exactly
    ec = metrics.EnergyCounters.from_array(slab, STEP, ctx)
    off = synth.even_node_offsets(nodes, S // nodes)
    ps = ec.power(off, range_ms=300_000)
    (wmin, wmean, wmax), q, st, fl = ps.rollup(60, 0.95)
    want32, _ = oracle_lib.counters_rate(v, Counters.params(STEP, 20))
    omin, omean, omax = oracle_lib.window_stats(want32, 60)
    oq, ost, ofl = oracle_lib.quantile_by_node(want32, off, 0.95)
    same = lambda a, b: np.all((a == b) | (np.isnan(a) & np.isnan(b)))
    assert same(wmin, omin) and same(wmax, omax) and np.allclose(wmean, omean, rtol=1e-5, equal_nan=True)
    assert np.allclose(q, oq, rtol=1e-5, equal_nan=True) and np.array_equal(st["count"], ost["count"])
    ps.close()
    ec.close()


def test_full_width_slab_linear_counters(ctx):
    """Size-independent property at a bench-sized shard: a counter that climbs p joules per second has rate p at
    every evaluation, whatever the range; checked on the device-side float32 slab through its min/max/p95."""
    S, T = 20_000, 10_099
    c = ctx.counters_create(S, T)
    c.generate(5, step_ms=STEP)   # no resets, no gaps: rate = the series' power +- jitter/range
    prm = Counters.params(STEP, 20)
    dst = ctx.series_create(S, c.rate_len(prm), synth.even_node_offsets(S // 8, 8))
    c.rate(prm, dst)
    t = ctx.timing()
    assert t["kernels"]["counter_rate"] > 0
    q, st, fl = dst.quantile_by_node(0.95)
    assert st["count"].sum() == S * (T - 19)
    assert st["min"].min() > 47.0 and st["max"].max() < 304.0  # a freshly booted counter may be clamped by up to 1/20
    # per-series spread: the jitter is < step/4 of energy, i.e. under 1/4/19 of the power over a 20-sample range
    rows = dst.download(0, 64)
    mid = np.median(rows, axis=1, keepdims=True)
    assert (np.abs(rows - mid) <= mid * (0.5 / 19) * 1.01 + 1e-3).all()
    dst.destroy()
    c.destroy()


def test_gap_statistic_and_tiers_agree_at_bench_size(ctx, switch):
    """The statistic that picks the kernel, and — at a bench-sized shard, where the oracle is too slow — the two kernels
    against each other: same float32 bits on a slab with resets and 2 % of the samples missing."""
    S, T = 20_000, 10_099
    c = ctx.counters_create(S, T)
    c.generate(9, step_ms=STEP)
    miss, seen = c.gap_stat()
    assert miss == 0 and S * T // 64 * 0.9 < seen <= S * T // 64 + 16
    c.generate(9, step_ms=STEP, reset_period=2000, missing_per_million=20_000)
    miss, seen = c.gap_stat()
    assert 0.018 < miss / seen < 0.022
    prm = Counters.params(STEP, 20)
    dst = ctx.series_create(S, c.rate_len(prm), synth.even_node_offsets(S // 8, 8))
    rows = {}
    for tier in ("1", "2", "0"):
        switch("IGPU_RATE_TIER", tier)
        c.rate(prm, dst)
        rows[tier] = np.concatenate([dst.download(0, 256), dst.download(S - 256, 256)]).view(np.uint32)
    assert np.array_equal(rows["1"], rows["2"]) and np.array_equal(rows["0"], rows["2"])
    assert np.isnan(rows["2"].view(np.float32)).mean() < 1e-6  # 2 % gaps leave (almost) no 20-sample range with < 2 samples
    dst.destroy()
    c.destroy()
    small = ctx.counters_create(3, 100)   # a small slab is looked at whole
    v = np.arange(300, dtype=np.float64).reshape(3, 100)
    v[1, 10:20] = np.nan
    small.upload(v)
    assert small.gap_stat() == (10, 288)  # 18 whole 16-sample lines
    small.destroy()


def test_argument_errors(ctx):
    with pytest.raises(native.IgpuError) as e:
        ctx.counters_create(1, 0x7FFFFFF0)  # positions would not fit 31 bits (and 16 GB per row)
    assert e.value.status == -6
    c = ctx.counters_create(4, 100)
    good = Counters.params(STEP, 20)
    dst = ctx.series_create(4, c.rate_len(good), synth.even_node_offsets(1, 4))
    wrong = ctx.series_create(4, 50, synth.even_node_offsets(1, 4))
    with pytest.raises(native.IgpuError) as e:
        c.rate(good, wrong)
    assert e.value.status == -1 and "destination slab" in str(e.value)
    for bad in (Counters.params(0, 20, first_eval=19), Counters.params(STEP, 0, first_eval=19),
                Counters.params(STEP, 20, eval_stride=0), Counters.params(STEP, 20, options=64)):
        with pytest.raises(native.IgpuError) as e:
            c.rate(bad, dst)
        assert e.value.status == -1
    c.rate(good, dst)
    for h in (dst, wrong, c):
        h.destroy()
