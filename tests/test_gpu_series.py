"""GPU parity (through the C ABI): power series — generator, window statistics, per-node quantile.
min / max / counts are bit-exact; mean / quantile / sums are within 1e-5 relative (north star)."""
import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import metrics, synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def close(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    return np.all(both_nan | (np.abs(a - b) <= rtol * np.maximum(np.abs(b), 1e-30)))


def same_or_nan(a, b):
    return np.all((a == b) | (np.isnan(a) & np.isnan(b)))


def test_device_generator_is_bit_identical_to_host(ctx):
    s = ctx.series_create(40, 5000, synth.even_node_offsets(10, 4))
    s.generate(synth.DEFAULT_SEED, global_first=123, nan_per_million=1500)
    dev = s.download(0, 40)
    host = synth.synth_series(40, 5000, global_first=123, nan_per_million=1500)
    assert np.array_equal(dev.view(np.uint32), host.view(np.uint32))
    s.destroy()


@pytest.mark.parametrize(
    "n_series,T,W",
    [
        (24, 5760, 60),    # C1 shape: 24 h of 15 s samples, fast path PW=4 L=15
        (7, 86400, 60),    # one day of 1 s samples
        (5, 5000, 60),     # partial tail window, T % W != 0
        (9, 4096, 20),     # PW=4 L=5, K>1 rounds per tile
        (3, 7200, 300),    # PW=4 L=75 -> staged kernel with a run-time chunk length, 2-stage ring
        (5, 14400, 1440),  # PW=32 L=45 (a day of minutes)
        (4, 9000, 180),    # PW=4 L=45, T % W != 0: chunks cut by the end of the row
        (3, 8148, 388),    # PW=4 L=97, the longest staged chunk
        (3, 8080, 404),    # PW=4 L=101 -> generic kernel
        (600, 2400, 300),  # more tiles than CTAs: the ring wraps (2 stages)
        (4, 3000, 64),     # power of two, T % W != 0: float4 units, PW=1 L=16
        (3, 4096, 128),    # float4 units PW=2 L=16
        (3, 5120, 256),    # float4 units PW=4 L=16
        (3, 10240, 1024),  # float4 units PW=16 L=16
        (3, 8192, 4096),   # float4 units PW=32 L=32 > 24 -> generic
        (3, 3072, 768),    # float4 units PW=32 L=6 (even, not a power of two)
        (5, 2000, 8),      # float4 units PW=1 L=2
        (5, 2000, 16),     # float4 units PW=1 L=4
        (5, 2000, 32),     # float4 units PW=1 L=8
        (2, 9000, 3000),   # 750 float4s = 2 x 375: no staged shape -> generic kernel
        (700, 1280, 64),   # more tiles than CTAs
        (6, 999, 7),       # T % 4 != 0 -> generic kernel
        (2, 50, 100),      # window longer than the series
        (3, 2048, 4),      # L=1
        (2, 12000, 992),   # PW=32 L=31, largest fast-path window
        (6, 8640, 30),     # windows that are not a multiple of 4 samples take the staged kernel too: 2 x 15 (5 m at a 10 s step)
        (6, 5000, 30),     # ... with the last window cut by the row end (20 of 30 samples)
        (5, 4000, 10),     # 2 x 5, K = 3 rounds per tile
        (5, 4000, 6),      # 2 x 3
        (5, 4500, 15),     # 1 x 15
        (5, 3500, 7),      # 1 x 7
        (4, 9000, 90),     # 2 x 45: run-time chunk length
        (4, 9000, 150),    # 2 x 75
        (3, 9004, 194),    # 2 x 97
        (3, 8080, 202),    # 2 x 101: past the longest staged chunk -> generic kernel
        (4, 3000, 2),      # 2 x 1, K = 16
        (4, 3000, 1),      # every sample its own window
        (5, 5761, 60),     # rows that are not a multiple of 4 samples (a query_range of N steps returns N + 1 points): the
        (4, 8641, 30),     # slab pads its rows to a 16-byte pitch with NaN, so these take the staged kernel as well
        (3, 1001, 7),
        (3, 4099, 64),
        (2, 86401, 1440),
    ],
)
def test_window_stats_match_oracle(ctx, oracle_lib, n_series, T, W):
    x = synth.synth_series(n_series, T, seed=W, nan_per_million=3000)
    if n_series > 2:
        x[1, : min(T, 3 * W)] = np.nan  # whole windows missing
    ps = metrics.PowerSeries.from_array(x, [0, n_series], ctx)
    mn, me, mx = ps.window_stats(W)
    omn, ome, omx = oracle_lib.window_stats(x, W)
    assert same_or_nan(mn, omn) and same_or_nan(mx, omx)
    assert close(me, ome)
    ps.close()


@pytest.mark.parametrize("W", [60, 300, 64, 512])
def test_window_stats_generic_equals_fast(ctx, switch, W):
    x = synth.synth_series(16, 8640 if W == 60 else 9000, seed=5)
    ps = metrics.PowerSeries.from_array(x, [0, 16], ctx)
    fast = ps.window_stats(W)
    switch("IGPU_WS_FORCE_GENERIC")
    gen = ps.window_stats(W)
    assert np.array_equal(fast[0], gen[0]) and np.array_equal(fast[2], gen[2])  # min / max: order-free, bit-equal
    assert close(fast[1], gen[1], 1e-6)  # mean: the two kernels add in different orders (float64 sums, float32 result)
    ps.close()


@pytest.mark.parametrize("T", [1, 2, 3, 5, 999, 5761])
def test_rows_of_any_length_round_trip_and_generate(ctx, oracle_lib, T):
    """Rows are stored at a pitch of a multiple of 4 samples; what the caller uploads, downloads and generates is unchanged."""
    S = 7
    x = synth.synth_series(S, T, seed=T, nan_per_million=50000)
    s = ctx.series_create(S, T, synth.even_node_offsets(1, S))
    s.upload(x)
    assert same_or_nan(s.download(0, S), x)
    s.upload(x[2:5] + 1.0, first=2)
    assert same_or_nan(s.download(2, 3), x[2:5] + 1.0) and same_or_nan(s.download(5, 2), x[5:])
    s.generate(11, global_first=3, nan_per_million=2000)
    want = oracle_lib.generate_series(S, T, 11, 3, 2000)
    assert same_or_nan(s.download(0, S), want)
    qv, st, fl = s.quantile_by_node(0.5)
    assert st["count"][0] == int((~np.isnan(want)).sum())   # the padding is never counted
    assert s.quantile_fleet(0.5) == oracle_lib.quantile_fleet(want, 0.5)
    s.destroy()


def node_offsets_ragged(n_series):
    # nodes with 0, 1, 2, 8, ... series
    cuts = [0, 0, 1, 3, 11, 12, n_series]
    return np.array(sorted(min(c, n_series) for c in cuts), np.uint32)


@pytest.mark.parametrize("q,mode", [(0.95, 0), (0.5, 0), (0.0, 0), (1.0, 0), (0.999, 0), (0.95, 1), (0.01, 1), (1.0, 1)])
def test_quantile_matches_oracle(ctx, oracle_lib, q, mode):
    x = synth.synth_series(24, 20000, seed=21, nan_per_million=2000)
    x[11, :] = np.nan  # a node whose only series is entirely missing
    off = node_offsets_ragged(24)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    qv, st, fleet = ps.quantile_by_node(q, mode)
    oq, ost, ofleet = oracle_lib.quantile_by_node(x, off, q, mode)
    assert close(qv, oq, 1e-12), (qv, oq)  # the select is exact; only the interpolation rounds
    assert np.array_equal(st["count"], ost["count"])
    assert same_or_nan(st["min"], ost["min"]) and same_or_nan(st["max"], ost["max"])
    assert close(st["sum"], ost["sum"])
    assert fleet[K.F_COUNT] == ofleet[K.F_COUNT] and fleet[K.F_NODES] == ofleet[K.F_NODES]
    assert fleet[K.F_MIN] == ofleet[K.F_MIN] and fleet[K.F_MAX] == ofleet[K.F_MAX] and close(fleet[K.F_SUM], ofleet[K.F_SUM])
    ps.close()


def test_quantile_heavy_ties_and_constants(ctx, oracle_lib):
    T = 30000
    x = synth.synth_series(8, T, seed=3)
    x[0, :] = 150.0                                   # constant series (idle / clipped at TDP)
    x[1, :] = np.where(np.arange(T) % 10 < 9, 250.0, x[1])  # 90 % ties exactly at the cap
    x[2, :] = np.round(x[2])                          # few distinct values
    x[3, :] = np.sort(x[3])                           # monotone ramp
    x[4, :] = np.where(np.arange(T) % 7 == 0, 0.0, 120.0)   # periodic with the sampling stride's enemies
    x[5, :] = -x[5]                                   # negative values (key order across the sign)
    x[6, :5] = [np.inf, -np.inf, 0.0, -0.0, 1e-38]
    off = np.arange(9, dtype=np.uint32)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    for q, mode in [(0.95, 0), (0.5, 0), (0.9, 0), (0.95, 1), (0.1, 0)]:
        qv, st, _ = ps.quantile_by_node(q, mode)
        oq, ost, _ = oracle_lib.quantile_by_node(x, off, q, mode)
        assert close(qv, oq, 1e-12), (q, mode, qv, oq)
        assert np.array_equal(st["count"], ost["count"])
    ps.close()


def test_quantile_tiny_and_large_nodes(ctx, oracle_lib):
    # nodes from 1 sample to ~1.2e6 samples (several scan items, sample stride > 1)
    T = 1
    x = synth.synth_series(300, T, seed=8)
    off = np.array([0, 1, 3, 10, 300], np.uint32)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    qv, _, _ = ps.quantile_by_node(0.95, 0)
    assert close(qv, oracle_lib.quantile_by_node(x, off, 0.95, 0)[0], 1e-12)
    ps.close()
    x = synth.synth_series(10, 120000, seed=9, nan_per_million=500)
    off = np.array([0, 10], np.uint32)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    for q in (0.95, 0.5, 0.05):
        qv, st, _ = ps.quantile_by_node(q, 0)
        oq, ost, _ = oracle_lib.quantile_by_node(x, off, q, 0)
        assert close(qv, oq, 1e-12) and st["count"][0] == ost["count"][0]
    ps.close()


def test_series_c1_shape_end_to_end(ctx, oracle_lib):
    """BASELINE config 1 series: 50 nodes x 4 chips x 5,760 samples (24 h of 15 s samples)."""
    x = synth.synth_series(200, 5760)
    off = synth.even_node_offsets(50, 4)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    mn, me, mx = ps.window_stats(20)  # 5-minute windows
    omn, ome, omx = oracle_lib.window_stats(x, 20)
    assert np.array_equal(mn, omn) and np.array_equal(mx, omx) and close(me, ome)
    qv, st, fleet = ps.quantile_by_node(0.95)
    oq, ost, ofleet = oracle_lib.quantile_by_node(x, off, 0.95)
    assert close(qv, oq, 1e-12) and close(st["sum"], ost["sum"]) and close(fleet[K.F_SUM], ofleet[K.F_SUM])
    ps.close()


def test_full_rate_shape_properties(ctx, oracle_lib):
    """A slab far larger than L2 (2.2e8 samples) generated on the device: size-independent
    properties plus exact parity on rows sampled back to the host."""
    n_nodes, chips, T = 320, 8, 86400
    s = ctx.series_create(n_nodes * chips, T, synth.even_node_offsets(n_nodes, chips))
    s.generate(synth.DEFAULT_SEED, 0, 100)
    s.window_stats(60, fetch=False)
    qv, st, fleet = s.quantile_by_node(0.95)
    # property: min <= mean <= max per window, on sampled rows; exact parity against the oracle there
    for first in (0, 1283, n_nodes * chips - 8):
        rows = s.download(first, 8)
        mn, me, mx = s.read_window_stats(60, first, 8)
        omn, ome, omx = oracle_lib.window_stats(rows, 60)
        assert same_or_nan(mn, omn) and same_or_nan(mx, omx) and close(me, ome)
        assert np.all((mn <= me + 1e-3) & (me <= mx + 1e-3) | np.isnan(me))
    for node in (0, 77, n_nodes - 1):
        seg = s.download(node * chips, chips)
        oq, ost, _ = oracle_lib.quantile_by_node(seg, np.array([0, chips], np.uint32), 0.95)
        assert close(qv[node], oq[0], 1e-12) and st["count"][node] == ost["count"][0]
        assert st["min"][node] == ost["min"][0] and st["max"][node] == ost["max"][0] and close(st["sum"][node], ost["sum"][0])
    # property: quantile within [min, max]; fleet vector = reduction of node stats
    assert np.all((qv >= st["min"]) & (qv <= st["max"]))
    assert fleet[K.F_COUNT] == st["count"].sum() and fleet[K.F_MIN] == st["min"].min() and fleet[K.F_MAX] == st["max"].max()
    assert close(fleet[K.F_SUM], st["sum"].sum())
    s.destroy()


def test_quantile_key_range_not_a_multiple_of_the_bin_width(ctx, oracle_lib):
    """Regression: the histogram refinement picks shift so that the LARGEST bin index (width-1)>>shift
    stays below 2048; a key range of 2048*2^s + x (x >= 1) used to index one bin past the table."""
    T = 6000
    base = np.float32(150.0)
    lo_key = np.array([base], np.float32).view(np.uint32)[0]
    for extra in (1, 2, 4097, 8191):
        hi = np.array([lo_key + np.uint32(2048 * 8192 + extra - 1)], np.uint32).view(np.float32)[0]
        x = np.linspace(base, hi, T, dtype=np.float32)[None, :].copy()
        x[0, 0], x[0, -1] = base, hi
        ps = metrics.PowerSeries.from_array(x, [0, 1], ctx)
        for q in (0.95, 0.5):
            qv, _, _ = ps.quantile_by_node(q)
            assert close(qv, oracle_lib.quantile_by_node(x, np.array([0, 1], np.uint32), q)[0], 1e-12)
        ps.close()


# ---- igpu_series_rollup: window statistics + quantile in one read ------------------------------------


@pytest.mark.parametrize(
    "n_series,T,W,q,mode",
    [
        (24, 5760, 60, 0.95, 0),    # fused kernel, L = 15
        (24, 5000, 60, 0.95, 0),    # tail window cut by the row end
        (9, 4096, 20, 0.5, 0),      # L = 5, K = 3 rounds per tile
        (12, 2048, 4, 0.95, 1),     # L = 1, K = 16
        (8, 12000, 992, 0.95, 0),   # L = 31 > 15: falls back to the two separate passes
        (6, 999, 7, 0.9, 0),        # T % 4 != 0: falls back
        (16, 86400, 60, 0.999, 0),  # a day of 1 s samples
        (12, 8640, 30, 0.95, 0),    # W % 4 != 0: 2 x 15, the chunk-fused single-read kernel (5 m at a 10 s step)
        (12, 5000, 30, 0.5, 0),     # ... last window cut
        (9, 4000, 10, 0.95, 0),     # 2 x 5
        (9, 4500, 15, 0.95, 1),     # 1 x 15
        (9, 3500, 7, 0.9, 0),       # 1 x 7
        (8, 9000, 90, 0.95, 0),     # 2 x 45: chunk too long for the fused kernel -> staged window kernel + scan
        (12, 5761, 60, 0.95, 0),    # N + 1 points per row: padded pitch, thread-per-window kernel
        (7, 86401, 300, 0.95, 0),   # ... with sub-windows
        (9, 4321, 30, 0.5, 0),      # ... chunk-fused kernel; 4321 = 144 x 30 + 1: the last window holds one sample
        (9, 2999, 7, 0.9, 1),
    ],
)
def test_rollup_matches_oracle_and_separate_calls(ctx, oracle_lib, n_series, T, W, q, mode):
    x = synth.synth_series(n_series, T, seed=W + n_series, nan_per_million=2500)
    x[1, : min(T, 2 * W)] = np.nan
    off = node_offsets_ragged(n_series)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    (mn, me, mx), qv, st, fleet = ps.rollup(W, q, mode)
    omn, ome, omx = oracle_lib.window_stats(x, W)
    oq, ost, ofleet = oracle_lib.quantile_by_node(x, off, q, mode)
    assert same_or_nan(mn, omn) and same_or_nan(mx, omx) and close(me, ome)
    assert close(qv, oq, 1e-12), (qv, oq)
    assert np.array_equal(st["count"], ost["count"]) and same_or_nan(st["min"], ost["min"]) and same_or_nan(st["max"], ost["max"])
    assert close(st["sum"], ost["sum"]) and close(fleet[K.F_SUM], ofleet[K.F_SUM]) and fleet[K.F_COUNT] == ofleet[K.F_COUNT]
    # and bit-identical window planes / quantiles to the two separate entry points
    smn, sme, smx = ps.window_stats(W)
    sq, sst, _ = ps.quantile_by_node(q, mode)
    assert same_or_nan(mn, smn) and same_or_nan(me, sme) and same_or_nan(mx, smx)
    assert same_or_nan(qv, sq) and np.array_equal(st["count"], sst["count"])
    ps.close()


def test_rollup_ties_constants_and_many_nodes(ctx, oracle_lib):
    T = 28800
    x = synth.synth_series(64, T, seed=31)
    x[0, :] = 150.0
    x[1, :] = np.where(np.arange(T) % 10 < 9, 250.0, x[1])
    x[2, :] = np.round(x[2])
    x[3, :] = np.sort(x[3])
    x[5, :] = -x[5]
    off = np.arange(0, 65, 1, dtype=np.uint32)  # one series per node: node changes at every row
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    for q, mode in ((0.95, 0), (0.5, 0), (0.95, 1)):
        (mn, me, mx), qv, st, _ = ps.rollup(60, q, mode)
        oq, ost, _ = oracle_lib.quantile_by_node(x, off, q, mode)
        assert close(qv, oq, 1e-12), (q, mode)
        assert np.array_equal(st["count"], ost["count"]) and close(st["sum"], ost["sum"])
    omn, ome, omx = oracle_lib.window_stats(x, 60)
    assert np.array_equal(mn, omn) and np.array_equal(mx, omx) and close(me, ome)
    ps.close()


def test_rollup_full_rate_shape_properties(ctx, oracle_lib):
    """2.2e8 device-generated samples through the fused kernel; exact parity on rows copied back."""
    n_nodes, chips, T = 320, 8, 86400
    s = ctx.series_create(n_nodes * chips, T, synth.even_node_offsets(n_nodes, chips))
    s.generate(synth.DEFAULT_SEED + 1, 0, 100)
    _, qv, st, fleet = s.rollup(60, 0.95, 0, fetch_windows=False)
    for first in (0, 1283, n_nodes * chips - 8):
        rows = s.download(first, 8)
        mn, me, mx = s.read_window_stats(60, first, 8)
        omn, ome, omx = oracle_lib.window_stats(rows, 60)
        assert same_or_nan(mn, omn) and same_or_nan(mx, omx) and close(me, ome)
    for node in (0, 77, 160, n_nodes - 1):
        seg = s.download(node * chips, chips)
        oq, ost, _ = oracle_lib.quantile_by_node(seg, np.array([0, chips], np.uint32), 0.95)
        assert close(qv[node], oq[0], 1e-12) and st["count"][node] == ost["count"][0]
        assert st["min"][node] == ost["min"][0] and st["max"][node] == ost["max"][0] and close(st["sum"][node], ost["sum"][0])
    assert fleet[K.F_COUNT] == st["count"].sum() and close(fleet[K.F_SUM], st["sum"].sum())
    s.destroy()


# ---- BASELINE configs at full per-GPU size (device-generated; parity through rows copied back) -------


def _check_sampled(s, oracle_lib, qv, st, chips, nodes, rows, window=None):
    for node in nodes:
        seg = s.download(node * chips, chips)
        oq, ost, _ = oracle_lib.quantile_by_node(seg, np.array([0, chips], np.uint32), 0.95, threads=8)
        assert close(qv[node], oq[0], 1e-12), (node, qv[node], oq[0])
        assert st["count"][node] == ost["count"][0] and st["min"][node] == ost["min"][0] and st["max"][node] == ost["max"][0]
        assert close(st["sum"][node], ost["sum"][0])
    if window:
        for first in rows:
            block = s.download(first, 4)
            mn, me, mx = s.read_window_stats(window, first, 4)
            omn, ome, omx = oracle_lib.window_stats(block, window)
            assert same_or_nan(mn, omn) and same_or_nan(mx, omx) and close(me, ome)


def test_config_c4_full_size(ctx, oracle_lib):
    """BASELINE config 4 on one GPU: 10k nodes x 8 chips x 86,400 one-second samples = 6.912e9 f32 (27.6 GB),
    W = 60 downsampling + per-node p95 in one read; exact parity on sampled nodes and rows."""
    import torch

    if torch.cuda.mem_get_info()[0] < 40e9:
        pytest.skip("needs ~33 GB of free device memory")
    n_nodes, chips, T = 10_000, 8, 86_400
    s = ctx.series_create(n_nodes * chips, T, synth.even_node_offsets(n_nodes, chips))
    s.generate(synth.DEFAULT_SEED, 0, 50)
    _, qv, st, fleet = s.rollup(60, 0.95, 0, fetch_windows=False)
    _check_sampled(s, oracle_lib, qv, st, chips, nodes=(0, 6213, 6214, 9_999), rows=(0, 49_700, 79_996), window=60)
    assert fleet[K.F_COUNT] == st["count"].sum() and fleet[K.F_NODES] == n_nodes
    assert np.all((qv >= st["min"]) & (qv <= st["max"]))
    # the two separate entry points give the same numbers
    sq, sst, _ = s.quantile_by_node(0.95)
    assert np.array_equal(sq, qv) and np.array_equal(sst["count"], st["count"])
    s.destroy()


def test_config_c5_per_gpu_shard(ctx, oracle_lib):
    """BASELINE config 5, one rank's shard of the 8-GPU run: 1,250 nodes x 8 chips x 1.25e6 samples =
    1.25e10 f32 (50 GB), per-node p95 over 1e7-sample segments."""
    import torch

    if torch.cuda.mem_get_info()[0] < 70e9:
        pytest.skip("needs ~58 GB of free device memory")
    n_nodes, chips, T = 1_250, 8, 1_250_000
    s = ctx.series_create(n_nodes * chips, T, synth.even_node_offsets(n_nodes, chips))
    s.generate(synth.DEFAULT_SEED + 5, 3 * n_nodes * chips, 20)  # rank 3's slice of the fleet
    qv, st, fleet = s.quantile_by_node(0.95)
    _check_sampled(s, oracle_lib, qv, st, chips, nodes=(0, 429, 430, 1_249), rows=())
    assert fleet[K.F_COUNT] == st["count"].sum() and close(fleet[K.F_SUM], st["sum"].sum())
    assert np.all((qv >= st["min"]) & (qv <= st["max"]))
    s.destroy()


def test_quantile_exact_fallback_path(ctx, oracle_lib, switch):
    """Every node forced through the exact whole-segment select (the path taken when a bracket misses or
    a candidate list overflows): same answers as the bracketed path and the oracle."""
    x = synth.synth_series(24, 20000, seed=77, nan_per_million=3000)
    x[3, :] = 99.5
    x[5, :] = -x[5]
    off = node_offsets_ragged(24)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    switch("IGPU_Q_FORCE_FALLBACK")
    for q, mode in ((0.95, 0), (0.5, 0), (0.0, 0), (1.0, 0), (0.95, 1)):
        qv, st, _ = ps.quantile_by_node(q, mode)
        oq, ost, _ = oracle_lib.quantile_by_node(x, off, q, mode)
        assert close(qv, oq, 1e-12), (q, mode, qv, oq)
        assert np.array_equal(st["count"], ost["count"])
    (mn, me, mx), qv, st, _ = ps.rollup(60, 0.95, 0)
    assert close(qv, oracle_lib.quantile_by_node(x, off, 0.95, 0)[0], 1e-12)
    ps.close()


@pytest.mark.parametrize("n_chunks,lanes", [(1, 2), (2, 2), (5, 2), (4, 3)])
def test_streamed_rollup_equals_resident_rollup(ctx, n_chunks, lanes):
    """Chunks ping-ponging between contexts (upload of chunk i+1 overlapping rollup + download of chunk i) give the
    same per-node and per-window results as one resident slab; the merged fleet vector matches too."""
    from headlamp_intel_gpu_plugin_b200 import native

    nodes, chips, T, W = 6, 2, 2880, 60
    rows = nodes * chips
    x = synth.synth_series(n_chunks * rows, T, seed=31, nan_per_million=2000)
    whole = metrics.PowerSeries.from_array(x, synth.even_node_offsets(n_chunks * nodes, chips), ctx)
    (wmn, wme, wmx), q, st, fl = whole.rollup(W, 0.95)
    pinned = [native.PinnedBuffer(rows * T * 4) for _ in range(n_chunks)]
    chunks = []
    for i, pb in enumerate(pinned):
        v = pb.view(np.float32, (rows, T))
        v[...] = x[i * rows:(i + 1) * rows]
        chunks.append(v)
    nw = T // W
    outs = tuple(np.empty((n_chunks * rows, nw), np.float32) for _ in range(3))
    sr = metrics.StreamedSeriesRollup(rows, T, synth.even_node_offsets(nodes, chips), n_lanes=lanes)
    for _ in range(2):  # a second run reuses the lanes
        q2, st2, fl2 = sr.run(chunks, W, 0.95, out_windows=outs)
        assert np.array_equal(q2, q)
        assert np.array_equal(st2["count"], st["count"]) and same_or_nan(st2["min"], st["min"]) and same_or_nan(st2["max"], st["max"])
        assert close(st2["sum"], st["sum"], 1e-12)
        assert same_or_nan(outs[0], wmn) and same_or_nan(outs[2], wmx) and same_or_nan(outs[1], wme)
        assert close(fl2[:5], fl[:5], 1e-12)
    sr.close()
    for pb in pinned:
        pb.free()
    whole.close()


def test_range_view_numbers(ctx, oracle_lib):
    """Peak / cluster average / per-node numbers of the mocked-up range view, against the oracle module and against plain
    numpy on the same slab."""
    nodes, chips, T = 5, 3, 1800
    x = synth.synth_series(nodes * chips, T, seed=4, nan_per_million=3000)
    x[6:9] = np.nan  # node 2 never reported
    ps = metrics.PowerSeries.from_array(x, synth.even_node_offsets(nodes, chips), ctx)
    names = [f"gpu-node-{i:02d}" for i in range(nodes)]
    v = ps.range_view(60, 0.95, node_names=names, tdp_watts=[300.0] * nodes)
    per_node = x.reshape(nodes, -1)
    assert v["nodesReporting"] == 4 and v["peakNode"] == names[int(np.nanargmax(np.nanmax(np.where(np.isnan(per_node), -np.inf, per_node), axis=1)))]
    assert v["peakWatts"] == float(np.nanmax(x))
    assert abs(v["avgWattsCluster"] - np.nanmean(x.astype(np.float64))) <= 1e-9 * np.nanmean(x.astype(np.float64))
    assert np.isnan(v["nodes"][2]["avgWatts"]) and np.isnan(v["nodes"][2]["p95Watts"])
    for i in (0, 1, 3, 4):
        assert abs(v["nodes"][i]["avgWatts"] - np.nanmean(per_node[i].astype(np.float64))) < 1e-9 * 300
        assert v["nodes"][i]["peakWatts"] == float(np.nanmax(per_node[i]))
    want_util = np.mean([np.nanmean(per_node[i].astype(np.float64)) / 300.0 for i in (0, 1, 3, 4)]) * 100
    assert abs(v["avgTdpUtilizationPct"] - want_util) < 1e-9 * 100
    # the same numbers from the oracle's definitions (quantile_over_time, node statistics, fleet vector)
    oq, ost, ofl = oracle_lib.quantile_by_node(x, synth.even_node_offsets(nodes, chips), 0.95)
    for i in (0, 1, 3, 4):
        assert v["nodes"][i]["p95Watts"] == oq[i] and v["nodes"][i]["peakWatts"] == float(ost["max"][i])
        assert abs(v["nodes"][i]["avgWatts"] - ost["sum"][i] / ost["count"][i]) <= 1e-12 * 300
    assert v["peakWatts"] == ofl[K.F_MAX] and v["nodesReporting"] == int(ofl[K.F_NODES])
    assert abs(v["avgWattsCluster"] - ofl[K.F_SUM] / ofl[K.F_COUNT]) <= 1e-9 * 300
    fv = ps.range_view(60, 0.95, fleet_views=True)
    assert fv["p95WattsCluster"] == oracle_lib.quantile_fleet(x, 0.95)[0]
    _, ocurve = oracle_lib.window_rollup(x, synth.even_node_offsets(nodes, chips), 60)
    assert same_or_nan(fv["curve"]["peakWatts"], ocurve["max"]) and close(fv["curve"]["avgWatts"], ocurve["mean"])
    assert np.array_equal(fv["curve"]["samples"], ocurve["count"])
    ps.close()



# ---- fleet-level views (north star: "per-node and per-fleet windowed min/mean/max/p95") ---------------------------


@pytest.mark.parametrize("q,mode", [(0.95, 0), (0.5, 0), (0.0, 0), (1.0, 0), (0.9999, 0), (0.95, 1), (0.001, 1), (1.0, 1)])
def test_fleet_quantile_matches_oracle(ctx, oracle_lib, q, mode):
    x = synth.synth_series(37, 2999, seed=8, nan_per_million=4000)  # S * T odd: scalar tail after the 128-bit loads
    x[5] = np.nan
    x[11, :100] = -3.5  # negative keys sort below the positive ones
    s = ctx.series_create(37, 2999, node_offsets_ragged(37))
    s.upload(x)
    got, cnt = s.quantile_fleet(q, mode)
    want, wcnt = oracle_lib.quantile_fleet(x, q, mode)
    assert cnt == wcnt == int((~np.isnan(x)).sum())
    assert got == want  # float64 of exact order statistics and the same interpolation: identical bits
    s.destroy()


@pytest.mark.parametrize("q,mode", [(0.95, 0), (0.5, 0), (0.0, 0), (1.0, 0), (0.9999, 0), (0.0001, 0), (0.95, 1), (0.001, 1), (1.0, 1)])
@pytest.mark.parametrize("S,T", [(37, 2999), (64, 4096), (8, 86400)])
def test_fleet_quantile_bracketed_by_a_sample(ctx, oracle_lib, switch, S, T, q, mode):
    """The one-read path (sample -> bracket -> gather -> select), forced onto small slabs: same bits as the oracle, and the
    same as the three exact passes."""
    x = synth.synth_series(S, T, seed=S + 5, nan_per_million=4000)
    x[3] = np.nan
    x[5, :100] = -3.5
    s = ctx.series_create(S, T, synth.even_node_offsets(1, S))
    s.upload(x)
    switch("IGPU_FLEET_QUANTILE_MIN_SAMPLE", "1")
    got, cnt = s.quantile_fleet(q, mode)
    assert s.fleet_quantile_reads() == 1
    want, wcnt = oracle_lib.quantile_fleet(x, q, mode)
    assert cnt == wcnt and got == want
    switch("IGPU_FLEET_QUANTILE_EXACT")
    assert s.quantile_fleet(q, mode) == (got, cnt) and s.fleet_quantile_reads() == 3
    s.destroy()


def test_fleet_quantile_bracket_with_zeros_of_both_signs(ctx, oracle_lib, switch):
    """The one-read pass compares values, the selects compare order-preserving keys (-0.0 below +0.0): a fleet whose quantiles
    fall among zeros of both signs must still come out exact, whichever zero the sample's bracket ends on."""
    rng = np.random.default_rng(5)
    S, T = 16, 8192
    x = rng.choice(np.array([-2.0, -1.0, -0.0, 0.0, 0.0, -0.0, 1.0, 2.0], np.float32), size=(S, T))
    x[3, ::7] = np.nan
    s = ctx.series_create(S, T, synth.even_node_offsets(2, 8))
    s.upload(x)
    switch("IGPU_FLEET_QUANTILE_MIN_SAMPLE", "1")
    for q in (0.3, 0.5, 0.6, 0.26, 0.74, 0.0, 1.0):
        got, cnt = s.quantile_fleet(q)
        want, wcnt = oracle_lib.quantile_fleet(x, q)
        assert s.fleet_quantile_reads() == 1 and cnt == wcnt and got == want, (q, got, want)
    s.destroy()


def test_fleet_quantile_bracket_that_misses_falls_back(ctx, oracle_lib, switch):
    """A slab whose sampled lines say nothing about the rest: every 128-byte line the sample does NOT touch is lifted by
    1000 W (the sample's line choice is restated here).  The counts of the bracketed pass show that the wanted ranks are not
    inside the bracket, and the three exact passes take over: 4 reads, still the oracle's bits."""
    S, T = 16, 65536
    x = synth.synth_series(S, T, seed=2)
    s = ctx.series_create(S, T, synth.even_node_offsets(2, 8))
    s.upload(x)
    switch("IGPU_FLEET_QUANTILE_MIN_SAMPLE", "1")
    p95, _ = s.quantile_fleet(0.95)
    assert s.fleet_quantile_reads() == 1 and p95 == oracle_lib.quantile_fleet(x, 0.95)[0]

    def mix(v):  # fleet_series.cu fq_mix
        v = v.astype(np.uint64)
        v ^= v >> 16
        v = (v * 0x7FEB352D) & 0xFFFFFFFF
        v ^= v >> 15
        v = (v * 0x846CA68B) & 0xFFFFFFFF
        return v ^ (v >> 16)

    lines = x.reshape(-1, 32)
    pick = np.arange(lines.shape[0] // 64, dtype=np.uint64)
    sampled = pick * 64 + mix(pick) % 64
    lift = np.ones(lines.shape[0], bool)
    lift[sampled.astype(np.int64)] = False
    lines[lift] += 1000.0
    s.upload(x)
    got, cnt = s.quantile_fleet(0.95)
    assert s.fleet_quantile_reads() == 4
    assert (got, cnt) == oracle_lib.quantile_fleet(x, 0.95) and got > 1000.0
    s.destroy()


def test_fleet_quantile_ties_constants_and_empty(ctx, oracle_lib):
    x = np.full((6, 400), 250.0, np.float32)
    x[2, ::3] = 100.0
    x[4] = np.nan
    s = ctx.series_create(6, 400, synth.even_node_offsets(3, 2))
    s.upload(x)
    for q in (0.0, 0.05, 0.06, 0.5, 1.0):
        assert s.quantile_fleet(q)[0] == oracle_lib.quantile_fleet(x, q)[0]
    s.upload(np.full((6, 400), np.nan, np.float32))
    v, c = s.quantile_fleet(0.95)
    assert np.isnan(v) and c == 0
    s.destroy()


@pytest.mark.parametrize("n_nodes,chips,T,W", [(7, 4, 2880, 60), (5, 3, 1801, 60), (4, 8, 1440, 300), (3, 2, 500, 7), (6, 1, 640, 64)])
def test_node_and_fleet_window_rollup_match_oracle(ctx, oracle_lib, n_nodes, chips, T, W):
    x = synth.synth_series(n_nodes * chips, T, seed=10, nan_per_million=20000)
    x[chips:2 * chips, :W * 2] = np.nan  # node 1 has two empty windows
    x[:, W * 3:W * 4] = np.nan           # one window is empty fleet-wide
    off = synth.even_node_offsets(n_nodes, chips)
    s = ctx.series_create(n_nodes * chips, T, off)
    s.upload(x)
    node, fleet = s.window_rollup(W)
    onode, ofleet = oracle_lib.window_rollup(x, off, W)
    for got, want in ((node, onode), (fleet, ofleet)):
        assert np.array_equal(got["count"], want["count"])
        assert same_or_nan(got["min"], want["min"]) and same_or_nan(got["max"], want["max"])
        assert close(got["mean"], want["mean"])
    assert np.isnan(fleet["mean"][3]) and fleet["count"][3] == 0 and np.isnan(node["min"][1, 0])
    # the same planes follow from the per-series window statistics where no sample is missing
    s.destroy()


@pytest.mark.parametrize("n_series,T,W", [(28, 2880, 60), (32, 5000, 300), (13, 86400, 1440), (16, 5000, 64), (12, 20480, 1024), (9, 7204, 240),
                                          (14, 5761, 60), (13, 86401, 300)])  # the last two: N + 1 points per row (padded pitch)
def test_window_rollup_without_a_second_read(ctx, oracle_lib, switch, n_series, T, W):
    """After keep_window_counts() a rollup of the same window leaves per-series planes + valid-sample counts behind, and the
    node / fleet planes are folded from those: same counts / min / max as the pass over the slab, mean within float32
    rounding; any write into the slab, another window, or a shape the single-read kernel does not take brings the read back."""
    x = synth.synth_series(n_series, T, seed=W + n_series, nan_per_million=20000)
    x[1, : min(T, 2 * W)] = np.nan     # a chip with two empty windows
    x[:, 3 * W:4 * W] = np.nan         # a window that is empty fleet-wide
    off = node_offsets_ragged(n_series)
    s = ctx.series_create(n_series, T, off)
    s.upload(x)
    slab_node, slab_fleet = s.window_rollup(W)
    assert s.last_window_rollup_slab_reads == 1
    s.keep_window_counts(True)
    s.rollup(W, 0.95, 0, fetch_windows=False)
    node, fleet = s.window_rollup(W)
    assert s.last_window_rollup_slab_reads == 0
    onode, ofleet = oracle_lib.window_rollup(x, off, W)
    for got, want, slab in ((node, onode, slab_node), (fleet, ofleet, slab_fleet)):
        assert np.array_equal(got["count"], want["count"]) and np.array_equal(got["count"], slab["count"])
        assert same_or_nan(got["min"], want["min"]) and same_or_nan(got["max"], want["max"])
        assert close(got["mean"], want["mean"]) and close(got["mean"], slab["mean"], rtol=3e-7)
    assert np.isnan(fleet["mean"][3]) and fleet["count"][3] == 0
    # what brings the read back
    s.window_rollup(W // 2 if W % 2 == 0 else W + 1)
    assert s.last_window_rollup_slab_reads == 1
    s.window_rollup(W)
    assert s.last_window_rollup_slab_reads == 0   # (the kept planes were not touched by the call above)
    switch("IGPU_WINDOW_ROLLUP_FROM_SLAB")
    s.window_rollup(W)
    assert s.last_window_rollup_slab_reads == 1
    switch("IGPU_WINDOW_ROLLUP_FROM_SLAB", "0")
    x2 = x.copy()
    x2[0, :W] = 1000.0
    s.upload(x2)
    node2, _ = s.window_rollup(W)
    assert s.last_window_rollup_slab_reads == 1 and node2["max"][1, 0] == 1000.0   # series 0 is node 1 (node 0 has no series)
    s.rollup(W, 0.95, 0, fetch_windows=False)
    node3, _ = s.window_rollup(W)
    assert s.last_window_rollup_slab_reads == 0 and same_or_nan(node3["max"], node2["max"]) and np.array_equal(node3["count"], node2["count"])
    s.keep_window_counts(False)
    s.window_rollup(W)
    assert s.last_window_rollup_slab_reads == 1
    s.destroy()


def test_window_rollup_reads_the_slab_for_shapes_the_single_read_kernel_declines(ctx, oracle_lib):
    x = synth.synth_series(6, 500, seed=3, nan_per_million=10000)
    off = synth.even_node_offsets(3, 2)
    s = ctx.series_create(6, 500, off)
    s.upload(x)
    s.keep_window_counts(True)
    s.rollup(7, 0.95, 0, fetch_windows=False)   # W = 7: the chunk-fused kernel, which keeps no counts
    node, fleet = s.window_rollup(7)
    assert s.last_window_rollup_slab_reads == 1
    onode, ofleet = oracle_lib.window_rollup(x, off, 7)
    assert np.array_equal(node["count"], onode["count"]) and close(fleet["mean"], ofleet["mean"])
    s.destroy()


def test_fleet_views_on_device_generated_slab(ctx, oracle_lib):
    """A slab too large to sort on the host in a test: properties that do not depend on the size."""
    S, T = 4000, 86_400  # 345.6 M samples
    s = ctx.series_create(S, T, synth.even_node_offsets(S // 8, 8))
    s.generate(synth.DEFAULT_SEED)
    p95, cnt = s.quantile_fleet(0.95)
    qv, st, fl = s.quantile_by_node(0.95)
    assert cnt == S * T == int(fl[K.F_COUNT])
    assert s.quantile_fleet(0.0)[0] == fl[K.F_MIN] and s.quantile_fleet(1.0)[0] == fl[K.F_MAX]
    assert qv.min() <= p95 <= qv.max()  # the fleet's p95 lies between the smallest and the largest per-node p95
    assert s.quantile_fleet(0.5)[0] <= p95
    _, fleet = s.window_rollup(60, nodes=False)
    assert int(fleet["count"].sum()) == S * T and float(fleet["min"].min()) == fl[K.F_MIN] and float(fleet["max"].max()) == fl[K.F_MAX]
    mean = float((fleet["mean"].astype(np.float64) * fleet["count"]).sum() / (S * T))
    assert abs(mean - fl[K.F_SUM] / fl[K.F_COUNT]) <= 1e-6 * mean
    s.destroy()


@pytest.mark.parametrize(
    "n_series,T,W,q",
    [
        (16, 86400, 300, 0.95),   # the 5 m window at 1 s (metrics.ts:106): 5 sub-windows of 60, 60 of 64 threads busy
        (12, 5000, 300, 0.5),     # last window cut by the row end inside a sub-window (200 of 300 samples)
        (10, 86400, 1440, 0.95),  # 24 sub-windows of 60
        (9, 8640, 100, 0.9),      # 5 sub-windows of 20
        (8, 12000, 3000, 0.95),   # 50 sub-windows of 60: a shape the staged window kernel does not take at all
        (11, 7200, 120, 0.99),    # 2 sub-windows of 60
        (7, 7204, 240, 0.95),     # row length not a multiple of the tile
        (9, 86400, 64, 0.95),     # a power-of-two window: 16 float4s per thread, every thread starts at its own float4
        (8, 5000, 64, 0.5),       # ... with a last window of 8 samples (read in order)
        (8, 5000, 256, 0.95),     # 4 sub-windows of 64; the last window holds 2 whole sub-windows and 8 samples
        (6, 20480, 1024, 0.9),    # 16 sub-windows of 64
        (6, 9600, 192, 0.95),     # 3 sub-windows of 64, 63 of 64 threads busy
    ],
)
def test_rollup_single_read_for_windows_made_of_sub_windows(ctx, oracle_lib, switch, n_series, T, W, q):
    """Windows of 4 * NV * k samples take the thread-per-sub-window kernel: one read of the slab for window statistics and
    quantiles; results equal the oracle's and the two separate passes'."""
    x = synth.synth_series(n_series, T, seed=W + n_series, nan_per_million=3000)
    x[1, : min(T, 2 * W)] = np.nan
    x[2, W : W + 61] = np.nan  # a sub-window without a sample inside a window that has some
    off = node_offsets_ragged(n_series)
    ps = metrics.PowerSeries.from_array(x, off, ctx)
    launches0 = ctx.timing()["kernel_launches"]
    (mn, me, mx), qv, st, fleet = ps.rollup(W, q, 0)
    assert ctx.timing()["kernels"]["series_fused"] > 0 and ctx.timing()["kernel_launches"] - launches0 == 5  # sample, fused, finish, fallback, fleet: no separate window pass
    omn, ome, omx = oracle_lib.window_stats(x, W)
    oq, ost, ofleet = oracle_lib.quantile_by_node(x, off, q, 0)
    assert same_or_nan(mn, omn) and same_or_nan(mx, omx) and close(me, ome)
    assert close(qv, oq, 1e-12), (qv, oq)
    assert np.array_equal(st["count"], ost["count"]) and same_or_nan(st["min"], ost["min"]) and same_or_nan(st["max"], ost["max"])
    assert close(st["sum"], ost["sum"]) and close(fleet[K.F_SUM], ofleet[K.F_SUM]) and fleet[K.F_COUNT] == ofleet[K.F_COUNT]
    switch("IGPU_NO_WIDE")  # the two separate passes (or the chunk-per-thread fused kernel)
    (smn, sme, smx), sq, sst, _ = ps.rollup(W, q, 0)
    assert same_or_nan(mn, smn) and same_or_nan(mx, smx) and close(me, sme) and same_or_nan(qv, sq)
    ps.close()
