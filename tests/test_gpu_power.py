"""GPU parity (through the C ABI): the instant power join of metrics.ts and the fleet sums."""
import pytest

from headlamp_intel_gpu_plugin_b200 import metrics, synth
from oracle import ref_objects as ro

from .parity import nan_eq

pytestmark = pytest.mark.gpu


def mutate(ch, en, mx, un):
    en.append({"metric": {"chip": ch[0]["metric"]["chip"], "instance": "other"}, "value": [0, "7.5"]})  # last write wins
    en.append({"metric": {"chip": ch[1]["metric"]["chip"]}, "value": [0, "+Inf"]})                       # parseFloat -> NaN
    mx.append({"metric": {"chip": ch[2]["metric"]["chip"]}, "value": [0, "NaN"]})
    en.append({"metric": {}, "value": [0, "1"]})                                                         # no chip label
    un.append({"metric": {"instance": ch[5]["metric"]["instance"], "node": "renamed"}, "value": [0, "1"]})
    ch.append({"metric": {"chip": "lonely"}, "value": [0, "1"]})                                         # no instance


@pytest.mark.parametrize("n_nodes,chips,dirty", [(50, 4, False), (50, 4, True), (1, 1, False), (2000, 8, True)])
def test_join_matches_reference_restatement(ctx, n_nodes, chips, dirty):
    ch, en, mx, un = synth.synth_power_results(n_nodes, chips)
    if dirty:
        mutate(ch, en, mx, un)
    got = metrics.joinGpuMetrics(ch, en, mx, un, ctx)
    ref = ro.join_gpu_metrics(ch, en, mx, un)
    assert nan_eq(got["chips"], ref)
    fp = ro.fleet_power(ref)
    for k in ("totalWatts", "maxTotalWatts"):
        a, b = got["fleet"][k], fp[k]
        assert (a != a and b != b) or abs(a - b) <= 1e-5 * abs(b)
    assert nan_eq(got["fleet"]["pct"], fp["pct"])


def test_two_node_chip_label_collision(ctx):
    """the maps are keyed by `chip` only, so two nodes reusing a PCI address share one
    power value — the later sample (metrics.ts:130 Map.set)."""
    ch = [{"metric": {"chip": "0000:03:00_0", "instance": "a:9100"}, "value": [0, "1"]},
          {"metric": {"chip": "0000:03:00_0", "instance": "b:9100"}, "value": [0, "1"]}]
    en = [{"metric": {"chip": "0000:03:00_0", "instance": "a:9100"}, "value": [0, "100.5"]},
          {"metric": {"chip": "0000:03:00_0", "instance": "b:9100"}, "value": [0, "42.25"]}]
    un = [{"metric": {"instance": "a:9100", "nodename": "na"}, "value": [0, "1"]}, {"metric": {"instance": "b:9100", "nodename": "nb"}, "value": [0, "1"]}]
    got = metrics.joinGpuMetrics(ch, en, [], un, ctx)
    assert [c["powerWatts"] for c in got["chips"]] == [42.25, 42.25]
    assert [c["nodeName"] for c in got["chips"]] == ["na", "nb"]
    assert got["fleet"]["totalWatts"] == 84.5 and got["fleet"]["pct"] is None
    assert got["chips"] == ro.join_gpu_metrics(ch, en, [], un)


def test_empty_and_formatters(ctx):
    got = metrics.joinGpuMetrics([], [], [], [], ctx)
    assert got["chips"] == [] and got["fleet"]["totalWatts"] == 0 and got["fleet"]["pct"] is None
    assert metrics.formatWatts(45.3) == "45.3 W" and metrics.formatPercent(45.3, 120) == "38%" and metrics.formatPercent(1, 0) == "—"
    chips = [{"powerWatts": 45.3, "powerMaxWatts": 120.0}, {"powerWatts": None, "powerMaxWatts": 150.0}]
    fp = metrics.fleetPower(chips, ctx)
    assert fp == ro.fleet_power(chips)
