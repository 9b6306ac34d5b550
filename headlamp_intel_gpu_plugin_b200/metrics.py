"""Host side of the power path: the exported surface of the reference's src/api/metrics.ts with the
join and the fleet sums moved behind the C ABI, plus the windowed statistics the north star adds.

* `joinGpuMetrics` is the synchronous core of `fetchGpuMetrics` (metrics.ts:118-149): it takes the
  four Prometheus instant-vector results and returns the same `chips` array.  Service discovery and
  the HTTP round trips (metrics.ts:61-90, :101-116) are network I/O and stay with the caller.
* `fleetPower` is the reduce + PowerBar arithmetic of MetricsPage.tsx:330-332 / :50-53.
* `formatWatts` / `formatPercent` (metrics.ts:161-168) are string presentation and stay on the host.
* `PowerSeries` wraps the device-resident [series][time] slab: `window_stats` (min/mean/max
  downsampling) and `quantile_by_node` (p95) have no reference twin (metrics.ts has only instant
  queries); their semantics are PromQL's *_over_time, see include/igpu_rollup.h.

Everything numeric runs on the GPU through libigpu_rollup.so; there is no CPU fallback.
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

from . import columns as K
from . import jsnum
from . import native


def _metric(r: Any) -> dict:
    m = r.get("metric") if isinstance(r, dict) else None
    return m if isinstance(m, dict) else {}


class _Dict:
    """String -> dense id dictionary (host side of the dictionary encoding)."""

    def __init__(self):
        self.ids: Dict[Any, int] = {}
        self.names: List[Any] = []

    def get(self, key) -> int:
        i = self.ids.get(key)
        if i is None:
            i = self.ids[key] = len(self.names)
            self.names.append(key)
        return i


def marshalMetrics(chipResults: Sequence[dict], energyRateResults: Sequence[dict], powerMaxResults: Sequence[dict],
                   unameResults: Sequence[dict]) -> K.PowerColumns:
    """Prometheus result objects -> PowerColumns (ids + parseFloat'ed values)."""
    chips, insts, nodes = _Dict(), _Dict(), _Dict()
    none = K.NODE_NONE
    un_inst, un_node = [], []
    for r in unameResults:  # metrics.ts:120-124
        m = _metric(r)
        inst = m.get("instance")
        nodename = m.get("nodename")
        if nodename is None:
            nodename = m.get("node")
        if nodename is None:
            nodename = inst
        if jsnum.truthy(inst):
            un_inst.append(insts.get(inst))
            un_node.append(nodes.get(nodename))
    def values(results):
        ids, vals = [], []
        for r in results:  # metrics.ts:128-131 / :135-138
            chip = _metric(r).get("chip")
            ids.append(chips.get(chip) if jsnum.truthy(chip) else none)
            v = r.get("value") if isinstance(r, dict) else None
            vals.append(jsnum.parse_float(v[1]) if isinstance(v, (list, tuple)) and len(v) > 1 else math.nan)
        return ids, vals
    en_chip, en_val = values(energyRateResults)
    mx_chip, mx_val = values(powerMaxResults)
    c_chip, c_inst = [], []
    for r in chipResults:  # metrics.ts:141-144
        m = _metric(r)
        chip = m.get("chip")
        inst = m.get("instance")
        c_chip.append(chips.get("" if chip is None else chip))
        c_inst.append(insts.get("" if inst is None else inst))
    return K.PowerColumns(
        chip_chip=c_chip, chip_instance=c_inst, energy_chip=en_chip, energy_value=en_val, power_max_chip=mx_chip,
        power_max_value=mx_val, uname_instance=un_inst, uname_node=un_node, n_chip_ids=len(chips.names),
        n_instance_ids=len(insts.names), chip_names=chips.names, instance_names=insts.names, node_names=nodes.names,
    )


def _decode_chips(p: K.PowerColumns, out: dict) -> List[dict]:
    rows = []
    for i in range(len(p.chip_chip)):
        fl = int(out["chip_flags"][i])
        inst = p.instance_names[int(p.chip_instance[i])]
        rows.append(
            {
                "nodeName": p.node_names[int(out["chip_node"][i])] if fl & K.POWER_HAS_NODE else inst,
                "chip": p.chip_names[int(p.chip_chip[i])],
                "instance": inst,
                "powerWatts": float(out["chip_watts"][i]) if fl & K.POWER_HAS_WATTS else None,
                "powerMaxWatts": float(out["chip_max_watts"][i]) if fl & K.POWER_HAS_MAX else None,
            }
        )
    return rows


def joinGpuMetrics(chipResults, energyRateResults, powerMaxResults, unameResults, ctx: Optional[native.Context] = None) -> dict:
    """The join of fetchGpuMetrics (metrics.ts:118-149) plus the fleet sums, in one GPU call.
    Returns {'chips': [...GpuChipMetrics...], 'fleet': {'totalWatts', 'maxTotalWatts', 'pct'}}."""
    p = marshalMetrics(chipResults, energyRateResults, powerMaxResults, unameResults)
    ctx = ctx or native.default_context()
    out = ctx.power_join_sum(p)
    pct = float(out["fleet"][2])
    return {
        "chips": _decode_chips(p, out),
        "fleet": {
            "totalWatts": float(out["fleet"][0]),
            "maxTotalWatts": float(out["fleet"][1]),
            # PowerBar: null when maxTotal is not > 0 (MetricsPage.tsx:51,331)
            "pct": None if not (out["fleet"][1] > 0) else pct,
        },
    }


def fleetPower(chips: Sequence[dict], ctx: Optional[native.Context] = None) -> dict:
    """MetricsPage.tsx:330-332 over an already joined chips array (one row per chip)."""
    n = len(chips)
    ids = np.arange(n, dtype=np.uint32)
    has_w = np.array([c.get("powerWatts") is not None for c in chips], bool)
    has_m = np.array([c.get("powerMaxWatts") is not None for c in chips], bool)
    p = K.PowerColumns(
        chip_chip=ids, chip_instance=np.zeros(n, np.uint32),
        energy_chip=ids[has_w], energy_value=[c["powerWatts"] for c in chips if c.get("powerWatts") is not None],
        power_max_chip=ids[has_m], power_max_value=[c["powerMaxWatts"] for c in chips if c.get("powerMaxWatts") is not None],
        uname_instance=[], uname_node=[], n_chip_ids=max(n, 1), n_instance_ids=1,
    )
    out = (ctx or native.default_context()).power_join_sum(p)
    return {"totalWatts": float(out["fleet"][0]), "maxTotalWatts": float(out["fleet"][1]),
            "pct": None if not (out["fleet"][1] > 0) else float(out["fleet"][2])}


def formatWatts(w: float) -> str:
    """metrics.ts:161-163."""
    return f"{jsnum.to_fixed_1(w)} W"


def formatPercent(used: float, maximum: float) -> str:
    """metrics.ts:165-168."""
    if maximum <= 0:
        return "—"
    return f"{jsnum.number_to_string(jsnum.math_round((used / maximum) * 100))}%"


class PowerSeries:
    """Device-resident power samples of a fleet (or of one rank's node range)."""

    def __init__(self, n_series: int, n_time: int, node_series_off: Sequence[int], ctx: Optional[native.Context] = None):
        self.ctx = ctx or native.default_context()
        self.h = self.ctx.series_create(n_series, n_time, node_series_off)
        self.n_series, self.n_time = n_series, n_time
        self.n_nodes = len(node_series_off) - 1

    @classmethod
    def from_array(cls, samples: np.ndarray, node_series_off: Sequence[int], ctx: Optional[native.Context] = None) -> "PowerSeries":
        samples = np.ascontiguousarray(samples, np.float32)
        s = cls(samples.shape[0], samples.shape[1], node_series_off, ctx)
        s.h.upload(samples)
        return s

    def window_stats(self, window: int = 60, fetch: bool = True):
        """min / mean / max over non-overlapping windows of `window` samples per series."""
        return self.h.window_stats(window, fetch)

    def quantile_by_node(self, q: float = 0.95, mode: int = 0):
        """(quantile[n_nodes] float64, per-node stats, fleet vector)."""
        return self.h.quantile_by_node(q, mode)

    def rollup(self, window: int = 60, q: float = 0.95, mode: int = 0, fetch_windows: bool = True):
        """Both of the above in one read of the slab: ((min, mean, max) | None, quantile, stats, fleet)."""
        return self.h.rollup(window, q, mode, fetch_windows)

    def quantile_fleet(self, q: float = 0.95, mode: int = 0):
        """The quantile over every valid sample of the fleet (north star: per-fleet p95) -> (value, valid samples)."""
        return self.h.quantile_fleet(q, mode)

    def window_rollup(self, window: int = 60, nodes: bool = True):
        """Window min / mean / max / count over each node's chips and over the fleet (the range view's curves)."""
        return self.h.window_rollup(window, nodes)

    def range_view(self, window: int = 60, q: float = 0.95, node_names: Optional[Sequence[str]] = None, tdp_watts: Optional[Sequence[float]] = None,
                   fleet_views: bool = False) -> dict:
        """The numbers of the range view the plugin only mocks up (docs/screenshots/03-metrics.svg:26-45, 85-89):
        "Peak Power Draw" and the node it belongs to, "Avg Power (cluster)" over the nodes that reported, optional
        "Avg TDP Utilization", and per node avg / peak / p95 — all from one fused pass over the slab.  fleet_views=True
        adds the fleet's own quantile (exact; one more read of the slab for a fleet large enough to sample) and the curve the mock-up
        draws (window min / avg / peak over all chips), which is folded from the window planes the fused pass leaves on
        the device together with its valid-sample counts — no further read where the fused pass takes the window."""
        if fleet_views:
            self.h.keep_window_counts(True)
        _, qv, st, fleet = self.h.rollup(window, q, 0, fetch_windows=False)
        cnt = st["count"].astype(np.float64)
        with np.errstate(invalid="ignore", divide="ignore"):
            avg = np.where(cnt > 0, st["sum"] / cnt, np.nan)
        have = cnt > 0
        peak_node = int(np.nanargmax(np.where(have, st["max"], -np.inf))) if have.any() else None
        name = (lambda i: node_names[i] if node_names is not None else i)
        view = {
            "peakWatts": float(st["max"][peak_node]) if peak_node is not None else None,
            "peakNode": name(peak_node) if peak_node is not None else None,
            "avgWattsCluster": float(fleet[K.F_SUM] / fleet[K.F_COUNT]) if fleet[K.F_COUNT] > 0 else None,
            "nodesReporting": int(fleet[K.F_NODES]),
            "nodes": [{"node": name(i), "avgWatts": float(avg[i]), "peakWatts": float(st["max"][i]), "minWatts": float(st["min"][i]),
                       f"p{round(q * 100)}Watts": float(qv[i])} for i in range(self.n_nodes)],
        }
        if tdp_watts is not None:
            tdp = np.asarray(tdp_watts, np.float64)
            ok = have & (tdp > 0)
            view["avgTdpUtilizationPct"] = float(np.mean(avg[ok] / tdp[ok]) * 100.0) if ok.any() else None
        if fleet_views:
            view[f"p{round(q * 100)}WattsCluster"] = self.h.quantile_fleet(q, 0)[0]
            _, curve = self.h.window_rollup(window, nodes=False)
            view["curve"] = {"minWatts": curve["min"], "avgWatts": curve["mean"], "peakWatts": curve["max"], "samples": curve["count"]}
        return view

    def close(self):
        self.h.destroy()


class StreamedSeriesRollup:
    """Range-view rollup of a slab that lives in host memory, pipelined over PCIe.

    The slab is cut into equal node-range chunks (the same cut `shard.py` makes across GPUs, here made in time on one
    GPU).  Chunks ping-pong between two contexts — two streams — so chunk i+1 is uploading while chunk i is being
    rolled up and its window planes are downloading: PCIe is full duplex (measured on this box: 55.6 GB/s up alone,
    100 GB/s both ways at once, profiles/microbench/mb_pcie_r01.txt) and the kernels run at ~100x the link rate, so
    the step costs one pass of the upload link instead of upload + kernels + download.  Device memory: two chunk
    slabs, whatever the size of the fleet.  Per-node results are chunk-local and concatenated; the fleet vector is
    merged the way ranks' vectors are (`shard.merge_fleet_vectors`).
    """

    def __init__(self, chunk_series: int, n_time: int, chunk_node_series_off: Sequence[int], device: int = 0, n_lanes: int = 2):
        self.chunk_series, self.n_time = chunk_series, n_time
        self.chunk_nodes = len(chunk_node_series_off) - 1
        self.lanes = []
        for _ in range(n_lanes):
            ctx = native.Context(device)
            self.lanes.append((ctx, ctx.series_create(chunk_series, n_time, chunk_node_series_off)))

    def run(self, chunks: Sequence[np.ndarray], window: int = 60, q: float = 0.95, mode: int = 0, out_windows=None):
        """chunks: float32 [chunk_series][n_time] host arrays (pinned for asynchronous DMA).  out_windows: optional
        (min, mean, max) float32 arrays [len(chunks) * chunk_series][n_windows] to receive the window planes.
        Returns (quantile[all nodes], stats[all nodes], merged fleet vector)."""
        from . import shard

        n, nodes, rows = len(chunks), self.chunk_nodes, self.chunk_series
        qv = np.empty(n * nodes, np.float64)
        st = np.zeros(n * nodes, K.NODE_STATS_DTYPE)
        fleets = np.zeros((n, K.FLEET_LEN), np.float64)
        for i in range(min(len(self.lanes), n)):  # fill the pipe
            self.lanes[i][1].upload_async(chunks[i])
        for i in range(n):
            ctx, h = self.lanes[i % len(self.lanes)]
            w = [None, None, None] if out_windows is None else [o[i * rows:(i + 1) * rows] for o in out_windows]
            h.rollup_into(window, q, mode, w[0], w[1], w[2], qv[i * nodes:(i + 1) * nodes], st[i * nodes:(i + 1) * nodes], fleets[i])
            nxt = i + len(self.lanes)
            if nxt < n:  # this lane's slab is free again: queue its next chunk behind the other lane's upload
                h.upload_async(chunks[nxt])
        return qv, st, shard.merge_fleet_vectors(fleets)

    def close(self):
        for ctx, h in self.lanes:
            h.destroy()
            ctx.close()
        self.lanes = []


class EnergyCounters:
    """Device-resident cumulative energy counters (`node_hwmon_energy_joule_total`, joules) on a regular grid.

    Replaces the server-side `rate(node_hwmon_energy_joule_total[5m])` of metrics.ts:106-109 when the plugin fetches
    the raw counter as a `query_range` matrix: `power()` evaluates PromQL rate() on the device
    (Prometheus extrapolatedRate: counter resets, zero-point clamp, boundary extrapolation) and returns a
    PowerSeries that stays in HBM, ready for `rollup()`.
    """

    def __init__(self, n_series: int, n_time: int, step_ms: int = 15_000, ctx: Optional[native.Context] = None):
        self.ctx = ctx or native.default_context()
        self.h = self.ctx.counters_create(n_series, n_time)
        self.n_series, self.n_time, self.step_ms = n_series, n_time, step_ms

    @classmethod
    def from_array(cls, joules: np.ndarray, step_ms: int = 15_000, ctx: Optional[native.Context] = None) -> "EnergyCounters":
        """joules: float64 [series][time]; NaN where the step has no sample."""
        joules = np.ascontiguousarray(joules, np.float64)
        c = cls(joules.shape[0], joules.shape[1], step_ms, ctx)
        c.h.upload(joules)
        return c

    def power(self, node_series_off: Sequence[int], range_ms: int = 300_000, eval_stride: int = 1, closed_left: bool = False,
              first_eval: Optional[int] = None) -> PowerSeries:
        """rate(counter[range_ms]) at every eval_stride-th grid step, as a PowerSeries (watts)."""
        if range_ms % self.step_ms:
            raise ValueError(f"range {range_ms} ms is not a whole number of {self.step_ms} ms steps")
        prm = native.Counters.params(self.step_ms, range_ms // self.step_ms, first_eval, eval_stride,
                                     native.RATE_CLOSED_LEFT if closed_left else 0)
        n_out = self.h.rate_len(prm)
        if n_out == 0:
            raise ValueError("the slab is shorter than one range")
        out = PowerSeries(self.n_series, n_out, node_series_off, self.ctx)
        self.h.rate(prm, out.h)
        return out

    def close(self):
        self.h.destroy()


def marshalMatrix(response, start_ms: int, step_ms: int, n_time: int, labels: Sequence[str] = ("chip", "instance")):
    """Python marshal of a parsed Prometheus `query_range` response (dict with status/data.result, or the bare result
    list): (float64 [series][n_time] with NaN gaps, per-series label tuples, dropped sample count).
    Same contract as the native igpu_marshal_matrix_json, which tests hold equal to this function."""
    if isinstance(response, dict):
        if response.get("status") != "success":  # queryPrometheus, metrics.ts:73
            result = []
        else:
            data = response.get("data")
            result = data.get("result", []) if isinstance(data, dict) else []
    else:
        result = response or []
    values = np.full((len(result), n_time), np.nan)
    labs, dropped = [], 0
    for i, series in enumerate(result):
        metric = series.get("metric") or {}
        labs.append(tuple(metric.get(l) if isinstance(metric.get(l), str) else None for l in labels))
        for ts, text in series.get("values") or []:
            ms = round(float(ts) * 1000.0)
            rel = ms - start_ms
            if abs(float(ts) * 1000.0 - ms) < 1e-3 and rel >= 0 and rel % step_ms == 0 and rel // step_ms < n_time:
                if isinstance(text, str):  # parseFloat(r.value[1]), metrics.ts:130-137
                    values[i, rel // step_ms] = jsnum.parse_float(text)
                elif isinstance(text, (int, float)) and not isinstance(text, bool):
                    values[i, rel // step_ms] = float(text)
                else:
                    values[i, rel // step_ms] = np.nan
            else:
                dropped += 1
    return values, labs, dropped


def marshalMatrixJson(text, start_ms: int, step_ms: int, n_time: int, labels: Sequence[str] = ("chip", "instance"), threads: int = 0):
    """The same from the response TEXT, in native code (no json.loads)."""
    if isinstance(text, str):
        text = text.encode()
    return native.marshal_matrix_json(text, start_ms, step_ms, n_time, labels, threads)

