#!/usr/bin/env python
"""bench.py — fleet rollup throughput on B200 (BASELINE.json metric:
"pods+power samples aggregated/sec at 1/2/4/8 B200; % of HBM roofline").

One step = one pass of the hot path over one rank's shard of a synthetic fleet:
    K1  allocation rollup + classification over the C2 fleet shard (10k nodes / 1M pods per GPU)
    K3+K4  W=60 window min/mean/max AND per-node p95 over the C4 series shard (80,000 series x 86,400 f32
           per GPU) in one read (igpu_series_rollup)
units per step = pods + power samples.  Sharding is by node: every rank owns its own
node range, pods and series; the only exchange is the all-reduce of the small fleet totals, done by
the library over NCCL on the same stream.  scaling = weak (per-GPU shard fixed).

  value  : resident — inputs already in HBM when the timed region starts (CUDA events, max over ranks)
  e2e    : the same step through the public C-ABI calls with HOST buffers: every step copies the fleet
           columns and the whole series shard from pinned host memory, and copies the fleet outputs,
           window statistics and quantiles back (the series content is a pinned 1/10-shard chunk
           uploaded ten times, so that 8 ranks' host copies fit in RAM; the bytes moved are real)
  roofline: the dominant kernel's algorithmic bytes / its CUDA-event time, vs MEASURED_PEAKS.json
  cpu_baseline: the CPU oracle (C++ restatement of the reference, all host cores) on a bounded sample

`--impl reference` times the reference's own algorithm on the host cores.  The reference is a
TypeScript plugin with no native sources and Node.js is absent from this image, so that arm runs the
oracle port (oracle/rollup_oracle.cpp, OpenMP over all cores) — the only place besides tests and
cpu_baseline that executes oracle/.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# keep NCCL's version banner (NCCL_DEBUG=VERSION on the GPU boxes) off stdout — rank 0 prints exactly one JSON line.
# Must happen before torch / NCCL are loaded.
if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
    del os.environ["NCCL_DEBUG"]  # (NCCL prints the banner at WARN too; only an unset variable keeps it quiet)

import numpy as np  # noqa: E402

METRIC = "pods+power samples aggregated/sec"
UNIT = "rows+samples/s"
SEED = 20261106
WINDOW = 60
Q = 0.95


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    # workload shape per GPU (defaults = BASELINE configs C2 + C4); smaller values are for debugging only
    ap.add_argument("--nodes", type=int, default=10_000)
    ap.add_argument("--pods", type=int, default=1_000_000)
    ap.add_argument("--chips", type=int, default=8)
    ap.add_argument("--samples", type=int, default=86_400)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-series", type=int, default=0, help="--impl reference: series per pass (0 = the whole shard, the same config as the GPU arm)")
    ap.add_argument("--cpu-sample-series", type=int, default=8_000, help="series in the bounded sample of the GPU arm's cpu_baseline leg")
    ap.add_argument("--named-configs", default="auto", choices=["auto", "on", "off"],
                    help="BASELINE configs C3 / C4-strong / C5 + the all-reduce check under \"configs\" (auto: when WORLD_SIZE > 1)")
    ap.add_argument("--c3-nodes", type=int, default=100_000)
    ap.add_argument("--c3-pods", type=int, default=10_000_000)
    ap.add_argument("--c5-nodes-per-gpu", type=int, default=1_250)
    ap.add_argument("--c5-samples", type=int, default=1_250_000)
    ap.add_argument("--q", type=float, default=Q, help="quantile (0.95 = p95; 0.5 defeats the warp-level early-out: worst case)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-widened", action="store_true", help="skip the timing of the §8(f) kernels reported under \"widened\"")
    return ap.parse_args()


def workload_name(a):
    return (f"C2 alloc rollup ({a.nodes} nodes / {a.pods} pods, pod columns grouped by node by the host marshal) + C4 power series ({a.nodes * a.chips} series x "
            f"{a.samples} f32: W={WINDOW} window min/mean/max + per-node p95) per GPU")


def config_dict(a, world):
    """Identical for both arms: the driver compares the two lines' `config`."""
    samples = a.nodes * a.chips * a.samples
    return {"workload": workload_name(a), "parallelism": f"node-range shards x{world}, NCCL allreduce of fleet totals",
            "cache": f"inputs {4 * samples / 1e9:.1f} GB per GPU >> 126 MB L2 (no flush needed)", "seed": SEED,
            "window": WINDOW, "quantile": a.q}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------- CPU arm


def cpu_sample_inputs(a, threads, max_series):
    from headlamp_intel_gpu_plugin_b200 import synth
    from oracle import columnar as oracle

    cols = synth.synth_fleet(a.nodes, a.pods, SEED)
    n_series = min(max_series, a.nodes * a.chips) if max_series > 0 else a.nodes * a.chips
    n_series -= n_series % a.chips
    x = oracle.generate_series(n_series, a.samples, SEED, 0, 0, threads)
    off = synth.even_node_offsets(n_series // a.chips, a.chips)
    return cols, x, off


def cpu_step(cols, x, off, threads, q, wout=None):
    from oracle import columnar as oracle

    oracle.alloc_rollup(cols, threads)
    oracle.window_stats(x, WINDOW, threads, out=wout)
    oracle.quantile_by_node(x, off, q, 0, threads)


def run_cpu(a, steps, warmup, max_series):
    from oracle import columnar as oracle

    oracle.build()
    threads = os.cpu_count() or 1
    cols, x, off = cpu_sample_inputs(a, threads, max_series)
    units = cols.n_pods + x.size
    nw = (x.shape[1] + WINDOW - 1) // WINDOW
    wout = tuple(np.empty((x.shape[0], nw), np.float32) for _ in range(3))  # result planes allocated once, like the GPU arm's
    for _ in range(warmup):
        cpu_step(cols, x, off, threads, a.q, wout)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_step(cols, x, off, threads, a.q, wout)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    whole = x.shape[0] == a.nodes * a.chips
    sample = (f"full C2 fleet ({cols.n_pods} pods) + {'all ' if whole else ''}{x.shape[0]} of {a.nodes * a.chips} series x {a.samples} samples "
              f"({x.size / (a.nodes * a.chips * a.samples):.3f} of the series shard), {steps} timed passes")
    return units / dt, dt, threads, sample


def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # exactly the K timed / W warm-up passes asked for, each over the WHOLE per-GPU workload of the GPU arm (a pass is
    # ~2 s on 16 cores); the cap only guards against a request that could not end within minutes
    steps, warmup = max(1, min(a.steps, 60)), max(0, min(a.warmup, 20))
    value, dt, threads, sample = run_cpu(a, steps, warmup, a.cpu_series)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 (JS number) / f32 samples",
        "data": "synthetic", "config": config_dict(a, max(1, a.gpus)),
        "note": "reference = TypeScript, Node.js absent on this box: oracle port (C++ restatement of the reference's loops, OpenMP over all "
                "host cores) on one GPU's whole workload per pass; under torchrun rank 0 alone runs it",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0



# ------------------------------------------------------------------------------------- named BASELINE configs (N > 1)


def run_named_configs(a, ctx, rank, world, local, dev, barrier, peak):
    """BASELINE.json configs 3-5 on the ranks of this run, outside `value`, plus the multi-rank parity check:

    c3         100k nodes / 10M pods, every rank K1 over its node range (+ its share of the pods without a node)
    c4_strong  the 80,000 x 86,400 C4 slab split by node range over the ranks (strong scaling), fused W=60 + p95
    c5         1,250 nodes x 8 chips x 1.25e6 samples (50 GB) per rank, per-node p95 (igpu_series_quantile_by_node)
    allreduce_check  what the library's NCCL exchange returned on every rank, against (i) the same reduction of the
               ranks' local results done independently through torch.distributed, slot by slot, and (ii) the CPU oracle
               over a small whole fleet (all 96 int64 totals and the f64 fleet vector)
    Times are CUDA-event times, max over ranks."""
    import torch
    import torch.distributed as dist

    from headlamp_intel_gpu_plugin_b200 import columns as K
    from headlamp_intel_gpu_plugin_b200 import native, shard, synth

    def max_over_ranks(v):
        t = torch.tensor([float(v)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather(obj):
        if world == 1:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    stream = torch.cuda.current_stream()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx_local = native.Context(local, stream=stream.cuda_stream)  # same device, NO communicator: a rank's own partial results
    checks, res = {}, {}

    def timed(fn, reps, warm):
        for _ in range(warm):
            fn()
        kern = {}
        barrier()
        ev0.record(stream)
        for _ in range(reps):
            fn()
            for n, v in ctx.timing()["kernels"].items():
                kern[n] = kern.get(n, 0.0) + v / reps
        ev1.record(stream)
        barrier()
        return max_over_ranks(ev0.elapsed_time(ev1) / reps), {n: max_over_ranks(v) for n, v in sorted(kern.items())}

    def fleet_vector_from_node_stats(st):
        """[sum, count, min, max, nodes] of a rank, from the per-node statistics alone (float64 on the host)."""
        has = st["count"] > 0
        return [float(st["sum"].sum()), float(st["count"].sum()), float(st["min"][has].min()) if has.any() else float("nan"),
                float(st["max"][has].max()) if has.any() else float("nan"), float(has.sum())]

    def check_fleet_vector(name, lib_fleet, st):
        want = shard.merge_fleet_vectors([v + [0.0] * (K.FLEET_LEN - 5) for v in gather(fleet_vector_from_node_stats(st))])
        ok = (abs(lib_fleet[K.F_SUM] - want[K.F_SUM]) <= 1e-12 * abs(want[K.F_SUM]) and lib_fleet[K.F_COUNT] == want[K.F_COUNT]
              and lib_fleet[K.F_MIN] == want[K.F_MIN] and lib_fleet[K.F_MAX] == want[K.F_MAX] and lib_fleet[K.F_NODES] == want[K.F_NODES])
        checks[name] = bool(ok)

    # ---- C3 -----------------------------------------------------------------------------------------------------
    # node ranges with equal POD counts (the generator's pods are skewed towards the first nodes, like a real fleet's towards
    # its busy ones): equal node counts would leave rank 0 with 35 % of the pods at 8 ranks
    bounds = shard.balanced_node_ranges(synth.pods_per_node(a.c3_pods, a.c3_nodes, SEED), world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    cols = synth.synth_fleet(a.c3_nodes, a.c3_pods, SEED, node_range=(lo, hi), unscheduled_share=(rank, world))
    cols, _ = cols.grouped_by_node()
    cols.validate()
    pinned = native.pinned_copy_of_columns(cols)
    fleet = ctx.fleet_create(cols.n_nodes, cols.n_pods, cols.n_entries, cols.n_resources, cols.options)
    fleet.upload(pinned)
    ms, kern = timed(fleet.rollup, 10, 3)
    got = fleet.download()
    fleet.destroy()
    mine = ctx_local.alloc_rollup(cols).fleet_totals  # this rank's totals before any exchange
    indep = shard.allreduce_totals(mine)              # reduced over the ranks by torch.distributed, not by the library
    checks["c3_totals_96_slots_vs_torch_allreduce"] = bool(np.array_equal(got.fleet_totals, indep))
    pods_all = int(sum(gather(cols.n_pods)))
    k1 = kern.get("pod_rollup", 0.0) + kern.get("node_rollup", 0.0)
    c3_bytes = cols.nbytes_in() + cols.n_pods * (2 + (4 + 8 * cols.n_resources) * 2)
    res["c3"] = {"workload": f"{a.c3_nodes} nodes / {a.c3_pods} pods over {world} node-range shards of equal pod count (all pods: {pods_all}; "
                             f"rank 0: nodes {lo}-{hi}, {cols.n_pods} pods)",
                 "ms": ms, "kernels_ms": {n: kern[n] for n in ("pod_rollup", "node_rollup") if n in kern}, "pods_per_s": pods_all / (ms * 1e-3),
                 "pods_per_s_kernels_only": pods_all / (k1 * 1e-3) if k1 > 0 else None,
                 "rank0_algorithmic_bytes": int(c3_bytes), "rank0_frac": c3_bytes / (kern["pod_rollup"] * 1e-3) / 1e9 / peak if kern.get("pod_rollup") else None,
                 "gpu_pods": int(got.fleet_totals[K.T_GPU_PODS]), "allocated": int(got.fleet_totals[K.T_ALLOCATED])}
    del cols, pinned, got

    # ---- C4, strong scaling -----------------------------------------------------------------------------------------
    lo, hi = shard.node_range(a.nodes, rank, world)
    ns = (hi - lo) * a.chips
    series = ctx.series_create(ns, a.samples, synth.even_node_offsets(hi - lo, a.chips))
    series.generate(SEED, global_first=lo * a.chips)
    out = {}

    def c4():
        out["r"] = series.rollup(WINDOW, a.q, 0, fetch_windows=False)

    ms, kern = timed(c4, 5, 2)
    _, qv, st, fl = out["r"]
    check_fleet_vector("c4_strong_fleet_vector_vs_torch_gather", fl, st)
    total = a.nodes * a.chips * a.samples
    n_win = (a.samples + WINDOW - 1) // WINDOW
    b = 4 * ns * a.samples + 12 * ns * n_win + 4 * (hi - lo)
    dom = "series_fused"
    res["c4_strong"] = {"workload": f"{a.nodes * a.chips} series x {a.samples} f32 split by node range over {world} ranks ({ns} series on rank 0)",
                        "ms": ms, "kernels_ms": {n: kern[n] for n in ("series_fused", "quantile_sample", "quantile_finish", "quantile_tail") if kern.get(n)},
                        "samples_per_s": total / (ms * 1e-3),
                        "rank0_algorithmic_bytes": int(b), "rank0_frac": b / (kern[dom] * 1e-3) / 1e9 / peak,
                        "fleet": [float(v) for v in fl[:5]]}
    def c4q():
        out["fq"] = series.quantile_fleet(a.q, 0)

    ms_fq, _ = timed(c4q, 3, 1)
    res["c4_strong"]["fleet_p95"] = {"value": out["fq"][0], "valid_samples": int(out["fq"][1]), "ms": ms_fq,
                                     "note": "exact: a sampled bracket, one read of the slab that gathers the samples inside it, a radix select over those; histograms and counters all-reduced", "slab_reads": series.fleet_quantile_reads()}
    series.destroy()

    # ---- C5 -------------------------------------------------------------------------------------------------------
    n5, t5 = a.c5_nodes_per_gpu, a.c5_samples
    s5 = n5 * a.chips
    series = ctx.series_create(s5, t5, synth.even_node_offsets(n5, a.chips))
    series.generate(SEED + 5, global_first=rank * s5)

    def c5():
        out["r"] = series.quantile_by_node(a.q, 0)

    ms, kern = timed(c5, 3, 1)
    qv, st, fl = out["r"]
    check_fleet_vector("c5_fleet_vector_vs_torch_gather", fl, st)
    b = 4 * s5 * t5 + 4 * n5
    res["c5"] = {"workload": f"{n5} nodes x {a.chips} chips x {t5} samples per rank ({4 * s5 * t5 / 1e9:.1f} GB), {world} ranks = "
                             f"{world * s5 * t5:.3g} samples ({world}/8 of the 1e11-sample fleet)" if world != 8 else
                             f"{n5} nodes x {a.chips} chips x {t5} samples per rank ({4 * s5 * t5 / 1e9:.1f} GB), 8 ranks = {8 * s5 * t5:.3g} samples",
                 "ms": ms, "kernels_ms": {n: kern[n] for n in ("quantile_scan", "quantile_sample", "quantile_finish", "quantile_tail") if kern.get(n)},
                 "samples_per_s": world * s5 * t5 / (ms * 1e-3),
                 "rank_algorithmic_bytes": int(b), "scan_frac": b / (kern["quantile_scan"] * 1e-3) / 1e9 / peak if kern.get("quantile_scan") else None,
                 "whole_call_frac": b / (ms * 1e-3) / 1e9 / peak, "fleet": [float(v) for v in fl[:5]], "p95_rank0_node0": float(qv[0])}
    series.destroy()

    # ---- the exchange step against the CPU oracle, on a fleet small enough for rank 0 to roll up whole ---------------------
    NODES, CHIPS, T, PODS = 40 * world, 4, 2880, 20_000 * world
    whole = synth.synth_fleet(NODES, PODS, SEED + 9)
    mine, _rows = shard.shard_fleet(whole, rank, world)  # routes the pods without a node round-robin
    got = ctx.alloc_rollup(mine)
    lo, hi = shard.node_range(NODES, rank, world)
    series = ctx.series_create((hi - lo) * CHIPS, T, synth.even_node_offsets(hi - lo, CHIPS))
    series.generate(SEED, global_first=lo * CHIPS, nan_per_million=1000)
    _, _, _, fl = series.rollup(WINDOW, a.q, 0, fetch_windows=False)
    fq, fcnt = series.quantile_fleet(a.q, 0)          # a small fleet: three exact passes, ncclAllReduce(int64[4096]) after each
    _, fcurve = series.window_rollup(WINDOW, nodes=False)  # fleet window planes, all-reduced (one read of the slab)
    # the same two answers by the paths a large fleet takes: the sampled bracket (forced onto the small fleet), and the window
    # planes folded from what a rollup of the same window kept
    os.environ["IGPU_FLEET_QUANTILE_MIN_SAMPLE"] = "1"
    ctx.reload_env()
    fq_b, fcnt_b = series.quantile_fleet(a.q, 0)
    reads_b = series.fleet_quantile_reads()
    del os.environ["IGPU_FLEET_QUANTILE_MIN_SAMPLE"]
    ctx.reload_env()
    series.keep_window_counts(True)
    series.rollup(WINDOW, a.q, 0, fetch_windows=False)
    _, fcurve_k = series.window_rollup(WINDOW, nodes=False)
    reads_k = series.last_window_rollup_slab_reads
    series.destroy()
    same = gather((fq, fcnt, float(np.nansum(fcurve["mean"].astype(np.float64)))))
    checks["fleet_quantile_same_on_every_rank"] = all(v == same[0] for v in same)
    checks["fleet_quantile_bracketed_equals_exact_passes"] = bool(fq_b == fq and fcnt_b == fcnt and reads_b == 1)
    checks["fleet_window_planes_from_kept_counts_equal_the_read"] = bool(
        reads_k == 0 and np.array_equal(fcurve_k["count"], fcurve["count"]) and np.array_equal(fcurve_k["min"], fcurve["min"], equal_nan=True)
        and np.array_equal(fcurve_k["max"], fcurve["max"], equal_nan=True) and np.allclose(fcurve_k["mean"], fcurve["mean"], rtol=1e-6, equal_nan=True))
    if rank == 0:
        from oracle import columnar as oracle  # the checker, never the thing measured

        oracle.build()
        want = oracle.alloc_rollup(whole, threads=4)
        bad = [int(i) for i in np.nonzero(got.fleet_totals != want.fleet_totals)[0]]
        checks["small_fleet_totals_96_slots_vs_oracle"] = not bad
        if bad:
            checks["small_fleet_bad_slots"] = bad
        x = oracle.generate_series(NODES * CHIPS, T, SEED, 0, 1000, threads=4)
        _, _, ofl = oracle.quantile_by_node(x, synth.even_node_offsets(NODES, CHIPS), a.q, threads=4)
        checks["small_fleet_vector_vs_oracle"] = bool(np.allclose(fl[:5], ofl[:5], rtol=1e-12, atol=0))
        oq, ocnt = oracle.quantile_fleet(x, a.q, 0)
        checks["small_fleet_quantile_vs_oracle"] = bool(fq == oq and fcnt == ocnt)
        _, ocurve = oracle.window_rollup(x, synth.even_node_offsets(NODES, CHIPS), WINDOW)
        checks["small_fleet_window_planes_vs_oracle"] = bool(
            np.array_equal(fcurve["count"], ocurve["count"]) and np.array_equal(fcurve["min"], ocurve["min"], equal_nan=True)
            and np.array_equal(fcurve["max"], ocurve["max"], equal_nan=True) and np.allclose(fcurve["mean"], ocurve["mean"], rtol=1e-5, equal_nan=True))
    ctx_local.close()
    all_ok = all(all(v for k, v in c.items() if isinstance(v, bool)) for c in gather(checks))
    res["allreduce_check"] = "pass" if all_ok else "FAIL"
    res["allreduce_check_detail"] = checks
    return res


# ------------------------------------------------------------------------------------- GPU arm


def main_gpu(a):
    import torch
    import torch.distributed as dist

    from headlamp_intel_gpu_plugin_b200 import columns as K
    from headlamp_intel_gpu_plugin_b200 import native, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.current_stream()
    ctx = native.Context(local, stream=stream.cuda_stream)  # kernels run on torch's current stream
    if world > 1:
        # hand the library its own NCCL communicator (id exchanged over torch.distributed)
        uid = [ctx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- inputs: this rank's node range of an (N x nodes)-node fleet ---------------------------------
    n_series = a.nodes * a.chips
    cols = synth.synth_fleet(a.nodes, a.pods, SEED + rank)
    # the host marshal lays large fleets out with the pods of a node adjacent (k8s.rollupFleet, group_by_node): the same
    # fleet, the same results, one warp reduction + one set of atomics per node run on the device
    cols, _ = cols.grouped_by_node()
    cols.validate()
    pinned_cols = native.pinned_copy_of_columns(cols)
    pinned_out = native.pinned_outputs(cols.n_nodes, cols.n_pods, cols.n_resources, cols.options)
    fleet = ctx.fleet_create(cols.n_nodes, cols.n_pods, cols.n_entries, cols.n_resources, cols.options)
    fleet.upload(pinned_cols)
    series = ctx.series_create(n_series, a.samples, synth.even_node_offsets(a.nodes, a.chips))
    series.generate(SEED, global_first=rank * n_series)
    ctx.sync()
    samples = n_series * a.samples
    units = cols.n_pods + samples

    def step_resident():
        fleet.rollup()
        # window min/mean/max + per-node p95 in one read of the slab; syncs; copies N quantiles + node stats back
        return series.rollup(WINDOW, a.q, 0, fetch_windows=False)

    # ---- resident timing -----------------------------------------------------------------------------
    for _ in range(a.warmup):
        step_resident()
    kern_ms = {n: 0.0 for n in native.KERNEL_NAMES}
    launches0 = ctx.timing()["kernel_launches"]
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(a.steps):
        step_resident()
        for n, v in ctx.timing()["kernels"].items():
            kern_ms[n] += v
    ev1.record(stream)
    barrier()
    clock_info = clocks.stop() if rank == 0 else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = ctx.timing()["kernel_launches"] - launches0
    value = units * world * a.steps / (ms_total * 1e-3)
    kern_ms = {n: v / a.steps for n, v in kern_ms.items()}

    # ---- roofline of the dominant kernel ----------------------------------------------------------------
    peak, peak_src = peaks()
    n_win = (a.samples + WINDOW - 1) // WINDOW
    alg_bytes = {
        # in: 16E + 9P + 21N; out per pod: is_gpu 1 + gpu_containers 1 + (mask 4 + R x 8) for requests and again for limits
        "pod_rollup": cols.nbytes_in() + cols.n_pods * (2 + (4 + 8 * cols.n_resources) * (2 if cols.options & K.OPT_POD_LIMITS else 1)),
        "window_stats": 4 * samples + 12 * n_series * n_win,
        "quantile_scan": 4 * samples + 4 * a.nodes,
        "series_fused": 4 * samples + 12 * n_series * n_win + 4 * a.nodes,
    }
    kernels = {}
    for n, b in alg_bytes.items():
        if kern_ms[n] > 0:
            gbs = b / (kern_ms[n] * 1e-3) / 1e9
            kernels[n] = {"ms": kern_ms[n], "algorithmic_bytes": int(b), "achieved_gbs": gbs, "frac": gbs / peak}
    for n in ("node_rollup", "quantile_sample", "quantile_finish", "quantile_tail"):
        kernels[n] = {"ms": kern_ms[n]}
    if rank == 0 and "pod_rollup" in kernels:
        # what the device can do at K1's size: a plain device-to-device copy that moves as many bytes (read + written) as K1's
        # algorithmic traffic.  156 MB is far from the 4 GB transfer MEASURED_PEAKS.json quotes its peak on (launch, ramp, tail).
        nb = int(alg_bytes["pod_rollup"])
        src_t = torch.empty(nb // 2, dtype=torch.uint8, device=dev)
        dst_t = torch.empty_like(src_t)
        for _ in range(5):
            dst_t.copy_(src_t)
        ts = []
        for _ in range(15):
            ev0.record(stream)
            dst_t.copy_(src_t)
            ev1.record(stream)
            torch.cuda.synchronize()
            ts.append(ev0.elapsed_time(ev1))
        ts.sort()
        kernels["pod_rollup"]["same_traffic_copy_ms"] = ts[len(ts) // 2]
        kernels["pod_rollup"]["frac_of_same_traffic_copy"] = ts[len(ts) // 2] / kernels["pod_rollup"]["ms"]
        del src_t, dst_t
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tp):
        for n, t in json.load(open(tp)).items():
            if isinstance(t, dict) and (t.get("samples") == samples or t.get("pods") == cols.n_pods):
                traffic[n] = int(t["dram_bytes_read"] + t["dram_bytes_write"])
    # SURVEY §8(d) metric (i) and (ii) separately: the allocation rollup's kernels, and every kernel of the series pass
    k1_ms = kern_ms["pod_rollup"] + kern_ms["node_rollup"]
    series_ms = sum(kern_ms[n] for n in ("series_fused", "window_stats", "quantile_scan", "quantile_sample", "quantile_finish", "quantile_tail"))
    dom = max(alg_bytes, key=lambda n: kern_ms[n])
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                "frac": kernels[dom]["frac"], "traffic": traffic.get(dom), "algorithmic_bytes": int(alg_bytes[dom]),
                "traffic_source": "ncu --set full capture, profiles/ncu_traffic.json" if dom in traffic else None, "peak_source": peak_src,
                "frac_of_nominal_8TBs": kernels[dom]["achieved_gbs"] / 8000.0}

    # ---- the fused pass at the median (outside `value`): the warp-level early-out of the classification cannot fire there
    # (no lane's maximum stays below the lower splitter), so this is the kernel's data-independent worst case
    worst = None
    if rank == 0 and world == 1 and not a.no_widened and a.q != 0.5:
        for _ in range(2):
            series.rollup(WINDOW, 0.5, 0, fetch_windows=False)
        acc = {"series_fused": 0.0, "quantile_sample": 0.0, "quantile_finish": 0.0, "quantile_tail": 0.0}
        for _ in range(5):
            series.rollup(WINDOW, 0.5, 0, fetch_windows=False)
            for n in acc:
                acc[n] += ctx.timing()["kernels"][n] / 5
        if acc["series_fused"] > 0:
            b = alg_bytes["series_fused"]
            worst = {"ms": acc["series_fused"], "achieved_gbs": b / acc["series_fused"] / 1e6, "frac": b / acc["series_fused"] / 1e6 / peak,
                     "kernels_ms": acc}

    # ---- the fused pass on a slab with missing samples (real scrapes have gaps; outside `value`): windows that hold a NaN take
    # a second, NaN-aware fold of their sum and count.  The slab is regenerated for this and restored afterwards.
    with_gaps = None
    if rank == 0 and world == 1 and not a.no_widened:
        with_gaps = {}
        for ppm in (1000, 20_000):
            series.generate(SEED, global_first=rank * n_series, nan_per_million=ppm)
            for _ in range(2):
                series.rollup(WINDOW, a.q, 0, fetch_windows=False)
            acc_ms = 0.0
            for _ in range(3):
                series.rollup(WINDOW, a.q, 0, fetch_windows=False)
                acc_ms += ctx.timing()["kernels"]["series_fused"] / 3
            if acc_ms > 0:
                with_gaps[f"{ppm}_ppm_missing"] = {"ms": acc_ms, "frac": alg_bytes["series_fused"] / acc_ms / 1e6 / peak}
        series.generate(SEED, global_first=rank * n_series)
        ctx.sync()

    # ---- end to end: host buffers in, host buffers out, every step ------------------------------------------
    e2e = None
    if not a.no_e2e:
        parts = 10 if n_series % 10 == 0 and (n_series // 10) % a.chips == 0 else 1
        chunk_rows = n_series // parts
        host_chunk = native.PinnedBuffer(chunk_rows * a.samples * 4)
        hx = host_chunk.view(np.float32, (chunk_rows, a.samples))
        series.ctx.sync()
        hx[...] = series.download(0, chunk_rows)  # the first tenth of the shard, now in pinned host memory
        wbuf = native.PinnedBuffer(3 * n_series * n_win * 4)
        wviews = [wbuf.view(np.float32, (n_series, n_win), i * n_series * n_win * 4) for i in range(3)]
        h2d = pinned_cols.nbytes_in() + parts * hx.nbytes
        d2h = pinned_out.nbytes_out() + 3 * n_series * n_win * 4 + a.nodes * (8 + 24) + 2 * 8 * K.FLEET_LEN
        # The public call for a host-resident slab: node-range chunks ping-pong between two contexts so the upload of
        # chunk i+1 overlaps the rollup and the download of chunk i (metrics.StreamedSeriesRollup).  The resident slab
        # is released first: streaming needs two chunk slabs, not the shard.
        series.destroy()
        series = None
        from headlamp_intel_gpu_plugin_b200 import metrics, shard
        streamer = metrics.StreamedSeriesRollup(chunk_rows, a.samples, synth.even_node_offsets(a.nodes // parts, a.chips), device=local)
        chunk_list = [hx] * parts

        def step_e2e():
            fleet.upload(pinned_cols)
            fleet.rollup()
            fleet.download(pinned_out)
            qv, st_, fl = streamer.run(chunk_list, WINDOW, a.q, 0, out_windows=wviews)
            if world > 1:  # fleet vector over the ranks (the lanes carry no communicator)
                allv = [None] * world
                dist.all_gather_object(allv, fl.tolist())
                fl = shard.merge_fleet_vectors(allv)
            return qv

        step_e2e()
        barrier()
        ev0.record(stream)
        for _ in range(a.e2e_steps):
            step_e2e()
        ev1.record(stream)
        barrier()
        ms2 = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
        if world > 1:
            dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
        e2e = {"value": units * world * a.e2e_steps / (float(ms2.item()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": float(ms2.item()) / a.e2e_steps, "steps": a.e2e_steps,
               "note": f"series shard streamed as {parts} node-range chunks (each a copy of one pinned {hx.nbytes / 1e9:.2f} GB chunk) through "
                       "two contexts: upload of chunk i+1 overlaps rollup + download of chunk i"}
        streamer.close()
        host_chunk.free()
        wbuf.free()

    # ---- CPU baseline (rank 0, N = 1 only) ----------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        v, dt, threads, sample = run_cpu(a, 2, 1, a.cpu_sample_series)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample, "s_per_pass": dt}

    # ---- widened rows, outside the timed region and outside `value`: N3 rate() kernel ----------
    widened = None
    if rank == 0 and world == 1 and not a.no_widened:
        widened = {}
        if worst is not None:
            widened["series_fused_worst_case_q0.5"] = worst
        if with_gaps:
            widened["series_fused_with_missing_samples"] = with_gaps
        S5, T5, STEP5 = 80_000, 10_099, 15_000
        cn = ctx.counters_create(S5, T5)
        prm = native.Counters.params(STEP5, 20)
        pw = ctx.series_create(S5, cn.rate_len(prm), synth.even_node_offsets(S5 // 8, 8))
        nb = 8 * S5 * T5 + 4 * S5 * cn.rate_len(prm)
        # clean counters take the streaming kernel; with 2 % of the samples missing the device picks the gap kernel (rate.cu)
        for key, what, reset, missing in (("counter_rate", "no gaps", 0, 0),
                                          ("counter_rate_gaps", "2 % of the samples missing, a counter reset every 2,000 steps", 2000, 20_000)):
            cn.generate(SEED, 0, STEP5, reset, missing)
            for _ in range(3):
                cn.rate(prm, pw)
            acc = 0.0
            for _ in range(5):
                cn.rate(prm, pw)
                ctx.sync()
                acc += ctx.timing()["kernels"]["counter_rate"]
            widened[key] = {"workload": f"rate(counter[5m]) at a 15 s step, {S5} series x {T5} float64 samples, {what}", "ms": acc / 5,
                            "algorithmic_bytes": nb, "achieved_gbs": nb / (acc / 5) / 1e6, "frac": nb / (acc / 5) / 1e6 / peak}
        pw.destroy()
        cn.destroy()

    fleet.destroy()
    if series is not None:
        series.destroy()
        series = None
    configs = None
    if a.named_configs == "on" or (a.named_configs == "auto" and world > 1):
        configs = run_named_configs(a, ctx, rank, world, local, dev, barrier, peak)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64 (allocation) / f32 samples with f64 accumulation", "data": "synthetic",
            "config": config_dict(a, world),
            "pods_per_s": world * cols.n_pods / (k1_ms * 1e-3) if k1_ms > 0 else None,
            "samples_per_s": world * samples / (series_ms * 1e-3) if series_ms > 0 else None,
            "roofline": roofline, "kernels": kernels, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clock_info, "widened": widened, "configs": configs,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    args = parse_args()
    sys.exit(main_reference(args) if args.impl == "reference" else main_gpu(args))
