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
    ap.add_argument("--cpu-series", type=int, default=8_000, help="series in the bounded CPU sample")
    ap.add_argument("--q", type=float, default=Q, help="quantile (0.95 = p95; 0.5 defeats the warp-level early-out: worst case)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-widened", action="store_true", help="skip the timing of the §8(f) kernels reported under \"widened\"")
    return ap.parse_args()


def workload_name(a):
    return (f"C2 alloc rollup ({a.nodes} nodes / {a.pods} pods, pod columns grouped by node by the host marshal) + C4 power series ({a.nodes * a.chips} series x "
            f"{a.samples} f32: W={WINDOW} window min/mean/max + per-node p95) per GPU")


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


def cpu_sample_inputs(a, threads):
    from headlamp_intel_gpu_plugin_b200 import synth
    from oracle import columnar as oracle

    cols = synth.synth_fleet(a.nodes, a.pods, SEED)
    n_series = min(a.cpu_series, a.nodes * a.chips)
    n_series -= n_series % a.chips
    x = oracle.generate_series(n_series, a.samples, SEED, 0, 0, threads)
    off = synth.even_node_offsets(n_series // a.chips, a.chips)
    return cols, x, off


def cpu_step(cols, x, off, threads):
    from oracle import columnar as oracle

    oracle.alloc_rollup(cols, threads)
    oracle.window_stats(x, WINDOW, threads)
    oracle.quantile_by_node(x, off, Q, 0, threads)


def run_cpu(a, steps, warmup):
    from oracle import columnar as oracle

    oracle.build()
    threads = os.cpu_count() or 1
    cols, x, off = cpu_sample_inputs(a, threads)
    units = cols.n_pods + x.size
    for _ in range(warmup):
        cpu_step(cols, x, off, threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_step(cols, x, off, threads)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    sample = (f"full C2 fleet ({cols.n_pods} pods) + {x.shape[0]} of {a.nodes * a.chips} series x {a.samples} samples "
              f"({x.size / (a.nodes * a.chips * a.samples):.3f} of the series shard), {steps} timed passes")
    return units / dt, dt, threads, sample


def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # exactly the K timed / W warm-up passes asked for (a pass over the bounded sample is ~0.2 s on 16 cores); the cap only
    # guards against a request that could not end within minutes
    steps, warmup = max(1, min(a.steps, 400)), max(0, min(a.warmup, 100))
    value, dt, threads, sample = run_cpu(a, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 (JS number) / f32 samples",
        "data": "synthetic", "config": {"workload": workload_name(a), "note": "reference = TypeScript, Node.js absent: oracle port "
                                        "(C++ restatement, OpenMP) on a bounded sample of the same workload"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


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
        "pod_rollup": cols.nbytes_in() + cols.n_pods * (1 + 4 + 8 * cols.n_resources) * (2 if cols.options & K.OPT_POD_LIMITS else 1)
        - (cols.n_pods if cols.options & K.OPT_POD_LIMITS else 0),
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
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tp):
        for n, t in json.load(open(tp)).items():
            if isinstance(t, dict) and (t.get("samples") == samples or t.get("pods") == cols.n_pods):
                traffic[n] = int(t["dram_bytes_read"] + t["dram_bytes_write"])
    dom = max(alg_bytes, key=lambda n: kern_ms[n])
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                "frac": kernels[dom]["frac"], "traffic": traffic.get(dom), "algorithmic_bytes": int(alg_bytes[dom]),
                "traffic_source": "ncu --set full capture, profiles/ncu_traffic.json" if dom in traffic else None, "peak_source": peak_src,
                "frac_of_nominal_8TBs": kernels[dom]["achieved_gbs"] / 8000.0}

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
        v, dt, threads, sample = run_cpu(a, 2, 1)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample, "s_per_pass": dt}

    # ---- widened rows, outside the timed region and outside `value`: N3 rate() kernel ----------
    widened = None
    if rank == 0 and world == 1 and not a.no_widened:
        S5, T5, STEP5 = 80_000, 10_099, 15_000
        cn = ctx.counters_create(S5, T5)
        cn.generate(SEED, 0, STEP5, 0, 0)
        prm = native.Counters.params(STEP5, 20)
        pw = ctx.series_create(S5, cn.rate_len(prm), synth.even_node_offsets(S5 // 8, 8))
        for _ in range(3):
            cn.rate(prm, pw)
        acc = 0.0
        for _ in range(5):
            cn.rate(prm, pw)
            ctx.sync()
            acc += ctx.timing()["kernels"]["counter_rate"]
        nb = 8 * S5 * T5 + 4 * S5 * cn.rate_len(prm)
        widened = {"counter_rate": {"workload": f"rate(counter[5m]) at a 15 s step, {S5} series x {T5} float64 samples, no gaps", "ms": acc / 5,
                                    "algorithmic_bytes": nb, "achieved_gbs": nb / (acc / 5) / 1e6, "frac": nb / (acc / 5) / 1e6 / peak}}
        pw.destroy()
        cn.destroy()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64 (allocation) / f32 samples with f64 accumulation", "data": "synthetic",
            "config": {"workload": workload_name(a), "parallelism": f"node-range shards x{world}, NCCL allreduce of fleet totals",
                       "cache": f"inputs {4 * samples / 1e9:.1f} GB per GPU >> 126 MB L2 (no flush needed)", "seed": SEED},
            "roofline": roofline, "kernels": kernels, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clock_info, "widened": widened,
        }
        print(json.dumps(line))
    fleet.destroy()
    if series is not None:
        series.destroy()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    args = parse_args()
    sys.exit(main_reference(args) if args.impl == "reference" else main_gpu(args))
