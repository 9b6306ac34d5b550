"""Times the stand-alone series kernels on device-generated slabs (CUDA events inside the library).
usage: python scripts/kernel_times.py  -> one line per (shape, call)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from headlamp_intel_gpu_plugin_b200 import native, synth

ctx = native.Context(0)
PEAK = 6538.3


def run(name, n_nodes, chips, T, calls):
    s = ctx.series_create(n_nodes * chips, T, synth.even_node_offsets(n_nodes, chips))
    s.generate(synth.DEFAULT_SEED, 0, 0)
    S = n_nodes * chips * T
    for label, fn, kernels, nbytes in calls(s, S, n_nodes, chips, T):
        for _ in range(3):
            fn()
        acc = {k: 0.0 for k in kernels}
        reps = 5
        for _ in range(reps):
            fn()
            t = ctx.timing()["kernels"]
            for k in kernels:
                acc[k] += t[k]
        main = kernels[0]
        ms = acc[main] / reps
        others = " ".join(f"{k}={acc[k] / reps:.3f}" for k in kernels[1:])
        print(f"{name:10s} {label:22s} {main}={ms:.3f} ms  {nbytes / ms / 1e6:7.1f} GB/s  frac {nbytes / ms / 1e6 / PEAK:.3f}  {others}", flush=True)
    s.destroy()


def c4_calls(s, S, n_nodes, chips, T):
    nw = (T + 59) // 60
    yield "window_stats W=60", lambda: (s.window_stats(60, fetch=False), ctx.sync()), ["window_stats"], 4 * S + 12 * n_nodes * chips * nw
    yield "quantile p95", lambda: s.quantile_by_node(0.95), ["quantile_scan", "quantile_sample", "quantile_finish", "quantile_tail"], 4 * S + 4 * n_nodes
    yield "rollup W=60 p95", lambda: s.rollup(60, 0.95, 0, fetch_windows=False), ["series_fused", "quantile_sample", "quantile_finish", "quantile_tail"], 4 * S + 12 * n_nodes * chips * nw + 4 * n_nodes
    yield "rollup W=20 p95", lambda: s.rollup(20, 0.95, 0, fetch_windows=False), ["series_fused", "quantile_sample", "quantile_finish"], 4 * S + 12 * n_nodes * chips * ((T + 19) // 20) + 4 * n_nodes
    for W in (300, 1440, 120, 3000):
        yield f"rollup W={W} p95", (lambda W=W: s.rollup(W, 0.95, 0, fetch_windows=False)), ["series_fused", "quantile_sample", "quantile_finish", "quantile_tail"], 4 * S + 12 * n_nodes * chips * ((T + W - 1) // W) + 4 * n_nodes
    yield "rollup W=60 p50", lambda: s.rollup(60, 0.5, 0, fetch_windows=False), ["series_fused", "quantile_sample", "quantile_finish", "quantile_tail"], 4 * S + 12 * n_nodes * chips * nw + 4 * n_nodes
    yield "rollup W=300 p50", lambda: s.rollup(300, 0.5, 0, fetch_windows=False), ["series_fused", "quantile_sample", "quantile_finish", "quantile_tail"], 4 * S + 12 * n_nodes * chips * ((T + 299) // 300) + 4 * n_nodes
    yield "window_stats W=300", lambda: (s.window_stats(300, fetch=False), ctx.sync()), ["window_stats"], 4 * S + 12 * n_nodes * chips * ((T + 299) // 300)
    yield "window_stats W=1440", lambda: (s.window_stats(1440, fetch=False), ctx.sync()), ["window_stats"], 4 * S + 12 * n_nodes * chips * ((T + 1439) // 1440)
    yield "window_stats W=256", lambda: (s.window_stats(256, fetch=False), ctx.sync()), ["window_stats"], 4 * S + 12 * n_nodes * chips * ((T + 255) // 256)
    yield "window_stats W=1024", lambda: (s.window_stats(1024, fetch=False), ctx.sync()), ["window_stats"], 4 * S + 12 * n_nodes * chips * ((T + 1023) // 1024)
    for W in (30, 10, 90, 150):  # windows that are not a multiple of 4 samples
        yield f"window_stats W={W}", (lambda W=W: (s.window_stats(W, fetch=False), ctx.sync())), ["window_stats"], 4 * S + 12 * n_nodes * chips * ((T + W - 1) // W)
        yield f"rollup W={W} p95", (lambda W=W: s.rollup(W, 0.95, 0, fetch_windows=False)), ["series_fused", "window_stats", "quantile_scan", "quantile_sample", "quantile_finish", "quantile_tail"], 4 * S + 12 * n_nodes * chips * ((T + W - 1) // W) + 4 * n_nodes
    yield "window_stats W=64", lambda: (s.window_stats(64, fetch=False), ctx.sync()), ["window_stats"], 4 * S + 12 * n_nodes * chips * ((T + 63) // 64)


def c5_calls(s, S, n_nodes, chips, T):
    yield "quantile p95", lambda: s.quantile_by_node(0.95), ["quantile_scan", "quantile_sample", "quantile_finish", "quantile_tail"], 4 * S + 4 * n_nodes


def wall(label, fn, nbytes, reps=3):
    import time
    fn()
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    ctx.sync()
    ms = (time.perf_counter() - t0) / reps * 1e3
    print(f"{'C4':10s} {label:22s} wall={ms:.3f} ms  {nbytes / ms / 1e6:7.1f} GB/s  frac {nbytes / ms / 1e6 / PEAK:.3f}", flush=True)


def fleet_views():
    s = ctx.series_create(80_000, 86_400, synth.even_node_offsets(10_000, 8))
    s.generate(synth.DEFAULT_SEED, 0, 0)
    S = 80_000 * 86_400
    wall("quantile_fleet p95 (1 read)", lambda: s.quantile_fleet(0.95), 4 * S)
    assert s.fleet_quantile_reads() == 1
    wall("window_rollup W=60 fleet", lambda: s.window_rollup(60, nodes=False), 4 * S)
    wall("window_rollup W=300 fleet", lambda: s.window_rollup(300, nodes=False), 4 * S)
    # with the counts kept by the rollup: what keeping them costs the fused kernel, and the window rollup without a read
    s.keep_window_counts(True)
    for W in (60, 300):
        acc = 0.0
        for i in range(6):
            s.rollup(W, 0.95, 0, fetch_windows=False)
            acc += ctx.timing()["kernels"]["series_fused"] if i else 0.0
        print(f"{'C4':10s} rollup W={W} keeping counts: series_fused={acc / 5:.3f} ms", flush=True)
        wall(f"window_rollup W={W} fleet, from planes", lambda W=W: s.window_rollup(W, nodes=False), 4 * S)
        assert s.last_window_rollup_slab_reads == 0
        wall(f"window_rollup W={W} nodes+fleet, from planes", lambda W=W: s.window_rollup(W), 4 * S)
    s.destroy()


if os.environ.get("ONLY") != "c5":
    run("C4", 10_000, 8, 86_400, c4_calls)
    fleet_views()
run("C5-shard", 1_250, 8, 1_250_000, c5_calls)
