"""Host<->device transfer / launch-overhead crossover (BASELINE.json north star): at which fleet size does one call
through the C ABI with HOST buffers (upload + kernels + download inside the call) beat the CPU oracle port?
Writes one JSON document to stdout.  usage: python scripts/crossover.py > profiles/crossover_rNN.json"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import native, synth
from oracle import columnar as oracle

ctx = native.Context(0)
THREADS = oracle.num_threads()


def best(fn, reps=5):
    fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append(time.perf_counter() - t0)
    return min(t) * 1e3


alloc = []
for nodes, pods in [(50, 2_000), (200, 8_000), (1_000, 40_000), (3_000, 200_000), (10_000, 1_000_000)]:
    cols = synth.synth_fleet(nodes, pods)
    out = K.FleetOutputs.allocate(cols.n_nodes, cols.n_pods, cols.n_resources, cols.options)
    pin_c, pin_o = native.pinned_copy_of_columns(cols), native.pinned_outputs(cols.n_nodes, cols.n_pods, cols.n_resources, cols.options)
    row = {"nodes": nodes, "pods": pods, "cpu_1_thread_ms": best(lambda: oracle.alloc_rollup(cols, 1)),
           f"cpu_{THREADS}_threads_ms": best(lambda: oracle.alloc_rollup(cols, THREADS)),
           "gpu_call_pageable_ms": best(lambda: ctx.alloc_rollup(cols, out)),
           "gpu_call_pinned_ms": best(lambda: ctx.alloc_rollup(pin_c, pin_o))}
    ctx.alloc_rollup(pin_c, pin_o)
    t = ctx.timing()
    row["gpu_kernels_ms"] = t["kernel_ms"]
    row["h2d_bytes"], row["d2h_bytes"] = cols.nbytes_in(), out.nbytes_out()
    alloc.append(row)

series = []
for nodes, chips, T in [(10, 4, 5_760), (50, 4, 5_760), (200, 8, 5_760), (1_000, 8, 5_760), (1_000, 8, 86_400)]:
    n = nodes * chips
    x = oracle.generate_series(n, T, synth.DEFAULT_SEED, threads=THREADS)  # same generator, C++ twin (fast)
    off = synth.even_node_offsets(nodes, chips)
    pb = native.PinnedBuffer(x.nbytes)
    px = pb.view(np.float32, x.shape)
    px[...] = x
    s = ctx.series_create(n, T, off)

    def gpu(src):
        s.upload_async(src)
        return s.rollup(60, 0.95, 0, fetch_windows=True)

    def cpu(threads):
        oracle.window_stats(x, 60, threads=threads)
        oracle.quantile_by_node(x, off, 0.95, threads=threads)

    row = {"nodes": nodes, "series": n, "samples": n * T, "cpu_1_thread_ms": best(lambda: cpu(1), 3),
           f"cpu_{THREADS}_threads_ms": best(lambda: cpu(THREADS), 3), "gpu_call_pageable_ms": best(lambda: gpu(x), 3),
           "gpu_call_pinned_ms": best(lambda: gpu(px), 3), "h2d_bytes": int(x.nbytes)}
    series.append(row)
    s.destroy()
    pb.free()

print(json.dumps({"cpu_threads": THREADS, "note": "best-of-N wall time of one call, host buffers in and out; CPU = oracle port (OpenMP)",
                  "alloc_rollup": alloc, "series_rollup_W60_p95": series}, indent=1))
