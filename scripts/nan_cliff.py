import os, sys
sys.path.insert(0, ".")
from headlamp_intel_gpu_plugin_b200 import native, synth
ctx = native.Context(0)
s = ctx.series_create(80000, 86400, synth.even_node_offsets(10000, 8))
for ppm in (0, 1000, 20000):
    s.generate(synth.DEFAULT_SEED, 0, ppm)
    out = {}
    for W in (60, 300, 30):
        for _ in range(3):
            s.window_stats(W, fetch=False); ctx.sync()
        out[W] = round(ctx.timing()["kernels"]["window_stats"], 3)
    for _ in range(3):
        s.quantile_by_node(0.95)
    out["scan"] = round(ctx.timing()["kernels"]["quantile_scan"], 3)
    for _ in range(3):
        s.rollup(30, 0.95, 0, fetch_windows=False)
    out["fused30"] = round(ctx.timing()["kernels"]["series_fused"], 3)
    print(ppm, out)
