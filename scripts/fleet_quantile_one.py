"""One C4-shaped fleet quantile, a few calls (ncu target).  usage: python scripts/fleet_quantile_one.py [nodes] [chips] [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from headlamp_intel_gpu_plugin_b200 import native, synth

nodes, chips, T = (int(a) for a in (sys.argv[1:4] + ["10000", "8", "86400"][len(sys.argv) - 1:]))
ctx = native.Context(0)
s = ctx.series_create(nodes * chips, T, synth.even_node_offsets(nodes, chips))
s.generate(synth.DEFAULT_SEED, 0, 0)
for _ in range(3):
    print(s.quantile_fleet(0.95), s.fleet_quantile_reads(), round(ctx.timing()["kernel_ms"], 3), "ms")
