"""One C4-shaped fused rollup, a few calls (ncu target).  usage: [W=60] [Q=0.95] [NAN_PPM=0] python scripts/rollup_one.py [nodes] [chips] [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from headlamp_intel_gpu_plugin_b200 import native, synth

nodes, chips, T = (int(a) for a in (sys.argv[1:4] + ["10000", "8", "86400"][len(sys.argv) - 1:]))
ctx = native.Context(0)
s = ctx.series_create(nodes * chips, T, synth.even_node_offsets(nodes, chips))
s.generate(synth.DEFAULT_SEED, 0, int(os.environ.get("NAN_PPM", "0")))
W, Q = int(os.environ.get("W", "60")), float(os.environ.get("Q", "0.95"))
for _ in range(4):
    s.rollup(W, Q, 0, fetch_windows=False)
print({k: round(v, 4) for k, v in ctx.timing()["kernels"].items() if v})
