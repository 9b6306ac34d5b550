"""C2-shaped allocation rollup on resident columns, a few calls (ncu target).  usage: python scripts/fleet_one.py [nodes] [pods]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from headlamp_intel_gpu_plugin_b200 import native, synth

nodes, pods = (int(a) for a in (sys.argv[1:3] + ["10000", "1000000"][len(sys.argv) - 1:]))
ctx = native.Context(0)
cols = synth.synth_fleet(nodes, pods)
if os.environ.get("GROUPED"):
    cols, _ = cols.grouped_by_node()
f = ctx.fleet_create(cols.n_nodes, cols.n_pods, cols.n_entries, cols.n_resources, cols.options)
f.upload(cols)
acc = 0.0
for i in range(8):
    f.rollup()
    ctx.sync()
    if i >= 3:
        acc += ctx.timing()["kernels"]["pod_rollup"]
print("pod_rollup ms", acc / 5, "bytes", cols.nbytes_in())
