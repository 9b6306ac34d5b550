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

# the floor at this size: a plain device-to-device copy that moves the same number of bytes (read + written) as K1's
# algorithmic traffic — small transfers do not reach the 4 GB copy rate MEASURED_PEAKS.json quotes
if os.environ.get("FLOOR"):
    import torch

    total = cols.nbytes_in() + cols.n_pods * (2 + (4 + 8 * cols.n_resources) * 2)
    a = torch.empty(total // 2, dtype=torch.uint8, device="cuda")
    b = torch.empty_like(a)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for _ in range(5):
        b.copy_(a)
    ts = []
    for _ in range(20):
        ev[0].record()
        b.copy_(a)
        ev[1].record()
        torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    ts.sort()
    print(f"copy of {total / 1e6:.1f} MB of traffic: median {ts[10] * 1e3:.1f} us = {total / ts[10] / 1e6:.0f} GB/s, best {ts[0] * 1e3:.1f} us")
