import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
from headlamp_intel_gpu_plugin_b200 import native, synth
ctx = native.Context(0)
for nodes in (6100, 6300):
    s = ctx.series_create(nodes*8, 86400, synth.even_node_offsets(nodes, 8))
    s.generate(1, 0, 0); ctx.sync()
    for what in ("quantile", "window"):
        try:
            if what == "window":
                s.window_stats(60, fetch=False); ctx.sync()
            else:
                s.quantile_by_node(0.95)
            print(nodes, what, "ok", flush=True)
        except Exception as e:
            print(nodes, what, "FAIL", str(e)[:200], flush=True); sys.exit(0)
    s.destroy()
