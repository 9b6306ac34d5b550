"""One shape of igpu_counters_rate, a few launches (ncu target).  usage: python scripts/rate_one.py [S] [T] [K] [missing_ppm] [stride] [reset_period]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from headlamp_intel_gpu_plugin_b200 import native, synth
from headlamp_intel_gpu_plugin_b200.native import Counters

S, T, K, miss, stride, reset = (int(a) for a in (sys.argv[1:7] + ["40000", "10099", "20", "0", "1", "2000"][len(sys.argv) - 1:]))
ctx = native.Context(0)
c = ctx.counters_create(S, T)
c.generate(synth.DEFAULT_SEED, 0, 15_000, reset, miss)
prm = Counters.params(15_000, K, eval_stride=stride)
dst = ctx.series_create(S, c.rate_len(prm), synth.even_node_offsets(S // 8, 8))
for _ in range(4):
    c.rate(prm, dst)
ctx.sync()
print(ctx.timing()["kernels"]["counter_rate"], "ms")
