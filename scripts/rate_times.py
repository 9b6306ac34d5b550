"""Times igpu_counters_rate (csrc/rate.cu) on device-generated counters.  Algorithmic bytes: 8 B read per grid
sample + 4 B written per evaluation.  usage: python scripts/rate_times.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from headlamp_intel_gpu_plugin_b200 import native, synth
from headlamp_intel_gpu_plugin_b200.native import Counters

ctx = native.Context(0)
PEAK = 6538.3
STEP = 15_000


def run(label, S, T, K, stride=1, reset_period=0, missing=0, env=None):
    c = ctx.counters_create(S, T)
    c.generate(synth.DEFAULT_SEED, 0, STEP, reset_period, missing)
    prm = Counters.params(STEP, K, eval_stride=stride)
    n_out = c.rate_len(prm)
    dst = ctx.series_create(S, n_out, synth.even_node_offsets(S // 8, 8))
    for k, v in (env or {}).items():
        os.environ[k] = v
    ctx.reload_env()  # the switches are read when a context is created
    for _ in range(3):
        c.rate(prm, dst)
    ctx.sync()
    acc, reps = 0.0, 5
    for _ in range(reps):
        c.rate(prm, dst)
        ctx.sync()
        acc += ctx.timing()["kernels"]["counter_rate"]
    for k in env or {}:
        del os.environ[k]
    ctx.reload_env()
    ms = acc / reps
    nbytes = 8 * S * T + 4 * S * n_out
    print(f"{label:34s} S={S} T={T} K={K} stride={stride}: {ms:.3f} ms  {nbytes / ms / 1e6:7.1f} GB/s  frac {nbytes / ms / 1e6 / PEAK:.3f}", flush=True)
    dst.destroy()
    c.destroy()


run("clean, 5m range @15s", 80_000, 10_099, 20)
run("resets every 2000", 80_000, 10_099, 20, reset_period=2000)
run("0.1% missing + resets", 80_000, 10_099, 20, reset_period=2000, missing=1000)
run("0.1% missing, gap kernel forced", 80_000, 10_099, 20, reset_period=2000, missing=1000, env={"IGPU_RATE_TIER": "2"})
run("0.3% missing + resets", 80_000, 10_099, 20, reset_period=2000, missing=3000)
run("0.3% missing, streaming forced", 80_000, 10_099, 20, reset_period=2000, missing=3000, env={"IGPU_RATE_TIER": "1"})
run("2% missing + resets", 80_000, 10_099, 20, reset_period=2000, missing=20_000)
run("2% missing, streaming forced", 80_000, 10_099, 20, reset_period=2000, missing=20_000, env={"IGPU_RATE_TIER": "1"})
run("2% missing, no resets", 80_000, 10_099, 20, missing=20_000)
run("10% missing + resets", 80_000, 10_099, 20, reset_period=2000, missing=100_000)
run("clean, gap kernel forced", 80_000, 10_099, 20, env={"IGPU_RATE_TIER": "2"})
run("clean, streaming forced", 80_000, 10_099, 20, env={"IGPU_RATE_TIER": "1"})
run("stride 4 (1m step)", 80_000, 10_099, 20, stride=4)
run("long rows", 8_000, 100_019, 20)
run("5m range @5s: K=60, long-range kernel", 80_000, 10_099, 60)
run("1h range @15s: K=240, long-range kernel", 80_000, 10_099, 240)
run("K=240 + resets every 2000", 80_000, 10_099, 240, reset_period=2000)
run("K=60 + 0.1% missing + resets", 80_000, 10_099, 60, reset_period=2000, missing=1000)
run("K=60 + 2% missing + resets", 80_000, 10_099, 60, reset_period=2000, missing=20_000)
run("K=240 + 2% missing + resets", 80_000, 10_099, 240, reset_period=2000, missing=20_000)
run("generic kernel (forced)", 20_000, 10_099, 20, env={"IGPU_RATE_FORCE_GENERIC": "1"})
run("K=80 generic (forced)", 20_000, 10_099, 80, env={"IGPU_RATE_FORCE_GENERIC": "1"})
run("5m range @1s: K=300 (first sample from global memory)", 80_000, 10_099, 300)
run("K=300 generic (forced)", 20_000, 10_099, 300, env={"IGPU_RATE_FORCE_GENERIC": "1"})
