#!/bin/bash
# Sweeps the fused thread-per-window kernel's ring depth / private-slot count on the GPU box.
for cfg in ${SWEEP:-1:17 1:20 1:24 1:32}; do
  set -- ${cfg/:/ }
  IGPU_WIDE_STAGES=$1 IGPU_WIDE_SLOTS=$2 timeout 200 python bench.py --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/bench_b.log 2>&1
  python - "$cfg" <<'PY'
import json, sys
l = [x for x in open("gpurun_out/bench_b.log") if x.startswith("{")]
if l:
    d = json.loads(l[-1]); k = d["kernels"]
    print(sys.argv[1], "step", round(d["ms_per_step"], 3), "fused", round(k["series_fused"]["ms"], 3), "frac", round(k["series_fused"]["frac"], 3),
          "pod", round(k["pod_rollup"]["ms"], 4), "sample", round(k["quantile_sample"]["ms"], 3), "finish", round(k["quantile_finish"]["ms"], 3))
else:
    print(sys.argv[1], open("gpurun_out/bench_b.log").read()[-600:])
PY
done
