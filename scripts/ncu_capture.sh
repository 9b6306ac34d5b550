#!/bin/bash
# Captures for profiles/ (run on the GPU box, one GPU): `ncu --set full` of the kernels the bench times, at the bench's shapes,
# plus the launch list of one bench step.  Reports land in gpurun_out/; scripts/ncu_traffic.py digests them here.
# usage: bash scripts/ncu_capture.sh r02
R=${1:-r02}
NCU="ncu --set full --import-source on --clock-control none"
# C4: fused rollup (sample, wide, finish of the third call) — W = 60 p95, the bench's step
timeout 600 $NCU -k 'regex:series_rollup_wide|quantile_finish_kernel|quantile_sample' -s 6 -c 3 -o gpurun_out/${R}_c4 -f python scripts/rollup_one.py > gpurun_out/${R}_c4.log 2>&1
# the same at the median (the classification cannot be skipped) and at W = 300 (k threads per window)
Q=0.5 timeout 600 $NCU -k 'regex:series_rollup_wide' -s 2 -c 1 -o gpurun_out/${R}_c4_q50 -f python scripts/rollup_one.py > gpurun_out/${R}_c4_q50.log 2>&1
W=300 timeout 600 $NCU -k 'regex:series_rollup_wide' -s 2 -c 1 -o gpurun_out/${R}_c4_w300 -f python scripts/rollup_one.py > gpurun_out/${R}_c4_w300.log 2>&1
# C2: allocation rollup, pods grouped by node (the bench's layout)
GROUPED=1 timeout 300 $NCU -k 'regex:pod_rollup_ep' -s 4 -c 1 -o gpurun_out/${R}_k1 -f python scripts/fleet_one.py > gpurun_out/${R}_k1.log 2>&1
# launch list of the bench command (cold-cache, serialised times: shares of the step, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${R}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-widened > gpurun_out/${R}_launches.log 2>&1
ls -la gpurun_out/${R}_*
