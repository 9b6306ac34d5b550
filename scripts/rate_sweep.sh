#!/bin/bash
# sweeps the work-item length / grid size knobs of the rate kernel (env overrides read by igpu_counters_rate)
for seg in 16 32 64 128 400; do for grid in 3 6 12; do
  echo "seg=$seg grid=$grid: $(IGPU_RATE_SEG=$seg IGPU_RATE_GRID=$grid python scripts/rate_times.py 2>&1 | sed -n '1p;6p' | awk '{print $(NF-4), $NF}' | tr '\n' ' ')"
done; done
