"""GPU: bench.py on a tiny workload prints exactly one JSON line carrying every key of the measurement contract, for
both arms.  (Sizes are far below the BASELINE config, so the numbers mean nothing; the shape of the line is the test.)"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = ["--nodes", "200", "--pods", "20000", "--chips", "8", "--samples", "5760", "--cpu-series", "400", "--steps", "3", "--warmup", "3"]


def run(extra):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *SMALL, *extra], cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout[-1000:]
    return json.loads(lines[0])


@pytest.mark.gpu
def test_b200_arm_line():
    l = run(["--e2e-steps", "2"])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
              "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks"):
        assert k in l, k
    assert l["steps"] == 3 and l["warmup"] == 3 and l["n_gpus"] == 1 and l["scaling"] == "weak" and l["vs_baseline"] is None
    assert set(("bound", "achieved", "peak", "unit", "frac", "traffic")) <= set(l["roofline"]) and l["roofline"]["bound"] == "hbm"
    assert set(("value", "unit", "cores", "kind", "sample")) <= set(l["cpu_baseline"]) and l["cpu_baseline"]["kind"] == "port"
    assert set(("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step")) <= set(l["e2e"])
    assert l["e2e"]["h2d_bytes_per_step"] > 200 * 8 * 5760 * 4 and l["e2e"]["value"] < l["value"]
    assert l["gpu_launches"] >= 3 * 5 and "workload" in l["config"]
    assert set(("sm_mhz", "sm_max_mhz", "reasons")) <= set(l["clocks"])
    assert l["pods_per_s"] > 0 and l["samples_per_s"] > 0 and l["widened"]["series_fused_worst_case_q0.5"]["ms"] > 0
    # both arms describe the same workload with the same dict (the driver compares them)
    assert l["config"] == run(["--impl", "reference"])["config"]


@pytest.mark.gpu
def test_named_configs_and_allreduce_check_on_one_rank():
    """The N > 1 part of the line (BASELINE configs 3-5 + the exchange-step parity check), forced on at one rank and tiny sizes."""
    l = run(["--no-e2e", "--no-cpu-baseline", "--no-widened", "--named-configs", "on", "--c3-nodes", "300", "--c3-pods", "30000",
             "--c5-nodes-per-gpu", "20", "--c5-samples", "20000"])
    c = l["configs"]
    assert c["allreduce_check"] == "pass", c["allreduce_check_detail"]
    d = c["allreduce_check_detail"]
    for k in ("c3_totals_96_slots_vs_torch_allreduce", "c4_strong_fleet_vector_vs_torch_gather", "c5_fleet_vector_vs_torch_gather",
              "small_fleet_totals_96_slots_vs_oracle", "small_fleet_vector_vs_oracle", "fleet_quantile_same_on_every_rank",
              "small_fleet_quantile_vs_oracle", "small_fleet_window_planes_vs_oracle"):
        assert d[k] is True, k
    assert c["c3"]["pods_per_s"] > 0 and c["c4_strong"]["samples_per_s"] > 0 and c["c5"]["samples_per_s"] > 0


def test_reference_arm_line():
    l = run(["--impl", "reference"])
    assert l["impl"] == "reference" and l["steps"] == 3 and l["warmup"] == 3
    assert l["cpu_baseline"]["value"] == l["value"] == l["e2e"]["value"] and l["e2e"]["h2d_bytes_per_step"] == 0
    assert l["cpu_baseline"]["kind"] == "port" and l["cpu_baseline"]["cores"] >= 1
