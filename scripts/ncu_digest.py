"""Digest of one ncu report: headline metrics, opcode mix and the hottest SASS lines.
usage: python scripts/ncu_digest.py report.ncu-rep [iterations_for_per_iter_counts]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
iters = float(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units, r = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"]
for i, n in enumerate(h):
    if n in want or ("issue_stalled" in n and n.endswith("per_issue_active.ratio") and float(r[i] or 0) > 0.15):
        print(f"{n:90s} {units[i]:12s} {r[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, x in enumerate(rows) if "Source" in x][0]
h, body = rows[hi], rows[hi + 1:]
si = h.index("Source")
ie = [i for i, n in enumerate(h) if "Instructions Executed" in n][0]
ws = [i for i, n in enumerate(h) if "Warp Stall Sampling (All" in n][0]
body = [x for x in body if len(x) > ws]
tot = sum(float(x[ie] or 0) for x in body)
samp = sum(float(x[ws] or 0) for x in body)
op, ops = collections.Counter(), collections.Counter()
for x in body:
    m = x[si].split()
    k = (m[1] if m and m[0].startswith("@") else (m[0] if m else "")).split(".")[0]
    op[k] += float(x[ie] or 0)
    ops[k] += float(x[ws] or 0)
print(f"total inst {tot:.3g}" + (f"  = {tot / iters:.1f} per iteration" if iters else ""))
for k, v in op.most_common(16):
    print(f"  {k:8s} inst {v / tot * 100:5.1f}%" + (f" ({v / iters:5.1f}/iter)" if iters else "") + f"  stalls {ops[k] / samp * 100:5.1f}%")
print("hottest lines by stall samples:")
for x in sorted(body, key=lambda x: -float(x[ws] or 0))[:14]:
    print(f"  {float(x[ws] or 0) / samp * 100:5.1f}%  {x[si][:100]}")
