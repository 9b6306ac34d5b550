"""Instructions executed and stall samples per CUDA source line of one kernel in an ncu report (needs -lineinfo and --import-source on).
usage: python scripts/ncu_by_line.py report.ncu-rep kernel_name_substring [top_n]"""
import collections
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern],
                     capture_output=True, text=True).stdout
inst, stall, text = collections.Counter(), collections.Counter(), {}
fname, cols, cur = "", None, None
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].rsplit("/", 1)[-1]
    elif r[0] == "Line No":
        cols = {n: i for i, n in enumerate(r) if n not in ("Source",)}
        cols["src"], cols["sass"] = 1, 3
    elif cols and r[0].strip().isdigit():  # a source line; its SASS rows follow with an empty line number
        cur = (fname, int(r[0]))
        text[cur] = r[1].strip()
        if len(r) > cols["Instructions Executed"] and r[cols["Instructions Executed"]]:
            pass
    elif cols and cur and len(r) > cols["Instructions Executed"] and r[2].startswith("0x"):
        inst[cur] += float(r[cols["Instructions Executed"]] or 0)
        stall[cur] += float(r[cols["Warp Stall Sampling (All Samples)"]] or 0)
ti, ts = sum(inst.values()) or 1, sum(stall.values()) or 1
print(f"total warp instructions {ti:.4g}, stall samples {ts:.4g}")
for k, v in inst.most_common(top):
    print(f"{v / ti * 100:5.1f}% inst {stall[k] / ts * 100:5.1f}% stall  {k[0]}:{k[1]:<5d} {text.get(k, '')[:110]}")
