"""Per-source-line instruction counts and stall samples of one kernel from an ncu report (--import-source on):
  python tools/ncu_srcagg.py <report.ncu-rep> [kernel id, default 0] > summary.txt"""
import csv
import subprocess
import sys

rep = sys.argv[1]
kid = sys.argv[2] if len(sys.argv) > 2 else None
cmd = ["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"]
if kid is not None:
    cmd += ["--kernel-id", ":::" + kid] if not kid.isdigit() else []
rows = list(csv.reader(subprocess.run(cmd, capture_output=True, text=True).stdout.splitlines()))
cur, hdr, agg, kernel = None, None, {}, None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) == 2 and r[0] == "Function Name":
        if kernel is None:
            kernel = r[1]
        elif r[1] != kernel:
            break
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if len(r) < 8 or hdr is None or r[2] != "-":
        continue
    try:
        n, samp = int(r[7]), int(r[4])
    except ValueError:
        continue
    k = (cur, int(r[0]), r[1].strip()[:100])
    a = agg.setdefault(k, [0, 0])
    a[0] += n
    a[1] += samp
tot = sum(v[0] for v in agg.values()) or 1
tots = sum(v[1] for v in agg.values()) or 1
print("# kernel:", kernel)
print("# warp instructions executed (all lines):", tot, " stall samples:", tots)
print("# %inst %samples file:line source")
for k, v in sorted(agg.items(), key=lambda x: -x[1][0])[:60]:
    print("%5.1f %5.1f  %s:%d  %s" % (100.0 * v[0] / tot, 100.0 * v[1] / tots, k[0], k[1], k[2]))
