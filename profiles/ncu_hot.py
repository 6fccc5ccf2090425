"""Executed-instruction mix and top stall sites from `ncu --page source --csv`
(python profiles/ncu_hot.py report.ncu-rep)."""
import csv, io, re, subprocess, sys
from collections import Counter
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ci = {h: i for i, h in enumerate(hdr)}
mix, samples = Counter(), []
total = 0
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    n = int(r[ci["Instructions Executed"]] or 0)
    op = re.sub(r"^@!?U?P\w+\s+", "", r[ci["Source"]].strip()).split()[0].split(".")[0]
    mix[op] += n
    total += n
    samples.append((int(r[ci["Warp Stall Sampling (All Samples)"]] or 0), n, r[ci["Source"]].strip()[:90]))
print("warp instructions executed:", total)
for op, n in mix.most_common(22):
    print(f"  {op:10s} {n:12d} {100.0 * n / total:5.1f}%")
print("top stall sites (samples, executed, SASS):")
for s in sorted(samples, reverse=True)[:14]:
    print("  ", s)
