"""Top CUDA source lines of an .ncu-rep by executed warp instructions and stall samples
(ncu --page source --csv --print-source cuda,sass; lines with a line number are aggregates)."""
import csv
import io
import os
import subprocess
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
kf = ["-k", "regex:" + os.environ["NCU_KERNEL"]] if os.environ.get("NCU_KERNEL") else []  # NCU_KERNEL=fast_cells: only that kernel
out = subprocess.run(["ncu", "-i", path] + kf + ["--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fname = "?"
data = []
hdr = None
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = os.path.basename(r[1]); continue
    if r[0] == "Line No":
        hdr = r
        ci = hdr.index("Instructions Executed"); cw = hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if hdr is None or not r[0].strip().isdigit():
        continue
    try:
        data.append((int(r[ci] or 0), int(r[cw] or 0), fname, r[0], r[1]))
    except ValueError:
        pass
tot = sum(d[0] for d in data) or 1
tots = sum(d[1] for d in data) or 1
print(f"total warp inst {tot}  total stall samples {tots}")
for n, s, f, ln, src in sorted(data, reverse=True)[:top]:
    print(f"{100*n/tot:5.1f}% inst {100*s/tots:5.1f}% stall  {f}:{ln:>4}: {src.strip()[:110]}")

if len(sys.argv) > 3:  # phase ranges: file:lo-hi=name,...
    phases = []
    for spec in sys.argv[3].split(","):
        rng, name = spec.split("=")
        f, r = rng.split(":")
        lo, hi = map(int, r.split("-"))
        phases.append((f, lo, hi, name))
    agg = {}
    for n, s, f, ln, src in data:
        name = "other"
        for pf, lo, hi, pn in phases:
            if f == pf and lo <= int(ln) <= hi:
                name = pn
                break
        a = agg.setdefault(name, [0, 0])
        a[0] += n; a[1] += s
    for k, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{k:14s} {100*n/tot:5.1f}% inst  {100*s/tots:5.1f}% stall")
