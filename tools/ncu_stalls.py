"""Stall-reason ratios (warps stalled per issue) and key utilisations of .ncu-rep files."""
import csv, io, subprocess, sys
for path in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    d = dict(zip(r[0], r[2]))
    g = lambda k: d.get(k, "?")
    print("==", path.split("/")[-1], g("gpu__time_duration.sum"), "us  issue%", g("smsp__issue_active.avg.pct_of_peak_sustained_active")[:5],
          "warps%", g("sm__warps_active.avg.pct_of_peak_sustained_active")[:5], "regs", g("launch__registers_per_thread"),
          "l1tex%", g("l1tex__throughput.avg.pct_of_peak_sustained_active")[:5], "inst", g("smsp__inst_executed.sum"),
          "dramR", g("dram__bytes_read.sum"), "dramW", g("dram__bytes_write.sum"))
    st = {k.split("stalled_")[1].split("_per_issue")[0]: float(v) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled") and "per_issue_active" in k}
    print("   " + "  ".join(f"{k}={v:.2f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:8]))
