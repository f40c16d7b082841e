"""Turns one `ncu --set full` capture of a 1024-frame step (+ optionally one of the matcher scan) into the committed
summaries the bench reads:

  profiles/<tag>_ncu_full.csv   headline metrics per kernel launch
  profiles/traffic.json         dram__bytes_read.sum + dram__bytes_write.sum per launch (bench.py `roofline.traffic`)
  profiles/instr.json           executed warp instructions per launch and per 1024-frame step (bench.py `roofline_issue`)

  python tools/ncu_to_profiles.py <tag> <extractor.ncu-rep> [<matcher.ncu-rep>]
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "launch__grid_size", "launch__block_size",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__cycles_active.avg"]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    res = []
    for r in rows[2:]:
        d = {}
        for h, u, v in zip(hdr, units, r):
            d[h] = (v, u)
        res.append(d)
    return res


def num(d, k):
    v, u = d.get(k, ("", ""))
    try:
        return float(v.replace(",", "")) * SCALE.get(u, 1)
    except ValueError:
        return None


def short(name):
    name = re.sub(r"\(.*", "", name).replace("void ", "").replace("orbfront::", "")
    return re.sub(r"<.*", "", name)


def main():
    tag = sys.argv[1]
    launches = []
    for p in sys.argv[2:]:
        launches += raw(p)
    lines = ["# ncu --set full --clock-control none, one launch per row (%s); extractor launches process 1024 frames of 640x480, top2_scan 1000 queries x 10 M rows" % tag,
             "kernel," + ",".join(KEYS)]
    traffic, instr = {}, {"per_launch": {}, "per_step": {}}
    n_resize = 0
    for d in launches:
        k = short(d["Kernel Name"][0])
        lines.append(k + "," + ",".join("%s %s" % d.get(x, ("", "")) for x in KEYS))
        t = (num(d, "dram__bytes_read.sum") or 0) + (num(d, "dram__bytes_write.sum") or 0)
        w = num(d, "smsp__inst_executed.sum") or 0
        name = {"fast_cells_kernel": "fast", "blur_kernel": "blur", "orient_describe_kernel": "describe", "quadtree_pyramid_kernel": "quadtree",
                "quadtree_select_kernel": "quadtree_sorting", "top2_scan_kernel": "top2_scan"}.get(k)
        if k.startswith("resize_level"):
            n_resize += 1
            instr["per_step"]["pyramid"] = instr["per_step"].get("pyramid", 0) + w
            traffic["pyramid_all_levels"] = traffic.get("pyramid_all_levels", 0) + t
            if n_resize == 1:
                traffic["pyramid_level1"] = t
                instr["per_launch"]["pyramid_level1"] = w
            continue
        if name and name not in traffic:
            traffic[name] = t
            instr["per_launch"][name] = w
            if name == "quadtree_sorting":  # the sorting kernel redoes the levels the sort-free kernel handed back: same stage
                instr["per_step"]["quadtree"] = instr["per_step"].get("quadtree", 0) + w
                traffic["quadtree"] = traffic.get("quadtree", 0) + t
            elif name != "top2_scan":
                instr["per_step"][name] = instr["per_step"].get(name, 0) + w
    instr["per_step"]["total"] = sum(v for k, v in instr["per_step"].items() if k != "total")
    instr["_note"] = "smsp__inst_executed.sum from profiles/%s_ncu_full.csv: executed warp instructions per launch, and per 1024-frame step (pyramid = its %d launches)" % (tag, n_resize)
    traffic["_note"] = "dram__bytes_read.sum + dram__bytes_write.sum per launch from profiles/%s_ncu_full.csv (ncu --set full, 1024-frame launches; pyramid_level1 = the level-1 launch)" % tag
    open(os.path.join(ROOT, "profiles", tag + "_ncu_full.csv"), "w").write("\n".join(lines) + "\n")
    json.dump(traffic, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    json.dump(instr, open(os.path.join(ROOT, "profiles", "instr.json"), "w"), indent=1)
    print(json.dumps(instr, indent=1))
    print(json.dumps(traffic, indent=1))


if __name__ == "__main__":
    main()
