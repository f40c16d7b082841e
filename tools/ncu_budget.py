"""Instruction budget of every kernel in an .ncu-rep (captured with --import-source on, -lineinfo):
executed warp instructions split by SASS opcode class, and by CUDA source line range.

  python tools/ncu_budget.py rep.ncu-rep [--units kernel_regex=N ...] [--lines kernel_regex:file:lo-hi=name,...]

--units gives the number of work units (pixels, keypoints) one launch processes, so the table is
printed per unit (thread-instructions per unit = warp-instructions x average active threads).
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict, defaultdict

CLASSES = [
    ("LDS/STS", r"^(LDS|STS|LDSM|ATOMS)"),
    ("LDG/STG", r"^(LDG|STG|LD|ST|ATOMG|ATOM|RED|LDC|LDCU|ULDC|LDL|STL)\b"),
    ("SIMD(VABSDIFF/VIMNMX/IDP)", r"^(VABSDIFF4?|VIMNMX3?|IDP|IDP4A)"),
    ("SHFL/VOTE/MATCH", r"^(SHFL|VOTE|VOTEU|MATCH|REDUX)"),
    ("control", r"^(BRA|BSSY|BSYNC|EXIT|RET|CALL|WARPSYNC|BAR|NANOSLEEP|YIELD|BREAK|JMP|BRX|NOP|DEPBAR|ERRBAR|MEMBAR|S2R|S2UR|CS2R|BMOV|R2UR|UMOV|MOV|UIADD3|ULOP3|USHF|UIMAD|ULEA|USEL|UISETP|UFLO|UPOPC|UPRMT|R2P|P2R|PLOP3|UPLOP3|ISETP|SEL|FSEL)"),
    ("int ALU", r"^(IADD3?|IADD|IMAD|LOP3|LOP|SHF|LEA|PRMT|POPC|FLO|IMNMX|IABS|BREV|SGXT|I2I|IMUL|ISCADD|SHL|SHR|BMSK|LEPC)"),
    ("float/fp64", r"^(F|D|MUFU|I2F|F2I|I2FP|F2FP|HADD|HMUL|HFMA)"),
]


def classify(op):
    op = op.split(".")[0] if not op.startswith("IMAD.MOV") else "MOV"
    for name, pat in CLASSES:
        if re.match(pat, op):
            return name
    return "other"


def sass_page(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    kernels, cur, hdr = [], None, None
    for r in csv.reader(io.StringIO(out)):
        if not r:
            continue
        if r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            kernels.append(cur)
            hdr = None
            continue
        if r[0] == "Address":
            hdr = r
            continue
        if cur is None or hdr is None:
            continue
        cur["rows"].append(dict(zip(hdr, r)))
    return kernels


def main():
    path = sys.argv[1]
    units = {}
    for a in sys.argv[2:]:
        if a.startswith("--units"):
            continue
        if "=" in a and ":" not in a:
            k, v = a.split("=")
            units[k] = float(v)
    seen = OrderedDict()
    for k in sass_page(path):
        tot_w = tot_t = tot_p = 0
        by = defaultdict(lambda: [0, 0, 0])
        ops = defaultdict(int)
        for row in k["rows"]:
            try:
                w = int(row["Instructions Executed"] or 0)
                t = int(row["Thread Instructions Executed"] or 0)
                p = int(row["Predicated-On Thread Instructions Executed"] or 0)
            except ValueError:
                continue
            src = row["Source"].strip()
            m = re.match(r"^(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
            op = m.group(2) if m else "?"
            c = classify(op)
            by[c][0] += w; by[c][1] += t; by[c][2] += p
            ops[op.split(".")[0]] += w
            tot_w += w; tot_t += t; tot_p += p
        short = re.sub(r"\(.*", "", k["name"]).replace("void orbfront::", "").replace("orbfront::", "")
        key = (short, tot_w)
        if key in seen:
            continue
        seen[key] = 1
        n = None
        for pat, v in units.items():
            if re.search(pat, k["name"]):
                n = v
        print(f"\n== {short}: {tot_w/1e6:.1f} M warp-instr, {tot_t/1e6:.1f} M thread-instr "
              f"({tot_t/max(tot_w,1):.1f} threads/instr), predicated-off {100*(tot_t-tot_p)/max(tot_t,1):.1f} %"
              + (f"; per unit: {tot_w/n:.3f} warp-instr, {tot_t/n:.1f} thread-instr" if n else ""))
        for c, (w, t, p) in sorted(by.items(), key=lambda kv: -kv[1][0]):
            line = f"   {c:28s} {100*w/max(tot_w,1):5.1f} %  {w/1e6:9.1f} M"
            if n:
                line += f"  {t/n:7.2f} thread-instr/unit"
            print(line)
        top = sorted(ops.items(), key=lambda kv: -kv[1])[:14]
        print("   top opcodes: " + ", ".join(f"{o} {100*w/max(tot_w,1):.1f}%" for o, w in top))


if __name__ == "__main__":
    main()
