"""Per-source-line instruction counts / stall samples of one kernel from an .ncu-rep captured with --import-source on
(SASS rows summed per CUDA source line, so inlined helpers are attributed to their own file:line).
usage: python tools/ncu_lines.py <file.ncu-rep> <kernel regex> [min_pct] [kernel index]"""
import csv
import io
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + rx],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
kernels, cur = [], None          # a kernel = list of file sections
for r in rows:
    if r and r[0] == "File Path":
        cur = {"file": r[1], "rows": []}
    elif r and r[0] == "Function Name":
        if not kernels or kernels[-1]["name"] != r[1] or any(s["file"] == cur["file"] for s in kernels[-1]["secs"]):
            if not kernels or kernels[-1]["name"] != r[1]:
                kernels.append({"name": r[1], "secs": []})
        kernels[-1]["secs"].append(cur)
    elif r and r[0] == "Line No":
        cur["hdr"] = r
    elif cur is not None and "hdr" in cur and len(r) == len(cur["hdr"]):
        cur["rows"].append(r)
k = kernels[which]
print("#", k["name"][:120])
lines = []
for s in k["secs"]:
    h = s["hdr"]; ix = h.index("Instructions Executed"); sm = h.index("# Samples")
    line = None
    for r in s["rows"]:
        if r[2] == "-":
            line = [s["file"].split("/")[-1], r[0], r[1].strip(), 0, 0, 0]; lines.append(line)
        elif line is not None:
            line[3] += int(r[ix]) if r[ix].isdigit() else 0
            line[4] += int(r[sm]) if r[sm].isdigit() else 0
            line[5] += 1
tot = sum(l[3] for l in lines); smp = sum(l[4] for l in lines)
print(f"total warp instructions {tot}, samples {smp}")
for l in lines:
    if l[3] > tot * minpct / 100 or l[4] > smp * minpct / 100:
        print(f"{l[3] / tot * 100:5.1f}% inst {l[4] / max(smp, 1) * 100:5.1f}% smp {l[5]:4d} sass | {l[0]}:{l[1]:>4} {l[2][:110]}")
