"""Summarise ncu output for profiles/: (1) a launch list CSV (--metrics gpu__time_duration.sum) -> per-kernel share of
one step; (2) a .ncu-rep (--set full) -> the handful of counters DESIGN.md cites, one row per captured launch.
usage: python tools/ncu_summary.py launches <launches.csv> <first_kernel_substr> > profiles/x.md
       python tools/ncu_summary.py rep <file.ncu-rep> > profiles/y.csv"""
import csv
import subprocess
import sys

KEEP = ["Kernel Name", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def short(name):
    return name.replace("<unnamed>::", "").replace("void ", "").split("(")[0]


def launches(path, first):
    raw = list(csv.DictReader(l for l in open(path) if not l.startswith("==")))
    # one row per launch: a capture with several --metrics has one CSV line per (launch, metric)
    rows, by_id = [], {}
    for r in raw:
        d = by_id.get(r["ID"])
        if d is None:
            d = by_id[r["ID"]] = {"Kernel Name": r["Kernel Name"], "Grid Size": r["Grid Size"]}
            rows.append(d)
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    dram = any("dram__bytes_read.sum" in r for r in rows)
    starts = [i for i, r in enumerate(rows) if first in r["Kernel Name"]]
    print(f"# launch list: {path} ({len(rows)} launches; steps start at `{first}`)\n")
    print("ncu per-launch times are cold-cache and serialised: compare SHARES, not absolutes.\n")
    for si, s in enumerate(starts):
        e = starts[si + 1] if si + 1 < len(starts) else len(rows)
        tot = sum(r["gpu__time_duration.sum"] for r in rows[s:e])
        extra = " DRAM read MB | DRAM write MB |" if dram else ""
        print(f"## step {si}: {e - s} launches, {tot / 1e3:.1f} us\n\n| kernel | grid | us | share |{extra}\n|---|---|---|---|" + ("---|---|" if dram else ""))
        for r in rows[s:e]:
            v = r["gpu__time_duration.sum"]
            x = f" {r['dram__bytes_read.sum'] / 1e6:.1f} | {r['dram__bytes_write.sum'] / 1e6:.1f} |" if dram else ""
            print(f"| {short(r['Kernel Name'])} | {r['Grid Size']} | {v / 1e3:.1f} | {100 * v / tot:.1f}% |{x}")
        print()


def rep(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = [hdr.index(k) for k in KEEP if k in hdr]
    w = csv.writer(sys.stdout)
    w.writerow([hdr[i] for i in ix])
    w.writerow([units[i] for i in ix])
    for d in data:
        w.writerow([short(d[i]) if hdr[i] == "Kernel Name" else d[i] for i in ix])


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        rep(sys.argv[2])
