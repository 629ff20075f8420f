"""Key raw counters of every launch in an .ncu-rep (issue/pipe utilisation, stall reasons, DRAM bytes).
usage: python tools/ncu_raw.py <file.ncu-rep>"""
import csv
import io
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, data = rows[0], rows[2:]
want = ["gpu__time_duration.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum",
        "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_fmaheavy.sum", "sm__inst_executed_pipe_fmalite.sum",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__cycles_active.avg", "sm__cycles_elapsed.max",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
want += [h for h in hdr if "issue_stalled" in h and "per_issue_active" in h and "not_issued" not in h]
for d in data:
    print("=====", d[hdr.index("Kernel Name")][:90])
    for w in want:
        if w in hdr:
            v = d[hdr.index(w)]
            try:
                if float(v) == 0: continue
            except ValueError:
                pass
            print(f"  {w.replace('smsp__average_warps_issue_stalled_', 'stall ').replace('_per_issue_active.ratio', '')}: {v}")
