"""Turn an .ncu-rep into the metric,unit,value summary kept under profiles/ (last captured launch).
usage: python tools/ncu_summary.py report.ncu-rep out.csv"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[-1]
keep = ("gpu__time_duration", "dram__bytes", "dram__throughput", "gpu__dram_throughput", "l1tex__throughput", "l1tex__data_bank_conflicts",
        "l1tex__data_pipe_lsu_wavefronts", "l1tex__t_sector_hit_rate", "lts__t_sector_hit_rate", "lts__throughput", "sm__throughput",
        "sm__pipe_fp64_cycles_active", "sm__inst_executed_pipe_fp64", "sm__warps_active", "smsp__issue_active", "smsp__inst_executed.sum",
        "smsp__average_warps_issue_stalled", "launch__", "sass__inst_executed_local", "sm__inst_executed_pipe_tmem", "smsp__sass_inst_executed_op_tmem")
with open(sys.argv[2], "w") as f:
    f.write("metric,unit,value\n")
    for h, u, v in zip(hdr, units, vals):
        if h.startswith(keep) and "Not Issued" not in h:
            f.write("%s,%s,%s\n" % (h, u, v))
