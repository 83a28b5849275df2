"""Key per-kernel numbers from an .ncu-rep (ncu --set full): time, warp instructions, pipe use, occupancy, DRAM bytes, top stalls."""
import csv
import subprocess
import sys

rep = sys.argv[1]
rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
hdr = rows[0]
keys = {
    "us": "gpu__time_duration.sum",
    "warp_inst": "smsp__inst_executed.sum",
    "alu%": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "fma%": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "fmaheavy%": "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
    "fp16%": "sm__inst_executed_pipe_fp16.avg.pct_of_peak_sustained_active",
    "lsu%": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "issue%": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "warps%": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "regs": "launch__registers_per_thread",
    "dram_rd": "dram__bytes_read.sum",
    "dram_wr": "dram__bytes_write.sum",
    "bank_conf": "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
}
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print(d["Kernel Name"][:60], {k: d.get(v) for k, v in keys.items()})
    st = sorted(((float(d[h] or 0), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
                 for h in hdr if "smsp__average_warps_issue_stalled" in h and "per_issue_active" in h and "not_issued" not in h), reverse=True)[:7]
    print("   stalls:", ", ".join(f"{n} {v:.2f}" for v, n in st))
