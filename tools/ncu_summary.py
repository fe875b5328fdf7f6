"""Summarise an .ncu-rep (read with `ncu -i … --page raw --csv`) into a markdown table.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep "title" > profiles/xyz.md
"""
import csv
import io
import subprocess
import sys

rep, title = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
keys = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"]
keys = [k for k in keys if k in hdr]
units = rows[1]
print(f"# {title}\n")
print(f"Source: `{rep}` (ncu --set full --clock-control none). One row per captured launch.\n")
print("| " + " | ".join(k.split(".TriageCompute.")[-1] + (f" [{units[hdr.index(k)]}]" if units[hdr.index(k)] else "") for k in keys) + " |")
print("|" + "---|" * len(keys))
for r in rows[2:]:
    print("| " + " | ".join(r[hdr.index(k)][:70].replace("|", "/") for k in keys) + " |")
