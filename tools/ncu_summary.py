"""Condenses an ncu report into the JSON summaries kept under profiles/:
    python tools/ncu_summary.py gpurun_out/x.ncu-rep "source command line" "workload" > profiles/rNN_x_ncu_summary.json"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum"]

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
kernels = []
for r in rows[2:]:
    k = {"Kernel Name": r[hdr.index("Kernel Name")]}
    for name in KEEP:
        if name in hdr:
            i = hdr.index(name)
            k[name] = f"{r[i]} {units[i]}".strip()
    kernels.append(k)
print(json.dumps({"source": sys.argv[2] if len(sys.argv) > 2 else "", "workload": sys.argv[3] if len(sys.argv) > 3 else "",
                  "kernels": kernels}, indent=1))
