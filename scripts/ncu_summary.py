"""Turns an `ncu --set full` report of one K1 launch into the files bench.py and the judge read:
    python scripts/ncu_summary.py gpurun_out/prof_valid.ncu-rep profiles/r02 [states_per_launch]
writes <dir>/ncu_full_k_valid_dfs.raw.csv (the raw page) and <dir>/k1_ncu_summary.json (per-launch counters)."""
import csv
import io
import json
import os
import subprocess
import sys

rep, outdir = sys.argv[1], sys.argv[2]
states = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
os.makedirs(outdir, exist_ok=True)
with open(os.path.join(outdir, "ncu_full_k_valid_dfs.raw.csv"), "w") as f:
    f.write(raw)
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
col = {h: i for i, h in enumerate(hdr)}


def get(name, scale=True):
    v = float(vals[col[name]].replace(",", ""))
    u = units[col[name]]
    if scale:
        v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(u, 1.0)
    return v


out = {
    "kernel": vals[col["Kernel Name"]], "states_per_launch": states,
    "duration_s_under_ncu": get("gpu__time_duration.sum"),
    "dram_bytes_per_launch": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
    "dram_bytes_read": get("dram__bytes_read.sum"), "dram_bytes_write": get("dram__bytes_write.sum"),
    "warp_instructions_per_launch": get("smsp__inst_executed.sum"),
    "threads_per_instruction": get("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "branch_uniformity_pct": get("smsp__sass_average_branch_targets_threads_uniform.pct"),
    "registers_per_thread": get("launch__registers_per_thread"),
    "warps_active_per_sm": get("sm__warps_active.avg.per_cycle_active"),
    "l1_hit_pct": get("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": get("lts__t_sector_hit_rate.pct"),
    "sm_throughput_pct": get("sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    "sm_count": 148,
}
for k in ("smsp__issue_active.avg.pct", "smsp__inst_executed.avg.per_cycle_active", "sm__inst_executed.avg.per_cycle_elapsed",
          "dram__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
          "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
          "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"):
    if k in col:
        out[k] = get(k, scale=False)
with open(os.path.join(outdir, "k1_ncu_summary.json"), "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps(out, indent=1))
