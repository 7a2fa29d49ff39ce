#!/bin/bash
# Summarise gpurun_out/step_full.ncu-rep into profiles/ (tracked): key raw metrics per kernel + top stall instructions.
set -eu
REP=${1:-gpurun_out/step_full.ncu-rep}
ncu -i "$REP" --page raw --csv > /tmp/ncu_raw.csv
python - <<'PY'
import csv
rows = list(csv.reader(open("/tmp/ncu_raw.csv")))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__cycles_active.avg",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__cycles_elapsed.max", "smsp__inst_executed.avg.per_cycle_active",
        "sm__inst_executed_pipe_uniform.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_membar_per_warp_active.pct"]
idx = {h: i for i, h in enumerate(hdr)}
with open("profiles/ncu_step_full_summary.md", "w") as f:
    f.write("# ncu --set full, one capture per kernel of the flagship step (eager launches, --clock-control none)\n\n")
    f.write("Durations here include ncu's cache flush / cold L2 and are NOT bench numbers; in-graph kernel times are in profiles/README.md.\n\n")
    for r in data:
        f.write(f"## {r[idx['Kernel Name']][:90]}\n\n| metric | value | unit |\n|---|---|---|\n")
        for w in want[1:]:
            if w in idx:
                f.write(f"| {w} | {r[idx[w]]} | {units[idx[w]]} |\n")
        tens = [h for h in hdr if "tensor" in h or "tmem" in h or "utc" in h.lower()]
        for h in tens[:12]:
            if h not in want:
                f.write(f"| {h} | {r[idx[h]]} | {units[idx[h]]} |\n")
        f.write("\n")
print("wrote profiles/ncu_step_full_summary.md")
PY
ncu -i "$REP" --page source --csv > /tmp/ncu_src.csv 2>/dev/null || true
python tools/ncu_top_stalls.py /tmp/ncu_src.csv 12 > profiles/ncu_step_full_top_stalls.txt || true
wc -l profiles/ncu_step_full_summary.md profiles/ncu_step_full_top_stalls.txt
