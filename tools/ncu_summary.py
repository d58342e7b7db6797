"""Condense `ncu -i X.ncu-rep --page raw --csv` into the metrics DESIGN.md / bench.py cite.
usage: ncu -i rep --page raw --csv | python tools/ncu_summary.py label0,label1,... > profiles/summary.csv"""
import csv, sys
KEEP = ["Kernel Name", "gpu__time_duration.sum", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
        "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__grid_size", "launch__block_size"]
rows = list(csv.reader(sys.stdin))
hdr, units, data = rows[0], rows[1], rows[2:]
labels = sys.argv[1].split(",") if len(sys.argv) > 1 else [f"launch{i}" for i in range(len(data))]
print("metric,unit," + ",".join(labels))
for k in KEEP:
    if k in hdr:
        i = hdr.index(k)
        print(",".join([k, units[i]] + [r[i].replace(",", "") for r in data]))
