#!/bin/bash
# On the GPU box: f64 work and duration of the mapgen kernel (ncu after a plain run of the same command).
mkdir -p gpurun_out
M=gpu__time_duration.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__inst_executed.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_write.sum,dram__bytes_read.sum,launch__registers_per_thread
python scripts/mapgen_ncu.py > /dev/null 2>&1 &&
ncu --clock-control none --metrics $M -k regex:mapgen_kernel --csv --log-file gpurun_out/r02_mapgen_ncu.csv python scripts/mapgen_ncu.py > gpurun_out/r02_mapgen_ncu.log 2>&1
echo "exit $?"
python - <<P
import csv, collections
rows = list(csv.reader(open("gpurun_out/r02_mapgen_ncu.csv")))
h = next(r for r in rows if "Metric Name" in r)
m, v, u = h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
acc = collections.defaultdict(list)
for r in rows:
    if len(r) > v and r[0].isdigit():
        acc[r[m] + " [" + r[u] + "]"].append(float(r[v].replace(",", "")))
for k, x in acc.items():
    print("%-80s %18.1f (n=%d)" % (k, sum(x) / len(x), len(x)))
P
