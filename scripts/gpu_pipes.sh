#!/bin/bash
# On the GPU box: which execution pipes the two kernels load (per-kernel capture of steady-state steps, caches not
# flushed; instruction counts do not depend on the kernels overlapping).  Follows a plain run of the same command.
tag=${1:-r02}
mkdir -p gpurun_out
M=gpu__time_duration.sum,sm__inst_executed.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_uniform.sum,sm__inst_executed_pipe_cbu.sum,sm__inst_executed_pipe_adu.sum,smsp__inst_executed_op_shared_st.sum,smsp__inst_executed_op_shared_ld.sum,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,sm__cycles_active.avg
CMD="python scripts/ncu_steady.py"
$CMD > /dev/null 2> gpurun_out/${tag}_pipes_plain.err &&
ncu --cache-control none --clock-control none --metrics $M --profile-from-start off -c 8 --csv --log-file gpurun_out/${tag}_pipes_kernels.csv $CMD > gpurun_out/${tag}_pipes_kernels.log 2>&1
echo "pipes capture exit $?"; tail -3 gpurun_out/${tag}_pipes_kernels.log
python - <<P
import csv, collections
rows = list(csv.reader(open("gpurun_out/${tag}_pipes_kernels.csv")))
h = next(r for r in rows if "Metric Name" in r)
k, m, v = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
acc = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows:
    if len(r) > v and r[0].isdigit():
        acc[r[k][:40]][r[m]].append(float(r[v].replace(",", "")))
for kn, d in acc.items():
    print(kn)
    for mn, vals in d.items():
        print("   %-70s %16.1f  (n=%d)" % (mn, sum(vals) / len(vals), len(vals)))
P
