#!/bin/bash
# On the GPU box: the steady-state DRAM capture (scripts/ncu_steady.py).  Every ncu run follows a plain run of the same command.
tag=${1:-r02}
mkdir -p gpurun_out
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum,gpu__time_duration.sum
CMD="python scripts/ncu_steady.py"
$CMD > gpurun_out/${tag}_steady_plain.json 2> gpurun_out/${tag}_steady_plain.err &&
ncu --cache-control none --clock-control none --metrics $M --profile-from-start off -c 24 --csv --log-file gpurun_out/${tag}_steady_kernels.csv $CMD > gpurun_out/${tag}_steady_kernels.log 2>&1
echo "kernel capture exit $?"; cat gpurun_out/${tag}_steady_plain.json
$CMD > /dev/null 2>&1 &&
ncu --replay-mode range --cache-control none --clock-control none --metrics $M --csv --log-file gpurun_out/${tag}_steady_range.csv $CMD > gpurun_out/${tag}_steady_range.log 2>&1
echo "range capture exit $?"; tail -3 gpurun_out/${tag}_steady_range.log; tail -5 gpurun_out/${tag}_steady_range.csv | cut -c1-300
