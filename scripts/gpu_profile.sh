#!/bin/bash
# ncu evidence on the GPU box: (1) launch list with device time of every kernel of a short bench run,
# (2) one --set full capture of the mesher kernel.  Each ncu run follows a plain run of the same command.
tag=${1:-r1}
mkdir -p gpurun_out
CMD="python bench.py --steps 6 --warmup 3 --e2e-steps 2"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_list.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:analyse_kernel\|expand_kernel -s 16 -c 4 -f -o gpurun_out/${tag}_prof $CMD > gpurun_out/${tag}_ncu_full.log 2>&1
echo "full capture exit $?"
ls -la gpurun_out/
