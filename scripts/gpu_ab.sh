#!/bin/bash
# On the GPU box: parity tests against one variant library, then the streamed terrain bench of every variant,
# then isolated-call workloads with the pieces switch.  Usage: scripts/gpu_ab.sh <test-variant-name>
mkdir -p gpurun_out
v=$1
MM_LIB_PATH=$PWD/mimas_b200/variants/libmm_$v.so timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab_pytest_$v.log 2>&1; tail -4 gpurun_out/ab_pytest_$v.log
for f in mimas_b200/variants/libmm_*.so; do
  n=$(basename $f .so)
  for rep in 1 2; do
  MM_LIB_PATH=$PWD/$f timeout 300 python bench.py --quick --steps 600 --no-verify 2> gpurun_out/ab_$n.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$n', round(d['ms_per_step']*1e3,2), 'us/step frac', round(d['roofline']['frac'],4), 'iso', round(d['roofline']['isolated_batch_ms']*1e3,1))"
  done
done
