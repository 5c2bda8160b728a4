#!/bin/bash
# On the GPU box: parity tests of the in-tree library (every test under a time limit), then the streamed bench and the
# isolated-call workloads of the in-tree library and of every variant library.
tag=${1:-rX}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/${tag}_pytest.log 2>&1; tail -6 gpurun_out/${tag}_pytest.log
for f in mimas_b200/libmimas_meshgen_cuda.so mimas_b200/variants/libmm_*.so; do
  [ -f $f ] || continue
  n=$(basename $f .so)
  for rep in 1 2; do
  MM_LIB_PATH=$PWD/$f timeout 300 python bench.py --quick --steps 600 --no-verify 2> gpurun_out/${tag}_ab_$n.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$n', round(d['ms_per_step']*1e3,2), 'us/step frac', round(d['roofline']['frac'],4), 'iso', round(d['roofline']['isolated_batch_ms']*1e3,1))
except Exception as e: print('$n FAILED', e)"
  done
  for w in radius8 radius16 radius32 edits single; do
  MM_LIB_PATH=$PWD/$f timeout 300 python bench.py --workload $w --quick --no-verify 2>> gpurun_out/${tag}_ab_$n.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('   $n $w', round(d['ms_per_step']*1e3,1), 'us frac', round(d['roofline']['frac'],3))
except Exception as e: print('$n $w FAILED', e)"
  done
done
