#!/bin/bash
# On the GPU box: streamed terrain bench over run-time knobs and variant libraries.
mkdir -p gpurun_out
run() { env "$@" timeout 300 python bench.py --quick --steps 600 --no-verify 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$*', round(d['ms_per_step']*1e3,2), 'us/step frac', round(d['roofline']['frac'],4), 'iso', round(d['roofline']['isolated_batch_ms']*1e3,1))
except Exception as e: print('$* FAILED', e)"; }
for g in 3 4 5 6 7 8 10; do run MM_ANALYSE_GRID_PER_SM=$g; done
for f in mimas_b200/variants/libmm_*.so; do run MM_LIB_PATH=$PWD/$f; run MM_LIB_PATH=$PWD/$f MM_ANALYSE_GRID_PER_SM=6; done
