#!/bin/bash
# On the GPU box: quick device-resident bench of every library under mimas_b200/variants/.
mkdir -p gpurun_out
for f in mimas_b200/variants/libmm_*.so; do
  n=$(basename $f .so)
  MM_LIB_PATH=$PWD/$f timeout 300 python bench.py --quick --steps 400 --warmup 20 ${BENCH_ARGS} > gpurun_out/var_$n.json 2> gpurun_out/var_$n.err
  python - <<P
import json
try:
    d=json.loads(open("gpurun_out/var_$n.json").read().strip().splitlines()[-1])
    print("$n", round(d["ms_per_step"]*1e3,2), "us/step", round(d["roofline"]["frac"],4), "iso", d["roofline"].get("isolated_batch_ms"))
except Exception as e:
    print("$n", "FAILED", e)
P
done
