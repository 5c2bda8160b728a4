#!/bin/bash
# Runs on the GPU box (via gpurun): parity tests, smoke, bench, then the ncu launch list + one full capture.
# Usage: scripts/gpu_check.sh [tag]   -> logs under gpurun_out/<tag>_*
tag=${1:-r1}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total,clocks.max.sm --format=csv > gpurun_out/${tag}_gpu.txt 2>&1
python -c "import os; print('cpus', os.cpu_count())" >> gpurun_out/${tag}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/${tag}_smoke.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
cat gpurun_out/${tag}_bench.json; tail -3 gpurun_out/${tag}_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>&1; cat gpurun_out/${tag}_bench_ref.json
