#!/bin/bash
# On an 8-GPU box (gpurun --gpus 8): scaling evidence.  terrain = weak scaling (BENCH/SCALE's workload), million = strong.
tag=${1:-r02}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/${tag}_topo8.txt 2>&1; python -c "import os; print('cpus', os.cpu_count())" >> gpurun_out/${tag}_topo8.txt
port=29500
for n in 8 4 2; do
  port=$((port+1))
  timeout 600 $TR --nproc-per-node $n --master-port $port bench.py --gpus $n --workload million > gpurun_out/${tag}_bench_million_n$n.json 2> gpurun_out/${tag}_bench_million_n$n.err
  echo "million n=$n exit $?"; tail -1 gpurun_out/${tag}_bench_million_n$n.json | cut -c1-260
done
for n in 8 2; do
  port=$((port+1))
  timeout 600 $TR --nproc-per-node $n --master-port $port bench.py --gpus $n > gpurun_out/${tag}_bench_terrain_n$n.json 2> gpurun_out/${tag}_bench_terrain_n$n.err
  echo "terrain n=$n exit $?"; tail -1 gpurun_out/${tag}_bench_terrain_n$n.json | cut -c1-260
done
port=$((port+1)); timeout 300 $TR --nproc-per-node 8 --master-port $port scripts/timeline.py 48 > gpurun_out/${tag}_timeline_n8.jsonl 2> gpurun_out/${tag}_timeline_n8.err; grep period gpurun_out/${tag}_timeline_n8.jsonl | cut -c1-400
timeout 120 python scripts/timeline.py 48 | tail -1 > gpurun_out/${tag}_timeline_n1.jsonl; cat gpurun_out/${tag}_timeline_n1.jsonl
for n in 1 2 4 8; do port=$((port+1)); timeout 200 $TR --nproc-per-node $n --master-port $port scripts/d2h_scaling.py 2>/dev/null | grep whole_box | tee -a gpurun_out/${tag}_d2h_scaling.jsonl; done
