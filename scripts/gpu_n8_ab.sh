N=${1:-8}
MM_PIPE_MIN_CHUNKS=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('plain value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r01_bench_n$N.json 2> gpurun_out/r01_bench_n$N.err; cat gpurun_out/r01_bench_n$N.json
