for v in 0 1024 0 1024; do
MM_PIPE_MIN_CHUNKS=$v python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pipe_min', $v, 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
done
