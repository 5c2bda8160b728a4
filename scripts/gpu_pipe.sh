mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pipe_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pipe_pytest.log
tail -5 gpurun_out/pipe_pytest.log
for v in 0 1024; do
MM_PIPE_MIN_CHUNKS=$v timeout 300 python scripts/e2e_try.py 4096
MM_PIPE_MIN_CHUNKS=$v timeout 300 python scripts/e2e_try.py 2048
MM_PIPE_MIN_CHUNKS=$v timeout 300 python scripts/e2e_try.py 1024
done
timeout 600 python bench.py > gpurun_out/pipe_bench.json 2> gpurun_out/pipe_bench.err; echo "bench exit $?"; cat gpurun_out/pipe_bench.json
