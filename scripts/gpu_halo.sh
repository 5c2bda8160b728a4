mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/halo_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/halo_pytest.log
tail -5 gpurun_out/halo_pytest.log
timeout 600 python scripts/run_configs.py cfg_halo > gpurun_out/halo_cfg.jsonl 2> gpurun_out/halo_cfg.err; echo "cfg exit $?"; cat gpurun_out/halo_cfg.jsonl; tail -5 gpurun_out/halo_cfg.err
timeout 600 python bench.py > gpurun_out/halo_bench.json 2> gpurun_out/halo_bench.err; echo "bench exit $?"; cat gpurun_out/halo_bench.json
