mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_raster.py -m gpu -x -q > gpurun_out/raster_pytest.log 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/raster_pytest.log
timeout 600 python scripts/run_configs.py cfg_frame > gpurun_out/frame_cfg.jsonl 2> gpurun_out/frame_cfg.err; echo "cfg exit $?"; cat gpurun_out/frame_cfg.jsonl; tail -5 gpurun_out/frame_cfg.err
