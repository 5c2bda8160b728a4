for f in mimas_b200/variants/libmm_*.so; do
  echo $(basename $f) $(MM_LIB_PATH=$PWD/$f python scripts/run_configs.py cfg_frame 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print([round(f['raster_ms'],3) for f in d['frames']])")
done
