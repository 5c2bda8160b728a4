timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in 0 1 0 1; do
  if [ $v = 1 ]; then export MM_NO_FUSED_SINGLE=1; else unset MM_NO_FUSED_SINGLE; fi
  python scripts/run_configs.py cfg1 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('no_fused', $v, json.dumps({k:v for k,v in d.items() if k not in ('config','name')})[:600])"
done
