#!/bin/bash
# On the GPU box: every BASELINE.json config as a bench.py line (tag_bench_<workload>.json under gpurun_out/),
# then the CPU reference arm of the default workload.  Usage: scripts/gpu_bench_all.sh [tag] [workloads...]
tag=${1:-r02}; shift
wls=${@:-terrain single checkerboard16384 radius8 radius16 radius32 edits million mapgen}
mkdir -p gpurun_out
for w in $wls; do
  timeout 900 python bench.py --workload $w > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err
  echo "$w exit $?"; tail -c 600 gpurun_out/${tag}_bench_$w.err
  python - <<P
import json
try:
    d = json.loads(open("gpurun_out/${tag}_bench_$w.json").read().strip().splitlines()[-1])
    r = d["roofline"]
    print("  %-18s %10.3f ms/step %12.0f chunks/s  frac %.3f  iso %.4f ms  e2e %.0f chunks/s (%s)  cpu %s" % (
        "$w", d["ms_per_step"], d["value"], r["frac"], r["isolated_batch_ms"], d["e2e"]["value"], d["e2e"]["path"],
        d.get("cpu_baseline", {}).get("value")))
except Exception as e:
    print("  $w: no line:", e)
P
done
timeout 600 python bench.py --workload terrain --consume > gpurun_out/${tag}_bench_terrain_consumed.json 2> gpurun_out/${tag}_bench_terrain_consumed.err; echo "consume exit $?"
python -c "
import json
d=json.loads(open('gpurun_out/${tag}_bench_terrain_consumed.json').read().strip().splitlines()[-1])
print('  consumed: %.4f ms/step %.0f chunks/s' % (d['ms_per_step'], d['value']), d.get('consumed'))"
timeout 600 python bench.py --impl reference > gpurun_out/${tag}_bench_reference.json 2>&1; cat gpurun_out/${tag}_bench_reference.json | cut -c1-300
python - <<P > gpurun_out/${tag}_mapgen_time.json
import json, time, torch, mimas_b200
m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n = 4096
dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
m.gen_mapgen_device(78, (-16, -16, -2), (32, 32, 4), dc, do)
t = []
for _ in range(5):
    t0 = time.perf_counter(); m.gen_mapgen_device(78, (-16, -16, -2), (32, 32, 4), dc, do); t.append(time.perf_counter() - t0)
print(json.dumps({"what": "mm_gen_mapgen_device, 4096 chunks (32x32x4 region), host clock around the synchronous call incl. table build + upload",
                  "ms_median": 1e3 * sorted(t)[2], "chunks_per_s": n / sorted(t)[2]}))
P
cat gpurun_out/${tag}_mapgen_time.json
