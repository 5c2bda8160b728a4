mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum --clock-control none -k regex:"raster|clear_k|resolve" -c 100 --csv --log-file gpurun_out/raster_launches.csv python scripts/run_configs.py cfg_frame > gpurun_out/raster_ncu.log 2>&1; echo "ncu exit $?"
python - <<'P'
import csv
rows=[r for r in csv.reader(open('gpurun_out/raster_launches.csv')) if len(r)>10]
h=rows[0]; ki=h.index('Kernel Name'); mi=h.index('Metric Name'); vi=h.index('Metric Value'); ii=h.index('ID')
d={}
for r in rows[1:]:
    d.setdefault((int(r[ii]), r[ki].split('(')[0][-24:]),{})[r[mi]]=r[vi]
for k in sorted(d): 
    if k[0] in (1,2,3,31,32,33,61,62,63): print(k, d[k])
P
