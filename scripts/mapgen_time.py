import sys, time
sys.path.insert(0, "/root/repo")
import mimas_b200
m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n = 4096
dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
m.gen_mapgen_device(78, (-16, -16, -2), (32, 32, 4), dc, do)
t = []
for _ in range(7):
    t0 = time.perf_counter(); m.gen_mapgen_device(78, (-16, -16, -2), (32, 32, 4), dc, do); t.append(time.perf_counter() - t0)
print("ms median %.3f min %.3f" % (1e3 * sorted(t)[3], 1e3 * min(t)))
