import os, sys, time, json
import numpy as np
sys.path.insert(0, '/root/repo')
import mimas_b200
m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
hc, ho = mimas_b200.gen_terrain_host(78, (-16, -16, -2), (32, 32, 4))
sel = np.random.RandomState(1).choice(4096, n, replace=False) if n < 4096 else np.arange(4096)
hc, ho = np.ascontiguousarray(hc[sel]), np.ascontiguousarray(ho[sel])
pin_c, pin_o = m.host_alloc(n * 4096), m.host_alloc(n * 24)
pin_c[:] = hc.reshape(-1); pin_o[:] = ho.reshape(-1).view(np.uint8)
for _ in range(4): hm = m.mesh_batch_host_raw(pin_c.ctypes.data, pin_o.ctypes.data, n)
ts=[]
for _ in range(20):
    t0=time.perf_counter(); hm = m.mesh_batch_host_raw(pin_c.ctypes.data, pin_o.ctypes.data, n); ts.append(time.perf_counter()-t0)
print('n', n, 'mapped_max', os.environ.get('MM_MAPPED_MAX_CHUNKS'), 'median ms', 1e3*np.median(ts), 'chunks/s', n/np.median(ts), hm.n_intransparent)
