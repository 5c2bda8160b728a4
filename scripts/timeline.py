#!/usr/bin/env python
"""Device timeline of consecutive overlapped steps (run on the GPU box): where a step's time goes.
    python scripts/timeline.py [steps]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mimas_b200

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 24
m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n, rot = 4096, 6
dc = [torch.empty(n * 4096, dtype=torch.uint8, device="cuda") for _ in range(rot)]
do = [torch.empty(n * 3, dtype=torch.int64, device="cuda") for _ in range(rot)]
for r in range(rot):
    m.gen_terrain_device(78, (-16 + 32 * r, -16, -2), (32, 32, 4), dc[r].data_ptr(), do[r].data_ptr())
m.sync()
out = m.mesh_batch_device(dc[0].data_ptr(), do[0].data_ptr(), n)  # sizes the arenas
for i in range(10):
    m.mesh_batch_async(dc[i % rot].data_ptr(), do[i % rot].data_ptr(), n)
m.batch_finish()
m.debug_trace(steps)
for i in range(steps):
    m.mesh_batch_async(dc[i % rot].data_ptr(), do[i % rot].data_ptr(), n)
m.batch_finish()
t = m.debug_trace(0).astype(np.int64)
t0 = t[0, 0, 0]
print("step |  A.start  A.workdone  A.end |  E.start  E.pastwait  E.end | period(us)")
prev = None
for k in range(len(t)):
    a, e = (t[k, 0] - t0) / 1e3, (t[k, 1] - t0) / 1e3
    per = "" if prev is None else f"{e[3] - prev:7.1f}"
    prev = e[3]
    print(f"{k:4d} | {a[0]:8.1f} {a[2]:10.1f} {a[3]:7.1f} | {e[0]:8.1f} {e[1]:10.1f} {e[3]:7.1f} | {per}")
