#!/usr/bin/env python
"""Where a step of the consumed stream (bench.py --consume) goes: host time of each call of the loop."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, mimas_b200
from mimas_b200.render import Camera
m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n, rot = 4096, 6
dc = [torch.empty(n * 4096, dtype=torch.uint8, device="cuda") for _ in range(rot)]
do = [torch.empty(n * 3, dtype=torch.int64, device="cuda") for _ in range(rot)]
for r in range(rot):
    m.gen_terrain_device(78, bench.region_origin(0, r), bench.REGION, dc[r].data_ptr(), do[r].data_ptr())
m.sync(); m.ring_configure(3)
for r in range(3):
    m.mesh_batch_device(dc[r].data_ptr(), do[r].data_ptr(), n)
cam = Camera(pos=(8.0, 8.0, 24.0), pitch=-10.0, yaw=30.0, viewing_range=1e6).to_c()
tk = [m.submit(dc[i % rot].data_ptr(), do[i % rot].data_ptr(), n) for i in range(2)]
T = np.zeros((60, 3)); W = np.zeros(60)
for i in range(60):
    tw = time.perf_counter(); m.batch_wait(tk[0]); W[i] = time.perf_counter() - tw
    t0 = time.perf_counter(); m.draw_ticket_async(tk.pop(0), cam)
    t1 = time.perf_counter(); tk.append(m.submit(dc[i % rot].data_ptr(), do[i % rot].data_ptr(), n))
    t2 = time.perf_counter(); d = m.draw_collect()
    t3 = time.perf_counter(); T[i] = (t1 - t0, t2 - t1, t3 - t2)
print("batch_wait %.1f us" % (1e6 * W[10:].mean()))
print("us per step: draw_async %.1f  submit %.1f  collect %.1f   draw kernel_ms %.4f" % (*(1e6 * T[10:].mean(axis=0)), d.kernel_ms))
# the same without a consumer
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(60):
    m.submit(dc[i % rot].data_ptr(), do[i % rot].data_ptr(), n)
m.batch_finish(); print("no consumer: %.1f us per step" % (1e6 * (time.perf_counter() - t0) / 60))
