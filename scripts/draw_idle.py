#!/usr/bin/env python
"""One frame of the headless draw (cull + vertex shader) over a meshed 4096-chunk terrain batch on an otherwise idle
GPU: CUDA-event time of the frame's kernels and host time of the synchronous call (run on the GPU box)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mimas_b200  # noqa: E402
from mimas_b200.render import Camera  # noqa: E402

m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n = 4096
dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
m.gen_terrain_device(78, (-16, -16, -2), (32, 32, 4), dc, do)
m.mesh_batch_device(dc, do, n)
m.mesh_batch_device(dc, do, n)
cam = Camera(pos=(8.0, 8.0, 24.0), pitch=-10.0, yaw=30.0, viewing_range=1e6).to_c()
for _ in range(5):
    d = m.draw_batch(cam)
ks, ws = [], []
for _ in range(50):
    t0 = time.perf_counter()
    d = m.draw_batch(cam)
    ws.append(time.perf_counter() - t0)
    ks.append(d.kernel_ms)
nv = d.n_shaded_i + d.n_shaded_t
print("idle GPU frame: kernels %.1f us (median), call %.1f us, drawn %d of %d chunks, %d vertices, %.0f GB/s over 84 B per vertex" % (
    1e3 * np.median(ks), 1e6 * np.median(ws), d.n_drawn, n, nv, nv * 84 / (1e-3 * np.median(ks)) / 1e9))
m.close()
