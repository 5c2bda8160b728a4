#!/usr/bin/env python
"""mm_gen_mapgen_device over the bench's 32x32x4 region, three times (run on the GPU box, plain and under ncu:
scripts/gpu_mapgen_ncu.sh) -- what bounds the mapgen kernel: f64 instruction counts against its duration."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mimas_b200  # noqa: E402

m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n = 4096
dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
for _ in range(3):
    m.gen_mapgen_device(78, (-16, -16, -2), (32, 32, 4), dc, do)
m.sync()
m.close()
print("ok")
