#!/usr/bin/env python
"""Does L2 absorb part of the vertex stream because every step rewrites the same arena?  Streams the bench
workload through K contexts (K separate output arenas, one CUDA stream) in turn: with K arenas the same bytes
come round again only after K x 172 MB.  Prints the step time for K = 1, 2, 3."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mimas_b200  # noqa: E402

SEED, REGION, N, ROT = 78, (32, 32, 4), 4096, 12
stream = torch.cuda.Stream()
res = {}
for K in (1, 2, 3):
    ms = [mimas_b200.Mesher(0, mimas_b200.default_cache()) for _ in range(K)]
    for m in ms:
        m.set_stream(stream.cuda_stream)
    dc = [torch.empty(N * 4096, dtype=torch.uint8, device="cuda") for _ in range(ROT)]
    do = [torch.empty(N * 3, dtype=torch.int64, device="cuda") for _ in range(ROT)]
    for r in range(ROT):
        ms[0].gen_terrain_device(SEED, (-16 + 32 * r, -16, -2), REGION, dc[r].data_ptr(), do[r].data_ptr())
    ms[0].sync()
    with torch.cuda.stream(stream):
        for m in ms:  # size the arenas
            for r in range(ROT):
                m.mesh_batch_async(dc[r].data_ptr(), do[r].data_ptr(), N)
                m.batch_finish()
        steps = 300
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(stream)
            for i in range(steps):
                ms[i % K].mesh_batch_async(dc[i % ROT].data_ptr(), do[i % ROT].data_ptr(), N)
            e1.record(stream)
            torch.cuda.synchronize()
            t = e0.elapsed_time(e1) / steps
        for m in ms:
            m.batch_finish()
    res[f"arenas_{K}"] = round(t * 1e3, 2)
    for m in ms:
        m.close()
print(json.dumps({"us_per_step": res}))
