#!/usr/bin/env python
"""Where the analysis CTAs wait (run on the GPU box against a library built with -DMM_ANALYSE_PHASE_CLOCKS=1,
scripts/variants.sh): thread 0 of every analysis CTA sums the cycles between issuing a chunk's bulk load + the next
ticket's atomic and having the ticket back, the cycles it then still waits for the chunk, and its whole loop; per
launch the sums over all CTAs come back through mm_debug_trace.
    MM_LIB_PATH=mimas_b200/variants/libmm_clocks.so python scripts/phase_clocks.py [steps]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import mimas_b200  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 12
GHZ = 1.965
m = mimas_b200.Mesher(0, mimas_b200.default_cache())
n, rot = 4096, 6
dc = [torch.empty(n * 4096, dtype=torch.uint8, device="cuda") for _ in range(rot)]
do = [torch.empty(n * 3, dtype=torch.int64, device="cuda") for _ in range(rot)]
for r in range(rot):
    m.gen_terrain_device(78, (-16 + 32 * r, -16, -2), (32, 32, 4), dc[r].data_ptr(), do[r].data_ptr())
m.sync()
m.ring_configure(3)
for r in range(3):
    m.mesh_batch_device(dc[r].data_ptr(), do[r].data_ptr(), n)
pc, po = [t.data_ptr() for t in dc], [t.data_ptr() for t in do]
for alone in (False, True):
    m.submit_many(pc, po, [n] * rot, repeat=2)
    m.batch_finish()
    m.debug_trace(steps)
    if alone:  # every batch on an empty GPU
        for k in range(steps):
            m.mesh_batch_device(pc[k % rot], po[k % rot], n)
    else:
        m.submit_many(pc, po, [n] * rot, repeat=max(1, steps // rot))
        m.batch_finish()
    t = m.debug_trace(0).astype(np.uint64)
    for k in range(len(t)):
        a, e = t[k, 0], t[k, 1]
        w1 = int(a[1]) + 1 & (1 << 64) - 1
        atomic, load = w1 & 0xffffffff, w1 >> 32
        w6 = int(e[2])
        loop, chunks = w6 & 0xffffffff, w6 >> 32
        ctas = min(n, 148 * 7)
        print(json.dumps({"mode": "alone" if alone else "streamed", "step": k, "analysis_us": (int(a[2]) - int(a[0])) / 1e3,
                          "chunks": chunks, "per_chunk_us": {"ticket_back": atomic / GHZ / 1e3 / max(chunks, 1),
                                                             "then_chunk_there": load / GHZ / 1e3 / max(chunks, 1),
                                                             "whole_loop": loop / GHZ / 1e3 / max(chunks, 1)},
                          "loop_us_per_cta": loop / GHZ / 1e3 / ctas}))
m.close()
