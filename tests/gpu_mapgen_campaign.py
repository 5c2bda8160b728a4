#!/usr/bin/env python
"""A longer run of the device-mapgen parity check than tests/test_mapgen.py affords: random world seeds, region
positions (surface, deep underground, far from the origin), region shapes and ore / plant tables (0-8 ores, 0-16
plants, random 8-byte stream names, thresholds that fire rarely or often), mm_gen_mapgen_device against
oracle/mapgen_oracle.c byte for byte (chunks, origins, tree spawn counts).  Test infrastructure (GPU box); one JSON line.
    python tests/gpu_mapgen_campaign.py <first_case> <cases>"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mimas_b200  # noqa: E402
from mimas_b200 import mapgen as pm  # noqa: E402
from oracle import mapgen as om  # noqa: E402

first, cases = int(sys.argv[1]), int(sys.argv[2])


def name(rng):
    return bytes(int(v) for v in rng.randint(32, 127, size=8))


def random_params(rng):
    if rng.rand() < 0.25:
        return pm.default_mapgen_params()
    ids = rng.permutation(np.arange(1, 200))[:40].tolist()
    roles = {"air": 0, "water": ids[0], "sand": ids[1], "ground": ids[2], "ground_top": ids[3], "stone": ids[4]}
    ores = [pm.Ore(ids[5 + i], name(rng), name(rng), float(rng.choice([0.02, 0.063951, 0.083951, 0.11, 0.3])),
                   float(rng.choice([0.1, 0.3, 0.6, 0.9])), float(rng.choice([0.1, 0.4, 0.75, 2.0])),
                   float(rng.choice([0.05, 0.3, 0.5, 0.8])), int(rng.choice([-98, -60, -30, -10, 0, 5, 40])))
            for i in range(int(rng.randint(0, 9)))]
    plants = [pm.Plant(ids[14 + i], name(rng), float(rng.choice([0.0, 0.02, 0.2, 0.5, 0.8, 1.0])))
              for i in range(int(rng.randint(0, 17)))]
    return pm.MapgenParams(roles, ores, plants)


t0 = time.time()
bad, chunks_total, ids_seen = [], 0, set()
with mimas_b200.Mesher(0, mimas_b200.default_cache()) as m:
    for case in range(first, first + cases):
        rng = np.random.RandomState(case)
        seed = int(rng.randint(0, 2 ** 32)) | (int(rng.randint(0, 2 ** 32)) << 32)
        span = int(rng.choice([10, 300, 60000]))
        c0 = (int(rng.randint(-span, span)), int(rng.randint(-span, span)), int(rng.choice([-40, -8, -3, -2, -1, 0, 1, 3])))
        dims = tuple(int(v) for v in rng.randint(1, 6, size=3))
        n = dims[0] * dims[1] * dims[2]
        mp = random_params(rng)
        want_c, want_o, want_t = om.gen_region(seed, c0, dims, pm.to_c(mp))
        dc, do, dt = m.dev_alloc(n * 4096), m.dev_alloc(n * 24), m.dev_alloc(n * 4)
        m.gen_mapgen_device(seed, c0, dims, dc, do, mp, dt)
        got_c = m.download(dc, n * 4096)
        ok = (got_c.tobytes() == want_c.tobytes() and m.download(do, n * 24).tobytes() == want_o.tobytes()
              and m.download(dt, n * 4).view(np.uint32).tolist() == want_t.tolist())
        if not ok:
            bad.append(f"case {case}: seed {seed} c0 {c0} dims {dims}")
        chunks_total += n
        ids_seen.add(len(np.unique(want_c)))
        for p in (dc, do, dt):
            m.dev_free(p)
print(json.dumps({"what": "mm_gen_mapgen_device vs oracle/mapgen_oracle.c, byte compare of chunks, origins, tree spawn counts",
                  "first_case": first, "cases": cases, "chunks": chunks_total, "disagreements": len(bad), "examples": bad[:5],
                  "max_distinct_ids_in_a_region": max(ids_seen), "seconds": round(time.time() - t0, 1)}))
