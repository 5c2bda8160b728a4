#!/usr/bin/env python
"""Is the bench's synthetic look-alike (csrc/terrain_gen.h) representative of real mapgen chunks?  Meshes the
same 32x32x4 region (config 2's shape) from the restatement of mimas-server's mapgen phase one
(oracle/mapgen_oracle.c; parity to the reference unpinned) and from the look-alike with the CPU oracle and writes
both histograms of vertices per chunk plus the shares of empty chunks.  CPU only (a few minutes).

    python scripts/mapgen_histogram.py > profiles/r02_mapgen_histogram.json
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mimas_b200  # noqa: E402
import oracle  # noqa: E402
from mimas_b200 import mapgen as pm  # noqa: E402
from oracle import mapgen as om  # noqa: E402
from tests.util import oracle_cache  # noqa: E402

C0, DIMS, SEED = (-16, -16, -2), (32, 32, 4), 78
EDGES = [0, 1, 300, 600, 1200, 2400, 4800, 9600, 200000]


def stats(chunks, offs):
    r = oracle.mesh_batch(chunks, offs, oracle_cache(), threads=os.cpu_count() or 1)
    v = (r["count_i"] + r["count_t"]).astype(np.int64)
    ids, cnt = np.unique(chunks, return_counts=True)
    return {"chunks": len(chunks), "vertices_per_chunk_mean": float(v.mean()), "vertex_bytes_per_chunk_mean": 36 * float(v.mean()),
            "empty_chunk_share": float((v == 0).mean()), "all_air_chunk_share": float((chunks == 0).all(axis=1).mean()),
            "transparent_vertex_share": float(r["count_t"].sum() / max(v.sum(), 1)),
            "histogram_edges": EDGES, "histogram": np.histogram(v, bins=EDGES)[0].tolist(),
            "per_z_layer_mean": [float(v[offs[:, 2] == z].mean()) for z in sorted(set(offs[:, 2].tolist()))],
            "block_id_shares": {int(i): float(c / chunks.size) for i, c in zip(ids, cnt) if c / chunks.size > 1e-4}}


if __name__ == "__main__":
    rc, ro, trees = om.gen_region(SEED, C0, DIMS, pm.to_c(pm.default_mapgen_params()))
    lc, lo = mimas_b200.gen_terrain_host(SEED, C0, DIMS)
    print(json.dumps({"region": {"c0": C0, "dims": DIMS, "seed": SEED},
                      "mapgen_restatement (oracle/mapgen_oracle.c, phase one only, no trees)": stats(rc, ro),
                      "bench_look_alike (csrc/terrain_gen.h)": stats(lc, lo),
                      "tree_spawn_points_in_region": int(trees.sum())}, indent=1))
