#!/usr/bin/env python
"""A longer run of the CUDA-vs-oracle fuzz than tests/test_gpu_parity.py affords (160 tables x 8 chunks on every GPU
test run): `tables` random tables (tests/fuzz_util.py: ids sharing textures, fallback and crossed ids) x 64 random
chunks each x origins up to 2^62, meshed through the C ABI as one batch and byte-compared with the C oracle; every
16th chunk also through the fused single-chunk launch.  Test infrastructure (run on the GPU box); one JSON line.
    python tests/gpu_fuzz_campaign.py <first_table> <tables>"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mimas_b200  # noqa: E402
import oracle  # noqa: E402
from tests import fuzz_util  # noqa: E402
from tests.util import assert_batch_equal  # noqa: E402

first, tables = int(sys.argv[1]), int(sys.argv[2])
PER = 64
t0 = time.time()
bad, quads = [], 0
with mimas_b200.Mesher(0) as m:
    for t in range(first, first + tables):
        rng = np.random.RandomState(t)
        rows, fb = fuzz_util.random_table(rng)
        chunks = np.stack([fuzz_util.random_chunk(rng, len(rows)) for _ in range(PER)])
        offs = np.array([[fuzz_util.OFFSETS[i] for i in rng.randint(0, len(fuzz_util.OFFSETS), size=3)] for _ in range(PER)],
                        dtype=np.int64)
        m.set_cache(mimas_b200.TextureIdCache.from_table_rows(rows, fb))
        want = oracle.mesh_batch(chunks, offs, oracle.cache_from_entries(rows, len(rows), fb), threads=os.cpu_count() or 4)
        quads += (len(want["intransparent"]) + len(want["transparent"])) // 6
        try:
            assert_batch_equal(m.mesh_batch(offs, chunks), want)
            for k in range(0, PER, 16):
                one = m.mesh_for_chunk(offs[k], chunks[k])
                a, n = int(want["first_i"][k]), int(want["count_i"][k])
                assert one.intransparent.tobytes() == want["intransparent"][a:a + n].tobytes()
                a, n = int(want["first_t"][k]), int(want["count_t"][k])
                assert one.transparent.tobytes() == want["transparent"][a:a + n].tobytes()
        except AssertionError as e:
            bad.append(f"table {t}: {str(e)[:200]}")
print(json.dumps({"what": "CUDA (C ABI: mm_mesh_batch_host + mm_mesh_for_chunk) vs oracle/meshgen_oracle.c, byte compare",
                  "first_table": first, "tables": tables, "chunks": tables * PER, "quads": int(quads),
                  "disagreements": len(bad), "examples": bad[:5], "seconds": round(time.time() - t0, 1)}))
