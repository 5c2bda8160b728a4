#!/usr/bin/env python
"""Single-chunk latency of the crate-shaped call from a COMPILED host (tests/cxx/host_mirror_test latency): the two
spawn chunks of config 1 and three known-answer chunks, one mesh_for_chunk call each (incl. the copies into the
caller's vectors), next to the CPU mesher on one core.  Run on the GPU box; prints JSON lines."""
import json
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mimas_b200  # noqa: E402
import oracle  # noqa: E402
from tests.util import load_golden, oracle_cache  # noqa: E402

G = load_golden()
cases = [mimas_b200.gen_terrain_host(78, (3, 2, 1), (1, 1, 1)), mimas_b200.gen_terrain_host(78, (3, 2, 0), (1, 1, 1))]
chunks = [c[0][0] for c in cases] + [G.cases[n]["chunk"] for n in ("layered_terrain", "random_dense_ids", "checkerboard_even")]
offs = [c[1][0] for c in cases] + [G.cases[n]["offs"] for n in ("layered_terrain", "random_dense_ids", "checkerboard_even")]
path = os.path.join(tempfile.mkdtemp(), "cases.bin")
with open(path, "wb") as f:
    rows = G.entries[:G.n_known]
    f.write(struct.pack("<II4H", 0x54434D4D, len(rows), *G.fallback, 0))
    for r in rows:
        f.write(struct.pack("<5H", *r))
    f.write(struct.pack("<I", len(chunks)))
    for c, o in zip(chunks, offs):
        f.write(np.asarray(o, dtype=np.int64).tobytes())
        f.write(np.asarray(c, dtype=np.uint8).tobytes())
        f.write(struct.pack("<QQ", 0, 0))
subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "cxx"), "-s"], check=True)
r = subprocess.run([os.path.join(ROOT, "tests", "cxx", "host_mirror_test"), "latency", path], capture_output=True, text=True)
oc = oracle_cache(G)
for line, c, o in zip(r.stdout.splitlines(), chunks, offs):
    d = json.loads(line)
    cpu = sorted(oracle.mesh_batch(c[None], np.asarray(o)[None], oc, threads=1, discard=True)["seconds"] for _ in range(200))[100]
    d["cpu_1core_us_median"] = 1e6 * cpu
    print(json.dumps(d))
sys.stderr.write(r.stderr)
