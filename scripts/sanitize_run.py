#!/usr/bin/env python
"""A small tour of every kernel for compute-sanitizer (memcheck / racecheck / synccheck / initcheck)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mimas_b200, oracle
from mimas_b200.render import Camera
from tests.util import load_golden, oracle_cache, product_cache, assert_batch_equal, checkerboard_chunks
g = load_golden()
oc = oracle_cache(g)
rng = np.random.RandomState(4)
with mimas_b200.Mesher(0, product_cache(g)) as m:
    tc, to = mimas_b200.gen_terrain_host(78, (-3, -3, -2), (6, 6, 4))
    rc = rng.randint(0, 256, size=(40, 4096)).astype(np.uint8)
    ro = (rng.randint(-100, 100, size=(40, 3)) * 16).astype(np.int64)
    cc, co = checkerboard_chunks(3)
    chunks, offs = np.concatenate([tc, rc, cc]), np.concatenate([to, ro, co])
    assert_batch_equal(m.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, oc, threads=4))      # mapped path
    os.environ["X"] = "1"
    n = len(chunks)
    dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
    m.upload(dc, chunks); m.upload(do, offs)
    for _ in range(3):
        m.mesh_batch_async(dc, do, n)                                                                   # chained launches
    out = m.batch_finish()
    assert_batch_equal(m.copy_out(out), oracle.mesh_batch(chunks, offs, oc, threads=4))
    d = m.draw_batch(Camera(pos=(3.0, -2.0, 12.0), pitch=-20.0, yaw=40.0, viewing_range=90.0))      # culling + vertex shader
    assert d.n_drawn > 0
    m.mesh_for_chunk(offs[0], chunks[0])
    big_c, big_o = mimas_b200.gen_terrain_host(78, (-16, -16, -1), (32, 32, 1))
    big_c = np.concatenate([big_c] * 5)[:4500]; big_o = np.concatenate([big_o] * 5)[:4500]
    b = m.mesh_batch(big_o, big_c)                                                                     # copy path, two launch pairs
    assert len(b.ranges_i) == 4500
print("sanitize tour ok")
