#!/usr/bin/env python
"""A soak of the launch chain and the output ring in place of compute-sanitizer's racecheck (closed on this pool):
`iters` batches of very different content and size (terrain, random ids, checkerboards, nearly empty, single chunks
through the fused launch, zero-chunk batches) submitted in random order through rings of 2-5 slots with everything
the ring allows kept in flight, a random subset read back by ticket while later batches run, host-path calls and
synchronous calls mixed in; every read-back is byte-compared with the oracle's mesh of the same input.
Test infrastructure (run on the GPU box); one JSON line.     python tests/gpu_ring_soak.py <iters> [seed]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mimas_b200  # noqa: E402
import oracle  # noqa: E402
from tests.util import assert_batch_equal, checkerboard_chunks, load_golden, oracle_cache, product_cache  # noqa: E402

iters = int(sys.argv[1])
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rng = np.random.RandomState(seed)
G = load_golden()
oc = oracle_cache(G)
t0 = time.time()
bad, compared, compared_quads, host_calls = [], 0, 0, 0
with mimas_b200.Mesher(0, product_cache(G)) as m:
    sets = []
    for k in range(14):
        kind = k % 7
        if kind == 0:
            c, o = mimas_b200.gen_terrain_host(78, (-8 + 5 * k, -8, -2), (16, 16, 4))       # 1024 chunks
        elif kind == 1:
            c = rng.randint(0, 60, size=(200 + 30 * k, 4096)).astype(np.uint8)
            o = (rng.randint(-1000, 1000, size=(len(c), 3)) * 16).astype(np.int64)
        elif kind == 2:
            c, o = checkerboard_chunks(4 + k)
        elif kind == 3:
            c, o = np.zeros((700, 4096), np.uint8), (rng.randint(-50, 50, size=(700, 3)) * 16).astype(np.int64)
            c[::37, ::5] = 5
        elif kind == 4:
            c, o = mimas_b200.gen_terrain_host(78, (3 + k, 2, 1), (1, 1, 1))                  # one chunk: the fused launch
        elif kind == 5:
            c, o = mimas_b200.gen_terrain_host(5, (-30, 4 * k, -3), (32, 32, 4))             # a full 4096-chunk launch
        else:
            c, o = mimas_b200.gen_terrain_host(9, (k, k, -1), (24, 24, 9))                   # 5184 chunks: two launch pairs + carry
        c, o = np.ascontiguousarray(c), np.ascontiguousarray(o, dtype=np.int64)
        dc, do = m.dev_alloc(c.size), m.dev_alloc(o.size * 8)
        m.upload(dc, c)
        m.upload(do, o)
        sets.append((c, o, dc, do, oracle.mesh_batch(c, o, oc, threads=os.cpu_count() or 4)))
    small = [s for s in sets if len(s[0]) <= 256]
    slots = 3
    m.ring_configure(slots)
    live = []   # (ticket, set index) still in the ring
    for it in range(iters):
        r = rng.rand()
        if r < 0.004:
            slots = int(rng.randint(2, 6))
            m.ring_configure(slots)
            live = []
        elif r < 0.03 and small:      # the reference's call pattern between ring batches: host in, host out
            s = small[rng.randint(len(small))]
            try:
                assert_batch_equal(m.mesh_batch(s[1], s[0]), s[4])
            except AssertionError as e:
                bad.append(f"iter {it}: host call: {str(e)[:160]}")
            host_calls += 1
            live = live[-(slots - 1):] if slots > 1 else []
            continue
        k = int(rng.randint(len(sets)))
        c, o, dc, do, want = sets[k]
        n = len(c) if rng.rand() > 0.03 else 0
        t = m.submit(dc, do, n)
        live.append((t, k if n else -1))
        live = live[-slots:]
        if rng.rand() < 0.35 and live:
            tk, ks = live[int(rng.randint(len(live)))]
            try:
                got = m.copy_out_ticket(tk)
                if ks >= 0:
                    assert_batch_equal(got, sets[ks][4])
                    compared_quads += (len(got.intransparent) + len(got.transparent)) // 6
                else:
                    assert len(got.intransparent) == 0 and len(got.transparent) == 0
                compared += 1
            except AssertionError as e:
                bad.append(f"iter {it}: ticket {tk} of set {ks}: {str(e)[:160]}")
    for tk, ks in live:
        got = m.copy_out_ticket(tk)
        if ks >= 0:
            try:
                assert_batch_equal(got, sets[ks][4])
            except AssertionError as e:
                bad.append(f"tail: ticket {tk} of set {ks}: {str(e)[:160]}")
            compared += 1
print(json.dumps({"what": "launch chain + output ring soak, read-backs byte-compared with the oracle", "iters": iters, "seed": seed,
                  "read_backs_compared": compared, "quads_compared": int(compared_quads), "host_calls": host_calls,
                  "disagreements": len(bad), "examples": bad[:5], "seconds": round(time.time() - t0, 1)}))
