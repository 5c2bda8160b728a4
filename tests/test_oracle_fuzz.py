"""Differential fuzz of the three CPU formulations of mesh_for_chunk (mimas-meshgen/lib.rs:231-446) -- pin (4)
of DESIGN.md section 2.  The reference ships no vectors and cannot be run here, so the residual risk is a shared
misreading; the three formulations were written to different shapes (state machine in C, the reference's own
closures in Python, itertools.groupby over a pass table), and here they must agree byte for byte on random
chunks x random TABLES (block ids sharing textures: the Walker compares TextureId, lib.rs:253-258) x origins up
to 2^62 (`as f32` and `v - last_v` round, lib.rs:249,256).  CPU only."""
import os
from concurrent.futures import ProcessPoolExecutor

import numpy as np

from tests import fuzz_util

N_CASES = 512


def test_three_formulations_agree_on_random_tables_chunks_and_origins():
    workers = min(os.cpu_count() or 1, 16)
    with ProcessPoolExecutor(max_workers=workers) as ex:
        bad = [r for r in ex.map(fuzz_util.three_way, range(N_CASES), chunksize=8) if r is not None]
    assert not bad, "\n".join(bad[:10])


def test_fuzz_cases_cover_what_they_claim():
    """The generator really produces shared textures across different drawn ids, fallback ids, crossed ids,
    origins beyond 2^24 and both sparse and dense chunks."""
    shared = fallback_hit = crossed = huge = sparse = dense = 0
    for seed in range(N_CASES):
        rows, fb, chunk, offs = fuzz_util.make_case(seed)
        present = set(int(v) for v in np.unique(chunk))
        blocky = [(i, r) for i, r in enumerate(rows) if r[0] == 1 and i in present]
        sides = [r[1] for _, r in blocky] + ([fb[0]] if any(i >= len(rows) for i in present) else [])
        shared += len(sides) != len(set(sides))
        fallback_hit += any(i >= len(rows) for i in present)
        crossed += any(i < len(rows) and rows[i][0] == 2 for i in present)
        huge += any(abs(o) >= 2 ** 24 for o in offs)
        fill = np.count_nonzero(chunk) / 4096
        sparse += fill < 0.2
        dense += fill > 0.8
    assert shared > 150 and fallback_hit > 200 and crossed > 150 and huge > 300 and sparse > 20 and dense > 100, \
        (shared, fallback_hit, crossed, huge, sparse, dense)
