"""Random tables, chunks and origins for the differential fuzz of the three CPU formulations of the mesher
(tests/test_oracle_fuzz.py): the C oracle (Walker state machine, oracle/meshgen_oracle.c), the Python
transliteration that keeps the reference's closures (oracle/meshgen_oracle_py.py) and the independent
groupby mesher that never saw the Walker (tests/golden/make_kats.py).  CPU only, test infrastructure."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

# chunk origins (per component): zero, +-16k, around 2^24 where `as f32` starts to round, far beyond it
OFFSETS = [0, 16, -16, 16 * 1000, -16 * 1000, 2 ** 24 - 16, 2 ** 24, 2 ** 24 + 16, -(2 ** 24) - 16, -(2 ** 24) + 16,
           -(2 ** 24), 2 ** 25 - 16, 2 ** 40, -(2 ** 40), 2 ** 62]


def random_table(rng):
    """(rows, fallback): rows of (kind, sides, top, bottom, crossed) for 2..40 known ids whose face textures
    are drawn from a pool of 3..8 texture ids, so that DIFFERENT block ids sharing a texture is the norm --
    the Walker merges on TextureId, not on block id (mimas-meshgen/lib.rs:253-258).  Undrawn and crossed ids
    are mixed in; the table is shorter than 256 so that ids beyond it hit the fallback (lib.rs:105-109)."""
    n_known = int(rng.randint(2, 41))
    pool = [int(v) for v in rng.choice([0, 1, 2, 3, 7, 255, 256, 4095, 65534, 65535], size=rng.randint(3, 9), replace=False)]
    rows = []
    for i in range(n_known):
        u = rng.rand()
        if i == 0 and rng.rand() < 0.8 or u < 0.15:
            rows.append((0, 0, 0, 0, 0))
        elif u < 0.30:
            rows.append((2, 0, 0, 0, int(rng.choice(pool))))
        else:
            style = rng.randint(0, 3)
            s, t, b = (int(rng.choice(pool)) for _ in range(3))
            if style == 0:
                t = b = s                       # BlockTextureIds::uniform
            elif style == 1:
                b = t                           # new_tb
            rows.append((1, s, t, b, 0))
    fallback = tuple(int(rng.choice(pool)) for _ in range(3))
    return rows, fallback


def full_rows(rows, fallback):
    """The 256-row form tests/golden/make_kats.mesh takes: unknown ids are blocky with the fallback ids."""
    return list(rows) + [(1, fallback[0], fallback[1], fallback[2], 0)] * (256 - len(rows))


def random_chunk(rng, n_known):
    """4096 block ids in one of several shapes: iid noise of varying density, runs smeared along x / y (long
    merges), layers over z, or solid -- always from a palette that mixes known ids, air and unknown ids."""
    k = int(rng.randint(2, 9))
    palette = list(rng.randint(0, n_known, size=k)) + [0] * int(rng.randint(0, 4))
    if rng.rand() < 0.6:
        palette += [int(rng.randint(n_known, 256))] * int(rng.randint(1, 3))
    palette = np.array(palette, dtype=np.uint8)
    mode = int(rng.randint(0, 5))
    c = palette[rng.randint(0, len(palette), size=(16, 16, 16))]
    if mode == 0:
        c[rng.rand(16, 16, 16) < rng.choice([0.2, 0.5, 0.9, 0.98])] = 0
    elif mode == 1:   # runs along y
        w = int(rng.choice([2, 4, 8, 16]))
        c = np.repeat(c[:, ::w, :], w, axis=1)
    elif mode == 2:   # runs along x
        w = int(rng.choice([2, 4, 8, 16]))
        c = np.repeat(c[::w, :, :], w, axis=0)
    elif mode == 3:   # layers over z with a few holes
        c = np.broadcast_to(palette[rng.randint(0, len(palette), size=16)], (16, 16, 16)).copy()
        c[rng.rand(16, 16, 16) < 0.03] = 0
    # mode 4: dense iid
    return np.ascontiguousarray(c, dtype=np.uint8).reshape(4096)


def make_case(seed):
    rng = np.random.RandomState(seed)
    rows, fallback = random_table(rng)
    chunk = random_chunk(rng, len(rows))
    offs = tuple(int(rng.choice(len(OFFSETS))) for _ in range(3))
    offs = tuple(OFFSETS[i] for i in offs)
    return rows, fallback, chunk, offs


def three_way(seed):
    """Returns None when the three formulations agree byte for byte on case `seed`, else a description."""
    import make_kats as mk
    import oracle
    from oracle import meshgen_oracle_py as pyo
    from tests.util import first_diff
    rows, fallback, chunk, offs = make_case(seed)
    ci, ct = oracle.mesh_for_chunk(offs, chunk, oracle.cache_from_entries(rows, len(rows), fallback))
    gi, gt = mk.mesh(chunk, offs, full_rows(rows, fallback))
    pr, prt = pyo.mesh_for_chunk(offs, chunk, pyo.PyCache(rows, len(rows), fallback))
    pi, pt = pyo.to_structured(pr, oracle.VERTEX_DTYPE), pyo.to_structured(prt, oracle.VERTEX_DTYPE)
    for name, a, b in (("groupby vs C oracle, intransparent", gi, ci), ("groupby vs C oracle, transparent", gt, ct),
                       ("Python transliteration vs C oracle, intransparent", pi, ci),
                       ("Python transliteration vs C oracle, transparent", pt, ct)):
        d = first_diff(a.tobytes(), b.tobytes())
        if d is not None:
            return f"seed {seed} (offs {offs}, {len(rows)} known ids): {name}: {d}"
    return None
