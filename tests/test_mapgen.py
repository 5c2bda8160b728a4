"""Mapgen phase one (mimas-server/mapgen.rs:64-245).  PARITY TO THE REFERENCE IS UNPINNED: the function's
arithmetic lives in crates that are not in the reference checkout and no reference test pins mapgen output.  What IS
pinned: the PCG-XSH-RR 64/32 and XXH64 parts of the restatement (oracle/mapgen_oracle.c) against published /
independently computed vectors (tests/golden/mapgen_vectors.json); structural properties of the Perlin restatement;
and, on the GPU, that the CUDA generator (mm_gen_mapgen_device) equals the restatement byte for byte and feeds the
mesher without a host copy."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from mimas_b200 import mapgen as pm
from oracle import mapgen as om
from tests.util import assert_batch_equal, load_golden, oracle_cache, product_cache

HERE = os.path.dirname(os.path.abspath(__file__))
V = json.load(open(os.path.join(HERE, "golden", "mapgen_vectors.json")))
REF_TOML = "/root/reference/mimas-server/game-params.toml"


def test_pcg32_published_reference_vector():
    """pcg32_srandom(42, 54): the demo output of the PCG reference implementation, also rand_pcg's own test vector."""
    p = om.pcg32(42, 54)
    got = [om.load().mg_pcg32_next_u32(C.byref(p)) for _ in range(6)]
    assert got == V["pcg32_seed42_stream54"] == [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]
    # rand_core next_u64_via_u32: low half first; rand 0.8 Standard f64: 53 high bits
    p = om.pcg32(42, 54)
    assert om.load().mg_pcg32_next_u64(C.byref(p)) == (0x7b47f409 << 32) | 0xa15c02b7
    p = om.pcg32(42, 54)
    assert om.load().mg_pcg32_gen_f64(C.byref(p)) == float(((0x7b47f409 << 32) | 0xa15c02b7) >> 11) * 2.0 ** -53


def test_xxh64_and_pos_hash_against_the_xxhash_library():
    l = om.load()
    for v in V["xxh64"]:
        b = bytes.fromhex(v["hex"])
        assert l.mg_xxh64(b, len(b), v["seed"]) == v["digest"], v
    assert l.mg_xxh64(b"", 0, 0) == 0xEF46DB3751D8E999          # the algorithm's published empty-input value
    for v in V["pos_hash"]:
        assert om.pos_hash(v["pos"]) == v["hash"], v


def test_permutation_table_and_perlin_structure():
    for seed in (0, 1, 78, 0xdeadbeef):
        t = om.perm_table(seed)
        assert sorted(t.tolist()) == list(range(256))
    assert om.perm_table(1).tolist() != om.perm_table(2).tolist()
    t = om.perm_table(78)
    rng = np.random.RandomState(0)
    for _ in range(200):   # gradient noise vanishes on the lattice, stays in [-1, 1], and is continuous
        ix, iy, iz = (int(v) for v in rng.randint(-1000, 1000, size=3))
        assert om.perlin2(t, ix, iy) == 0.0 and om.perlin3(t, ix, iy, iz) == 0.0
        x, y, z = (float(v) for v in rng.uniform(-300, 300, size=3))
        a, b = om.perlin2(t, x, y), om.perlin3(t, x, y, z)
        assert -1.0 <= a <= 1.0 and -1.0 <= b <= 1.0
        assert abs(om.perlin2(t, x + 1e-7, y) - a) < 1e-5 and abs(om.perlin3(t, x, y, z + 1e-7) - b) < 1e-5
    vals = [om.perlin3(t, *rng.uniform(-50, 50, size=3)) for _ in range(2000)]
    assert abs(np.mean(vals)) < 0.05 and 0.1 < np.std(vals) < 0.5


def test_default_mapgen_tables_match_the_reference_toml():
    mp = pm.default_mapgen_params()
    assert [o.block for o in mp.ores] == [9, 10, 13, 14, 15, 16] and len(mp.plants) == 13
    assert mp.roles == {"air": 0, "water": 1, "sand": 2, "ground": 3, "ground_top": 11, "stone": 5}
    assert C.sizeof(pm.to_c(mp)) == om.load().mg_sizeof_params()
    golden = json.load(open(os.path.join(HERE, "golden", "default_mapgen_params.json")))
    assert mp.to_json() == {k: golden[k] for k in ("roles", "ores", "plants")}
    if os.path.exists(REF_TOML):   # build container only
        assert pm.load_mapgen_params_from_toml(open(REF_TOML).read()).to_json() == mp.to_json()


def test_restatement_regression_digests_and_terrain_statistics():
    """The frozen digests pin the restatement against accidental change; the statistics say the chunks look like a
    world: water below z = 0, ground and grass at the surface, ores in the stone, some plants and trees."""
    cp = pm.to_c(pm.default_mapgen_params())
    for r in V["regions"]:
        chunks, offs, trees = om.gen_region(r["seed"], r["c0"], r["dims"], cp)
        assert hashlib.sha256(chunks.tobytes()).hexdigest() == r["sha256_chunks"]
        assert int(trees.sum()) == r["trees"]
    chunks, offs, trees = om.gen_region(78, (-4, -4, -2), (8, 8, 4), cp)
    ids, cnt = np.unique(chunks, return_counts=True)
    h = dict(zip(ids.tolist(), cnt.tolist()))
    assert h[0] > 0 and h[1] > 0 and h[5] > h[3] > 0 and h[11] > 0 and h.get(9, 0) > 0
    assert sum(h.get(i, 0) for i in range(43, 56)) > 0 and trees.sum() > 0
    assert (offs[:, 2] == np.repeat([-32, -16, 0, 16], 64)).all()
    below = chunks[offs[:, 2] < 0]
    assert not (below == 0).all(axis=1).any()              # no all-air chunk below sea level (water fills)


@pytest.fixture(scope="module")
def mesher():
    import mimas_b200
    m = mimas_b200.Mesher(0, product_cache(load_golden()))
    yield m
    m.close()


@pytest.mark.gpu
@pytest.mark.parametrize("seed,c0,dims", [(78, (-4, -4, -2), (8, 8, 4)), (12345, (100, -77, -30), (3, 3, 3)),
                                          (78, (3, 2, 1), (1, 1, 1)), (2 ** 63 + 5, (-1000, 999, -3), (5, 2, 6))])
def test_device_mapgen_equals_the_restatement_and_feeds_the_mesher(seed, c0, dims, mesher):
    import oracle
    mp = pm.default_mapgen_params()
    n = dims[0] * dims[1] * dims[2]
    want_c, want_o, want_t = om.gen_region(seed, c0, dims, pm.to_c(mp))
    dc, do, dt = mesher.dev_alloc(n * 4096), mesher.dev_alloc(n * 24), mesher.dev_alloc(n * 4)
    try:
        mesher.gen_mapgen_device(seed, c0, dims, dc, do, mp, dt)
        got_c = mesher.download(dc, n * 4096).reshape(n, 4096)
        bad = np.nonzero((got_c != want_c).any(axis=1))[0]
        assert len(bad) == 0, f"{len(bad)} of {n} chunks differ, first {bad[0]} at offs {want_o[bad[0]]}"
        assert mesher.download(do, n * 24).tobytes() == want_o.tobytes()
        assert mesher.download(dt, n * 4).view(np.uint32).tolist() == want_t.tolist()
        # straight into the mesher, no host copy
        got = mesher.copy_out(mesher.mesh_batch_device(dc, do, n))
        assert_batch_equal(got, oracle.mesh_batch(want_c, want_o, oracle_cache(), threads=8))
    finally:
        for p in (dc, do, dt):
            mesher.dev_free(p)


@pytest.mark.gpu
def test_device_mapgen_with_other_tables(mesher):
    """Not the default game: two ores with thresholds that fire often, plants that fire often, other role ids."""
    mp = pm.MapgenParams({"air": 0, "water": 7, "sand": 9, "ground": 4, "ground_top": 2, "stone": 1},
                         [pm.Ore(20, b"oreAnois", b"oreApcg ", 0.11, 0.3, 0.2, 0.1, -10),
                          pm.Ore(21, b"oreBnois", b"oreBpcg ", 0.07, 0.5, 0.3, 0.3, 5)],
                         [pm.Plant(30, b"plantone", 0.5), pm.Plant(31, b"planttwo", 0.8)])
    seed, c0, dims = 99, (-3, -3, -3), (6, 6, 5)
    n = dims[0] * dims[1] * dims[2]
    want_c, want_o, want_t = om.gen_region(seed, c0, dims, pm.to_c(mp))
    dc, do, dt = mesher.dev_alloc(n * 4096), mesher.dev_alloc(n * 24), mesher.dev_alloc(n * 4)
    try:
        mesher.gen_mapgen_device(seed, c0, dims, dc, do, mp, dt)
        assert mesher.download(dc, n * 4096).tobytes() == want_c.tobytes()
        assert mesher.download(dt, n * 4).view(np.uint32).tolist() == want_t.tolist()
        ids = set(np.unique(want_c).tolist())
        assert {20, 21, 30, 31} <= ids
    finally:
        for p in (dc, do, dt):
            mesher.dev_free(p)
