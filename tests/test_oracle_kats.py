"""Pins the parity oracle (oracle/meshgen_oracle.c) against the independent known-answer
vectors (tests/golden/kats.npz, derived by tests/golden/make_kats.py) and against the
second restatement (oracle/meshgen_oracle_py.py).  CPU only."""
import numpy as np
import pytest

import oracle
from oracle import meshgen_oracle_py as pyo
from tests.util import first_diff, load_golden

G = load_golden()


@pytest.fixture(scope="module")
def cache():
    return oracle.cache_from_entries(G.entries, G.n_known, G.fallback)


@pytest.mark.parametrize("name", G.names)
def test_c_oracle_matches_kat(name, cache):
    c = G.cases[name]
    intr, trans = oracle.mesh_for_chunk(c["offs"], c["chunk"], cache)
    assert first_diff(intr.tobytes(), c["intr"]) is None
    assert first_diff(trans.tobytes(), c["trans"]) is None


SMALL = [n for n in G.names if len(G.cases[n]["intr"]) + len(G.cases[n]["trans"]) < 36 * 2000]


@pytest.mark.parametrize("name", SMALL + ["random_dense_ids"])
def test_python_transliteration_matches_c_oracle(name, cache):
    c = G.cases[name]
    pc = pyo.PyCache(G.entries, G.n_known, G.fallback)
    r, rt = pyo.mesh_for_chunk(c["offs"], c["chunk"], pc)
    intr, trans = oracle.mesh_for_chunk(c["offs"], c["chunk"], cache)
    assert first_diff(pyo.to_structured(r, oracle.VERTEX_DTYPE).tobytes(), intr.tobytes()) is None
    assert first_diff(pyo.to_structured(rt, oracle.VERTEX_DTYPE).tobytes(), trans.tobytes()) is None


def test_hand_written_single_stone_pass1(cache):
    """SURVEY.md §8c vector (2), typed in by hand: single stone (id 5 -> tex 6) at the origin;
    the first quad is pass 1 (bottom face)."""
    c = G.cases["single_stone_origin"]
    intr, _ = oracle.mesh_for_chunk(c["offs"], c["chunk"], cache)
    want_uv = [(0, 0), (1, 0), (0, 1), (1, 0), (1, 1), (0, 1)]
    want_p = [(0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0)]
    for i in range(6):
        v = intr[i]
        assert v["tex_ind"] == 6 and v["pad0"] == 0
        assert tuple(v["tex_pos"]) == want_uv[i]
        assert tuple(v["position"]) == want_p[i]
        assert tuple(v["normal"]) == (1, 1, 0)
    assert len(intr) == 36


def test_counts(cache):
    """quad counts that follow from the algorithm alone."""
    n = lambda name: len(oracle.mesh_for_chunk(G.cases[name]["offs"], G.cases[name]["chunk"], cache)[0])
    assert n("uniform_stone") == 96 * 6
    assert n("checkerboard_even") == 6 * 2048 * 6
    i, t = oracle.mesh_for_chunk(G.cases["one_sapling"]["offs"], G.cases["one_sapling"]["chunk"], cache)
    assert len(i) == 0 and len(t) == 24 and set(t["tex_ind"]) == {30}


def test_grass_block_face_textures(cache):
    """ground_with_grass uses tex 16 bottom, 15 top, 14 sides."""
    c = G.cases["ground_with_grass"]
    intr, _ = oracle.mesh_for_chunk(c["offs"], c["chunk"], cache)
    assert [int(intr[6 * q]["tex_ind"]) for q in range(6)] == [16, 14, 14, 15, 14, 14]


def test_negative_zero_never_appears(cache):
    """zsig(+-0) = +0.0 and run ends such as -16 + 16 give +0.0, so no -0.0 bit pattern
    (0x80000000) may occur in the common cases."""
    for name in ("negative_offs_uniform", "uniform_stone", "sapling_does_not_occlude"):
        c = G.cases[name]
        intr, trans = oracle.mesh_for_chunk(c["offs"], c["chunk"], cache)
        for arr in (intr, trans):
            w = np.frombuffer(arr.tobytes(), dtype=np.uint32)
            assert not (w == 0x80000000).any()


def test_batch_packing_equals_per_chunk(cache):
    names = ["layered_terrain", "all_air", "random_dense_ids", "one_sapling", "uniform_stone"]
    chunks = np.stack([G.cases[n]["chunk"] for n in names])
    offs = np.stack([G.cases[n]["offs"] for n in names])
    for threads in (1, 3):
        b = oracle.mesh_batch(chunks, offs, cache, threads=threads)
        assert b["intransparent"].tobytes() == b"".join(G.cases[n]["intr"] for n in names)
        assert b["transparent"].tobytes() == b"".join(G.cases[n]["trans"] for n in names)
        assert list(b["count_i"]) == [len(G.cases[n]["intr"]) // 36 for n in names]
    d = oracle.mesh_batch(chunks, offs, cache, threads=2, discard=True)
    assert d["vertices"] == sum(len(G.cases[n]["intr"]) + len(G.cases[n]["trans"]) for n in names) // 36
