"""The optional halo mode of the oracle (an extension: the reference meshes every chunk as if surrounded by
air, lib.rs:274-283; with halo slabs a face against a blocky cell of the neighbouring chunk is culled
too).  Checked against an independent numpy count of the visible unit faces of a whole region.  CPU only."""
import numpy as np
import pytest

import oracle
from tests.util import first_diff, load_golden, oracle_cache, region_to_chunks_and_halos

G = load_golden()
NORMALS = {(-1, 0, 0): 0, (1, 0, 0): 1, (0, -1, 0): 2, (0, 1, 0): 3, (0, 0, -1): 4, (0, 0, 1): 5}


@pytest.fixture(scope="module")
def cache():
    return oracle_cache(G)


def blocky_ids():
    """ids whose draw style is blocky (known ids that are not air and not crossed), from the fixture table."""
    rows = G.entries
    return [i for i, r in enumerate(rows) if r[0] == 1]


def visible_faces_numpy(vol, blocky):
    """[X,Y,Z,6] count of unit faces of each chunk that face a non-blocky cell of the REGION (air beyond)."""
    solid = np.isin(vol, blocky)
    pad = np.zeros(tuple(s + 2 for s in vol.shape), dtype=bool)
    pad[1:-1, 1:-1, 1:-1] = solid
    out = []
    for (dx, dy, dz) in NORMALS:
        nb = pad[1 + dx:pad.shape[0] - 1 + dx, 1 + dy:pad.shape[1] - 1 + dy, 1 + dz:pad.shape[2] - 1 + dz]
        vis = solid & ~nb
        X, Y, Z = (s // 16 for s in vol.shape)
        out.append(vis.reshape(X, 16, Y, 16, Z, 16).sum(axis=(1, 3, 5)))
    return np.stack(out, axis=-1)


def face_area_by_normal(verts):
    """Sum of quad areas per outward face direction of a vertex stream of axis-aligned quads.  The `normal`
    field is no use here (it holds the signs of the spanning vectors, lib.rs:133-149); the direction comes
    from the winding: the reference's triangles are clockwise seen from outside, so outward = -(e1 x e2)."""
    area = np.zeros(6)
    p = verts["position"].astype(np.float64).reshape(-1, 6, 3)
    cr = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])
    for c in cr:
        (ax,) = np.nonzero(c)
        n = [0, 0, 0]
        n[int(ax[0])] = -int(np.sign(c[ax[0]]))
        area[NORMALS[tuple(n)]] += abs(c[ax[0]])
    return area


def make_region(seed, dims=(2, 2, 2)):
    rng = np.random.default_rng(seed)
    b = blocky_ids()
    shape = tuple(16 * d for d in dims)
    vol = np.where(rng.random(shape) < 0.55, rng.choice(b, size=shape), 0).astype(np.uint8)
    vol[:, :20, :] = b[0]  # a thick floor: whole chunk faces vanish between chunks
    return vol


def test_halo_of_air_is_the_reference_mode(cache):
    c = G.cases["random_dense_ids"]
    a = oracle.mesh_batch(c["chunk"][None], c["offs"][None], cache)
    h = oracle.mesh_batch(c["chunk"][None], c["offs"][None], cache, halo=np.zeros((1, 6, 256), np.uint8))
    assert first_diff(a["intransparent"].tobytes(), h["intransparent"].tobytes()) is None
    assert first_diff(a["transparent"].tobytes(), h["transparent"].tobytes()) is None


def test_solid_chunk_inside_solid_neighbours_has_no_faces(cache):
    b = blocky_ids()[0]
    ch = np.full((1, 4096), b, np.uint8)
    offs = np.zeros((1, 3), np.int64)
    r = oracle.mesh_batch(ch, offs, cache, halo=np.full((1, 6, 256), b, np.uint8))
    assert r["count_i"][0] == 0 and r["count_t"][0] == 0
    # one open side: exactly the 16 strips of that face come back
    for side in range(6):
        halo = np.full((1, 6, 256), b, np.uint8)
        halo[0, side] = 0
        r = oracle.mesh_batch(ch, offs, cache, halo=halo)
        assert r["count_i"][0] + r["count_t"][0] == 16 * 6
        area = face_area_by_normal(np.concatenate([r["intransparent"], r["transparent"]]))
        want = np.zeros(6)
        want[side] = 256
        assert (area == want).all()


@pytest.mark.parametrize("seed", [1, 2])
def test_region_face_count_matches_numpy(seed, cache):
    vol = make_region(seed)
    chunks, offs, halo = region_to_chunks_and_halos(vol)
    want = visible_faces_numpy(vol, blocky_ids()).reshape(-1, 6)
    r = oracle.mesh_batch(chunks, offs, cache, halo=halo)
    ci, ct = oracle.count_batch(chunks, offs, cache, halo=halo)
    assert list(ci) == list(r["count_i"]) and list(ct) == list(r["count_t"])
    for k in range(chunks.shape[0]):
        vi = r["intransparent"][r["first_i"][k]:r["first_i"][k] + r["count_i"][k]]
        vt = r["transparent"][r["first_t"][k]:r["first_t"][k] + r["count_t"][k]]
        got = face_area_by_normal(np.concatenate([vi, vt]))
        assert (got == want[k]).all(), (k, got, want[k])
    # and without halos every chunk border is a surface: strictly more faces
    r0 = oracle.mesh_batch(chunks, offs, cache)
    assert int(r0["count_i"].sum() + r0["count_t"].sum()) > int(r["count_i"].sum() + r["count_t"].sum())


def test_box_halo_helper_layout():
    """halos_for_box (the numpy meaning of mm_gather_halos): slabs are the neighbours' touching layers."""
    from tests.util import halos_for_box
    rng = np.random.default_rng(0)
    nx, ny, nz = 3, 2, 2
    chunks = rng.integers(0, 256, size=(nx * ny * nz, 4096), dtype=np.uint8)
    halo = halos_for_box(chunks, (nx, ny, nz), outside=9)
    c = chunks.reshape(-1, 16, 16, 16)
    idx = lambda cx, cy, cz: cx + nx * (cy + ny * cz)
    i = idx(1, 0, 1)
    assert (halo[i, 0].reshape(16, 16) == c[idx(0, 0, 1)][15, :, :]).all()   # x-: [y, z]
    assert (halo[i, 1].reshape(16, 16) == c[idx(2, 0, 1)][0, :, :]).all()    # x+
    assert (halo[i, 2] == 9).all()                                            # y-: outside the box
    assert (halo[i, 3].reshape(16, 16) == c[idx(1, 1, 1)][:, 0, :]).all()    # y+: [x, z]
    assert (halo[i, 4].reshape(16, 16) == c[idx(1, 0, 0)][:, :, 15]).all()   # z-: [x, y]
    assert (halo[i, 5] == 9).all()
