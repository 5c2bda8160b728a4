"""Shared helpers for the test-suite (fixtures, byte comparison)."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

VERTEX_DTYPE = np.dtype([("tex_ind", "<u2"), ("pad0", "<u2"), ("tex_pos", "<f4", (2,)),
                         ("position", "<f4", (3,)), ("normal", "<f4", (3,))])


class Golden:
    def __init__(self):
        self.table = json.load(open(os.path.join(GOLDEN, "default_block_table.json")))
        z = np.load(os.path.join(GOLDEN, "kats.npz"))
        self.names = [str(n) for n in z["names"]]
        self.rows = [tuple(int(v) for v in r) for r in z["rows"]]
        self.cases = {n: dict(chunk=z[n + "/chunk"], offs=z[n + "/offs"],
                              intr=z[n + "/intr"].tobytes(), trans=z[n + "/trans"].tobytes())
                      for n in self.names}

    @property
    def n_known(self):
        return self.table["n_ids"]

    @property
    def fallback(self):
        return tuple(self.table["fallback"])

    @property
    def entries(self):
        return [tuple(r) for r in self.table["entries"]]


def load_golden():
    return Golden()


def first_diff(a: bytes, b: bytes):
    """Human-readable description of the first differing vertex of two vertex streams."""
    if len(a) != len(b):
        return f"length {len(a)} != {len(b)} bytes ({len(a) // 36} vs {len(b) // 36} vertices)"
    va = np.frombuffer(a, dtype=np.uint8).reshape(-1, 36)
    vb = np.frombuffer(b, dtype=np.uint8).reshape(-1, 36)
    bad = np.nonzero((va != vb).any(axis=1))[0]
    if len(bad) == 0:
        return None
    i = int(bad[0])
    sa = np.frombuffer(va[i].tobytes(), dtype=VERTEX_DTYPE)[0]
    sb = np.frombuffer(vb[i].tobytes(), dtype=VERTEX_DTYPE)[0]
    return f"{len(bad)} vertices differ; first at {i} (quad {i // 6}, corner {i % 6}): {sa} != {sb}"


# ---- helpers shared by the CPU-emulation and GPU parity tests ---------------------------

def oracle_cache(g=None):
    import oracle
    g = g or load_golden()
    return oracle.cache_from_entries(g.entries, g.n_known, g.fallback)


def product_cache(g=None):
    from mimas_b200.meshgen import TextureIdCache
    g = g or load_golden()
    return TextureIdCache.from_table_rows(g.entries, g.fallback)


def checkerboard_chunks(n, solid=5, other=0):
    """(x+y+z+parity)&1 ? solid : other, parity alternating per chunk,
    origins on a 32 x 32 x k grid of chunks."""
    idx = np.arange(4096)
    x, y, z = idx >> 8, (idx >> 4) & 15, idx & 15
    base = (x + y + z) & 1
    chunks = np.empty((n, 4096), dtype=np.uint8)
    offs = np.empty((n, 3), dtype=np.int64)
    for i in range(n):
        chunks[i] = np.where((base ^ (i & 1)) == 1, solid, other)
        offs[i] = (16 * (i % 32 - 16), 16 * ((i // 32) % 32 - 16), 16 * (i // 1024))
    return chunks, offs


def assert_batch_equal(got, want):
    """got: mimas_b200.BatchMesh; want: dict from oracle.mesh_batch."""
    assert list(got.ranges_i["vertex_count"]) == list(want["count_i"])
    assert list(got.ranges_t["vertex_count"]) == list(want["count_t"])
    assert list(got.ranges_i["first_vertex"]) == list(want["first_i"])
    assert list(got.ranges_t["first_vertex"]) == list(want["first_t"])
    d = first_diff(got.intransparent.tobytes(), want["intransparent"].tobytes())
    assert d is None, "intransparent: " + d
    d = first_diff(got.transparent.tobytes(), want["transparent"].tobytes())
    assert d is None, "transparent: " + d


# ---- halo extension (no reference counterpart; see oracle/meshgen_oracle.c blocked()) ------------

def region_to_chunks_and_halos(vol, outside=0):
    """vol: uint8 [16*X, 16*Y, 16*Z] block ids.  Returns (chunks [n,4096], offs [n,3], halo [n,6,256]) with
    chunk order cx-major, each halo slab the touching layer of the face neighbour (order x-,x+,y-,y+,z-,z+;
    x slabs [y,z], y slabs [x,z], z slabs [x,y]); `outside` fills what lies beyond the region."""
    X, Y, Z = (s // 16 for s in vol.shape)
    pad = np.full(tuple(s + 2 for s in vol.shape), outside, dtype=np.uint8)
    pad[1:-1, 1:-1, 1:-1] = vol
    chunks, offs, halos = [], [], []
    for cx in range(X):
        for cy in range(Y):
            for cz in range(Z):
                x0, y0, z0 = 16 * cx + 1, 16 * cy + 1, 16 * cz + 1
                chunks.append(pad[x0:x0 + 16, y0:y0 + 16, z0:z0 + 16].reshape(4096))
                offs.append((16 * cx, 16 * cy, 16 * cz))
                halos.append(np.stack([
                    pad[x0 - 1, y0:y0 + 16, z0:z0 + 16], pad[x0 + 16, y0:y0 + 16, z0:z0 + 16],
                    pad[x0:x0 + 16, y0 - 1, z0:z0 + 16], pad[x0:x0 + 16, y0 + 16, z0:z0 + 16],
                    pad[x0:x0 + 16, y0:y0 + 16, z0 - 1], pad[x0:x0 + 16, y0:y0 + 16, z0 + 16]]).reshape(6, 256))
    return (np.ascontiguousarray(np.stack(chunks)), np.array(offs, dtype=np.int64),
            np.ascontiguousarray(np.stack(halos)))


def halos_for_box(chunks, dims, outside=0):
    """Halo slabs [n, 6, 256] of a box of dims = (nx, ny, nz) chunks in gen_terrain order (chunk i at
    (i % nx, (i // nx) % ny, i // (nx*ny))) -- the numpy meaning of mm_gather_halos."""
    nx, ny, nz = dims
    c = np.asarray(chunks, dtype=np.uint8).reshape(nz, ny, nx, 16, 16, 16)  # [cz, cy, cx, x, y, z]
    vol = c.transpose(2, 3, 1, 4, 0, 5).reshape(nx * 16, ny * 16, nz * 16)
    _, _, halo = region_to_chunks_and_halos(vol, outside)                     # cx-major order
    halo = halo.reshape(nx, ny, nz, 6, 256).transpose(2, 1, 0, 3, 4)
    return np.ascontiguousarray(halo.reshape(nx * ny * nz, 6, 256))
