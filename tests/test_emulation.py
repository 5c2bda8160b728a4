"""CPU check of the kernel's per-thread phase functions (mimas_b200/csrc/mesh_core.cuh), compiled
with g++ by tests/emul/ and run as a serial emulation of one CTA, against the oracle.  This pins the
bit-level logic the CUDA kernel is built from without needing a GPU; the GPU tests then cover the
real kernels (tickets, barriers, the shared-memory prefix of the expand kernel, bulk loads and stores)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle
from tests.util import checkerboard_chunks, first_diff, load_golden, oracle_cache

HERE = os.path.dirname(os.path.abspath(__file__))
G = load_golden()


def kernel_table(rows, fallback):
    """rows (kind, sides, top, bottom, crossed) -> the packed 2 x u32 entries of mesh_core.cuh TabEntry,
    ids beyond the known ones = blocky fallback (what mm_table_from_cache + mm_set_block_table build)."""
    tab = np.zeros((256, 2), dtype=np.uint32)
    for i in range(256):
        kind, s, t, b, x = rows[i] if i < len(rows) else (1, fallback[0], fallback[1], fallback[2], 0)
        if kind == 2:
            tab[i, 0] = x
        elif kind == 1:
            tab[i, 0] = s | (t << 16)
        tab[i, 1] = (b | (kind << 16)) if kind else 0
    return tab


@pytest.fixture(scope="module")
def emul():
    subprocess.run(["make", "-C", os.path.join(HERE, "emul"), "-s"], check=True)
    lib = C.CDLL(os.path.join(HERE, "emul", "libemul_mesher.so"))

    def run(chunks, offs, tab, cap_quads=None, halo=None):
        chunks = np.ascontiguousarray(chunks, dtype=np.uint8).reshape(-1, 4096)
        n = len(chunks)
        offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(n, 3)
        cap = cap_quads if cap_quads is not None else 20480 * n
        bufs = [np.zeros(cap * 216 + 48, dtype=np.uint8) for _ in range(2)]
        ptrs = [b.ctypes.data + (-b.ctypes.data) % 16 for b in bufs]
        ri = np.zeros((n, 2), dtype=np.uint64)
        rt = np.zeros((n, 2), dtype=np.uint64)
        tot = np.zeros(2, dtype=np.uint64)
        args = [C.c_void_p(chunks.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_uint32(n),
                C.c_void_p(tab.ctypes.data), C.c_void_p(ptrs[0]), C.c_uint64(cap), C.c_void_p(ptrs[1]),
                C.c_uint64(cap), C.c_void_p(ri.ctypes.data), C.c_void_p(rt.ctypes.data), C.c_void_p(tot.ctypes.data)]
        if halo is None:
            lib.emul_mesh_batch(*args)
        else:
            hb = np.zeros(n * 1536 + 16, dtype=np.uint8)
            ho = (-hb.ctypes.data) % 16
            hb[ho:ho + n * 1536] = np.ascontiguousarray(halo, dtype=np.uint8).reshape(-1)
            lib.emul_mesh_batch_halo(*args, C.c_void_p(hb.ctypes.data + ho))
        outs = []
        for b, p, t in zip(bufs, ptrs, tot):
            o = p - b.ctypes.data
            outs.append(b[o:o + int(t) * 216].tobytes())
        return outs[0], outs[1], ri, rt
    run.lib = lib
    return run


def test_bit_transpose_and_equality_mask(emul):
    """The two SWAR building blocks of phases A and C against their naive meaning."""
    rng = np.random.RandomState(1)
    for _ in range(200):
        rows = rng.randint(0, 65536, size=16).astype(np.uint16)
        want = np.zeros(16, dtype=np.uint16)
        for r in range(16):
            for c in range(16):
                want[c] |= ((int(rows[r]) >> c) & 1) << r
        got = rows.copy()
        emul.lib.emul_transpose16(C.c_void_p(got.ctypes.data))
        assert list(got) == list(want)
    emul.lib.emul_eq_mask16.restype = C.c_uint32
    for _ in range(500):
        a = rng.randint(0, 256, size=16).astype(np.uint8)
        b = a.copy()
        flip = rng.rand(16) < rng.rand()
        b[flip] = rng.randint(0, 256, size=int(flip.sum()))
        b[0] ^= rng.choice([0, 0x80, 0x01, 0xff])
        want = sum(1 << z for z in range(16) if a[z] == b[z])
        assert emul.lib.emul_eq_mask16(C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data)) == want


TAB = kernel_table(G.entries, G.fallback)


@pytest.mark.parametrize("name", G.names)
def test_emulated_kernel_matches_kat(name, emul):
    c = G.cases[name]
    i, t, _, _ = emul(c["chunk"], c["offs"], TAB)
    assert first_diff(i, c["intr"]) is None
    assert first_diff(t, c["trans"]) is None


def test_emulated_batch_packing_and_ranges(emul):
    chunks = np.stack([G.cases[n]["chunk"] for n in G.names])
    offs = np.stack([G.cases[n]["offs"] for n in G.names])
    i, t, ri, rt = emul(chunks, offs, TAB)
    assert i == b"".join(G.cases[n]["intr"] for n in G.names)
    assert t == b"".join(G.cases[n]["trans"] for n in G.names)
    want = oracle.mesh_batch(chunks, offs, oracle_cache(G))
    assert list(ri[:, 0]) == list(want["first_i"]) and list(rt[:, 0]) == list(want["first_t"])
    assert list(ri[:, 1] & 0xFFFFFFFF) == list(want["count_i"])


def test_emulated_random_and_terrain(emul):
    import mimas_b200
    oc = oracle_cache(G)
    rng = np.random.RandomState(5)
    chunks = rng.randint(0, 256, size=(12, 4096)).astype(np.uint8)
    offs = rng.randint(-2 ** 40, 2 ** 40, size=(12, 3)).astype(np.int64)
    tc, to = mimas_b200.gen_terrain_host(78, (-3, -3, -2), (6, 6, 4))
    cc, co = checkerboard_chunks(3)
    chunks = np.concatenate([chunks, tc, cc])
    offs = np.concatenate([offs, to, co])
    want = oracle.mesh_batch(chunks, offs, oc, threads=4)
    i, t, _, _ = emul(chunks, offs, TAB)
    assert first_diff(i, want["intransparent"].tobytes()) is None
    assert first_diff(t, want["transparent"].tobytes()) is None


def test_emulated_shared_textures(emul):
    rows = [(0, 0, 0, 0, 0), (1, 7, 8, 9, 0), (1, 7, 8, 10, 0), (1, 11, 8, 9, 0), (1, 7, 12, 9, 0), (2, 0, 0, 0, 5)]
    fb = (7, 8, 9)
    rng = np.random.RandomState(3)
    chunks = np.array([0, 1, 2, 3, 4, 5, 77], dtype=np.uint8)[rng.randint(0, 7, size=(8, 4096))]
    chunks[4:] = np.array([1, 2, 3, 4, 77], dtype=np.uint8)[rng.randint(0, 5, size=(4, 4096))]
    offs = (rng.randint(-50, 50, size=(8, 3)) * 16).astype(np.int64)
    want = oracle.mesh_batch(chunks, offs, oracle.cache_from_entries(rows, len(rows), fb))
    i, t, _, _ = emul(chunks, offs, kernel_table(rows, fb))
    assert first_diff(i, want["intransparent"].tobytes()) is None
    assert first_diff(t, want["transparent"].tobytes()) is None


# ---- halo mode (extension; oracle: blocked() with halo slabs) ----------------------------------------

def test_emulated_halo_of_air_equals_reference_mode(emul):
    c = G.cases["random_dense_ids"]
    a = emul(c["chunk"], c["offs"], TAB)
    h = emul(c["chunk"], c["offs"], TAB, halo=np.zeros((1, 6, 256), np.uint8))
    assert a[0] == h[0] and a[1] == h[1]
    assert first_diff(h[0], c["intr"]) is None


def test_emulated_halo_regions_match_oracle(emul):
    from tests.util import region_to_chunks_and_halos
    oc = oracle_cache(G)
    rng = np.random.RandomState(11)
    # a region of every kind of id (blocky, crossed, undrawn, unknown) and random halos around random chunks
    vol = rng.randint(0, 70, size=(32, 32, 32)).astype(np.uint8)
    vol[rng.rand(32, 32, 32) < 0.4] = 0
    chunks, offs, halo = region_to_chunks_and_halos(vol, outside=3)
    rc = rng.randint(0, 256, size=(6, 4096)).astype(np.uint8)
    rh = rng.randint(0, 256, size=(6, 6, 256)).astype(np.uint8)
    rh[rng.rand(6, 6, 256) < 0.5] = 0
    ro = (rng.randint(-2 ** 30, 2 ** 30, size=(6, 3)) * 16).astype(np.int64)
    chunks = np.concatenate([chunks, rc])
    offs = np.concatenate([offs, ro])
    halo = np.concatenate([halo, rh])
    want = oracle.mesh_batch(chunks, offs, oc, threads=4, halo=halo)
    i, t, ri, rt = emul(chunks, offs, TAB, halo=halo)
    assert first_diff(i, want["intransparent"].tobytes()) is None
    assert first_diff(t, want["transparent"].tobytes()) is None
    assert list(ri[:, 1] & 0xFFFFFFFF) == list(want["count_i"])


def test_emulated_halo_shared_textures(emul):
    rows = [(0, 0, 0, 0, 0), (1, 7, 8, 9, 0), (1, 7, 8, 10, 0), (1, 11, 8, 9, 0), (1, 7, 12, 9, 0), (2, 0, 0, 0, 5)]
    fb = (7, 8, 9)
    rng = np.random.RandomState(4)
    pal = np.array([0, 1, 2, 3, 4, 5, 77], dtype=np.uint8)
    chunks = pal[rng.randint(0, 7, size=(8, 4096))]
    halo = pal[rng.randint(0, 7, size=(8, 6, 256))]
    offs = (rng.randint(-50, 50, size=(8, 3)) * 16).astype(np.int64)
    want = oracle.mesh_batch(chunks, offs, oracle.cache_from_entries(rows, len(rows), fb), halo=halo)
    i, t, _, _ = emul(chunks, offs, kernel_table(rows, fb), halo=halo)
    assert first_diff(i, want["intransparent"].tobytes()) is None
    assert first_diff(t, want["transparent"].tobytes()) is None


def test_emulated_kernel_on_fuzz_cases_with_random_tables(emul):
    """The per-lane kernel logic on the differential fuzz's cases (tests/fuzz_util.py): random tables in which
    different block ids share textures, fallback ids, crossed ids, origins up to 2^62."""
    from tests import fuzz_util
    for seed in range(1000, 1096):
        rows, fb, chunk, offs = fuzz_util.make_case(seed)
        wi, wt = oracle.mesh_for_chunk(offs, chunk, oracle.cache_from_entries(rows, len(rows), fb))
        i, t, _, _ = emul(chunk, offs, kernel_table(rows, fb))
        assert first_diff(i, wi.tobytes()) is None, f"seed {seed}"
        assert first_diff(t, wt.tobytes()) is None, f"seed {seed}"
