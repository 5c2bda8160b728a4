"""GPU parity tests: the CUDA mesher, called through the C ABI (ctypes -> libmimas_meshgen_cuda.so),
must be BIT-EXACT with the CPU oracle (oracle/meshgen_oracle.c, a restatement of
mimas-meshgen/lib.rs:231-446) -- integer / byte work, tolerance zero.  Run with -m gpu on a B200."""
import numpy as np
import pytest

import oracle
from tests.util import (assert_batch_equal, checkerboard_chunks, first_diff, load_golden, oracle_cache,
                        product_cache)

pytestmark = pytest.mark.gpu

G = load_golden()


@pytest.fixture(scope="module")
def mesher():
    import mimas_b200
    m = mimas_b200.Mesher(0, product_cache(G))
    yield m
    m.close()


@pytest.fixture(scope="module")
def ocache():
    return oracle_cache(G)


@pytest.mark.parametrize("name", G.names)
def test_known_answer_vectors_single_chunk(name, mesher):
    """mesh_for_chunk (lib.rs:231-232) on every committed known-answer vector."""
    c = G.cases[name]
    m = mesher.mesh_for_chunk(c["offs"], c["chunk"])
    assert first_diff(m.intransparent.tobytes(), c["intr"]) is None
    assert first_diff(m.transparent.tobytes(), c["trans"]) is None


def test_known_answer_vectors_as_one_batch(mesher):
    """Streams are packed in chunk order: one memcmp against the concatenated vectors."""
    chunks = np.stack([G.cases[n]["chunk"] for n in G.names])
    offs = np.stack([G.cases[n]["offs"] for n in G.names])
    b = mesher.mesh_batch(offs, chunks)
    assert b.intransparent.tobytes() == b"".join(G.cases[n]["intr"] for n in G.names)
    assert b.transparent.tobytes() == b"".join(G.cases[n]["trans"] for n in G.names)
    for i, n in enumerate(G.names):
        assert b.chunk(i).intransparent.tobytes() == G.cases[n]["intr"]
        assert b.chunk(i).transparent.tobytes() == G.cases[n]["trans"]


def test_empty_batch(mesher):
    b = mesher.mesh_batch(np.zeros((0, 3), np.int64), np.zeros((0, 4096), np.uint8))
    assert len(b.intransparent) == 0 and len(b.transparent) == 0 and len(b.ranges_i) == 0


@pytest.mark.parametrize("dims,c0", [((8, 8, 4), (-4, -4, -2)), ((5, 3, 2), (100, -77, -1)), ((1, 1, 1), (3, 2, 1))])
def test_terrain_regions_match_oracle(dims, c0, mesher, ocache):
    import mimas_b200
    chunks, offs = mimas_b200.gen_terrain_host(78, c0, dims)
    assert_batch_equal(mesher.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_random_ids_and_extreme_offsets(mesher, ocache):
    """All 256 ids (unknown ids -> fallback, lib.rs:105-109), crossed plants everywhere, and origins far
    beyond 2^24 where `as f32` and `v - last_v` round (lib.rs:249,256)."""
    rng = np.random.RandomState(7)
    chunks = rng.randint(0, 256, size=(48, 4096)).astype(np.uint8)
    offs = rng.randint(-2 ** 40, 2 ** 40, size=(48, 3)).astype(np.int64)
    offs[0] = (2 ** 62, -2 ** 62, 2 ** 24 - 8)
    offs[1] = (-2 ** 24 - 8, 2 ** 24 - 8, -16)
    assert_batch_equal(mesher.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_sparse_random_chunks(mesher, ocache):
    rng = np.random.RandomState(11)
    ids = np.array([0, 0, 0, 0, 5, 5, 3, 11, 2, 1, 9, 42, 51, 7, 6, 200], dtype=np.uint8)
    chunks = ids[rng.randint(0, len(ids), size=(64, 4096))]
    # long runs too: smear along y and x
    chunks[32:] = np.repeat(chunks[32:].reshape(32, 16, 4, 4, 16)[:, :, :, :1, :], 4, axis=3).reshape(32, 4096)
    offs = (rng.randint(-1000, 1000, size=(64, 3)) * 16).astype(np.int64)
    assert_batch_equal(mesher.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_checkerboard_worst_case_batch(mesher, ocache):
    """cfg3 shape: 12288 quads per chunk -> 12 descriptor rounds and 96 staging tiles per chunk."""
    chunks, offs = checkerboard_chunks(24)
    b = mesher.mesh_batch(offs, chunks)
    assert list(b.ranges_i["vertex_count"]) == [73728] * 24
    assert_batch_equal(b, oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_max_vertex_chunks(mesher, ocache):
    """Absolute maxima: all-crossed chunk (4096 plants = 98304 transparent vertices) and
    the solid/crossed checkerboard (122880 vertices)."""
    a = np.full(4096, 42, dtype=np.uint8)
    b, _ = checkerboard_chunks(1, solid=5, other=48)
    chunks = np.stack([a, b[0], a])
    offs = np.array([[0, 0, 0], [16, 0, 0], [-16, -16, -16]], dtype=np.int64)
    got = mesher.mesh_batch(offs, chunks)
    assert list(got.ranges_t["vertex_count"]) == [98304, 2048 * 24, 98304]
    assert int(got.ranges_i["vertex_count"][1]) + int(got.ranges_t["vertex_count"][1]) == 122880
    assert_batch_equal(got, oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_shared_textures_merge_across_block_ids(ocache):
    """Runs merge on equal *texture ids*, not block ids (Walker compares TextureId, lib.rs:253-258):
    a table where different blocks share some face textures but not others."""
    import mimas_b200
    rows = [(0, 0, 0, 0, 0), (1, 7, 8, 9, 0), (1, 7, 8, 10, 0), (1, 11, 8, 9, 0), (1, 7, 12, 9, 0), (2, 0, 0, 0, 5)]
    fb = (7, 8, 9)  # unknown ids share everything with id 1
    tic = mimas_b200.TextureIdCache.from_table_rows(rows, fb)
    oc = oracle.cache_from_entries(rows, len(rows), fb)
    rng = np.random.RandomState(3)
    chunks = np.array([0, 1, 2, 3, 4, 5, 77], dtype=np.uint8)[rng.randint(0, 7, size=(16, 4096))]
    chunks[8:] = np.array([1, 2, 3, 4, 77], dtype=np.uint8)[rng.randint(0, 5, size=(8, 4096))]
    offs = (rng.randint(-50, 50, size=(16, 3)) * 16).astype(np.int64)
    with mimas_b200.Mesher(0, tic) as m:
        assert_batch_equal(m.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, oc, threads=4))


def test_fuzz_cases_with_random_tables_through_the_c_abi():
    """The differential fuzz's generator (tests/fuzz_util.py) against the CUDA path: 160 random tables (different ids
    sharing textures, fallback and crossed ids) x chunks x origins up to 2^62, eight chunks per table in one batch."""
    import mimas_b200
    from tests import fuzz_util
    with mimas_b200.Mesher(0) as m:
        for t in range(160):
            rng = np.random.RandomState(5000 + t)
            rows, fb = fuzz_util.random_table(rng)
            chunks = np.stack([fuzz_util.random_chunk(rng, len(rows)) for _ in range(8)])
            offs = np.array([[fuzz_util.OFFSETS[i] for i in rng.randint(0, len(fuzz_util.OFFSETS), size=3)] for _ in range(8)],
                            dtype=np.int64)
            m.set_cache(mimas_b200.TextureIdCache.from_table_rows(rows, fb))
            want = oracle.mesh_batch(chunks, offs, oracle.cache_from_entries(rows, len(rows), fb), threads=4)
            assert_batch_equal(m.mesh_batch(offs, chunks), want)
            one = m.mesh_for_chunk(offs[3], chunks[3])          # and the fused single-chunk launch
            a, n = int(want["first_i"][3]), int(want["count_i"][3])
            assert one.intransparent.tobytes() == want["intransparent"][a:a + n].tobytes(), f"table {t}"


def test_repeated_launches_with_changing_batch_sizes(mesher, ocache):
    """Exercises arena growth + re-run and the alternating hand-over buffers across launches of changing size."""
    import mimas_b200
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        for dims in [(2, 2, 1), (6, 6, 2), (1, 1, 1), (7, 5, 3), (3, 3, 3), (7, 5, 3), (2, 1, 1)]:
            chunks, offs = mimas_b200.gen_terrain_host(5, (-3, -2, -1), dims)
            assert_batch_equal(m.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_device_resident_path_and_device_terrain_generator(mesher, ocache):
    import mimas_b200
    dims, c0 = (16, 8, 4), (-8, -4, -2)
    n = dims[0] * dims[1] * dims[2]
    chunks, offs = mimas_b200.gen_terrain_host(78, c0, dims)
    d_c, d_o = mesher.dev_alloc(n * 4096), mesher.dev_alloc(n * 24)
    try:
        mesher.gen_terrain_device(78, c0, dims, d_c, d_o)
        mesher.sync()
        assert mesher.download(d_c, n * 4096).tobytes() == chunks.tobytes()
        assert mesher.download(d_o, n * 24).tobytes() == offs.tobytes()
        want = oracle.mesh_batch(chunks, offs, ocache, threads=8)
        for _ in range(3):  # back-to-back async launches, last one read back
            mesher.mesh_batch_async(d_c, d_o, n)
        out = mesher.batch_finish()
        assert out.n_chunks == n and out.kernel_ms == 0  # async batches enqueue kernels only, no timing events
        assert_batch_equal(mesher.copy_out(out), want)
    finally:
        mesher.dev_free(d_c)
        mesher.dev_free(d_o)


def test_overlapping_batches_do_not_disturb_each_other(ocache):
    """Back-to-back mm_mesh_batch_async batches chain by programmatic dependent launch: the analysis of
    batch k+1 runs beside the expansion of batch k, on the other half of the hand-over buffers, and the
    expansion of k+1 may only touch the arenas once k has finished.  Alternate between batches of very
    different content and size; whichever comes last must be exact."""
    import mimas_b200
    from tests.util import checkerboard_chunks
    rng = np.random.RandomState(33)
    a_c, a_o = mimas_b200.gen_terrain_host(78, (-8, -8, -2), (16, 16, 4))          # 1024 terrain chunks
    b_c = rng.randint(0, 256, size=(700, 4096)).astype(np.uint8)                      # dense random ids, many plants
    b_o = (rng.randint(-1000, 1000, size=(700, 3)) * 16).astype(np.int64)
    c_c, c_o = checkerboard_chunks(40)                                                # few huge chunks
    e_c, e_o = np.zeros((3000, 4096), np.uint8), np.zeros((3000, 3), np.int64)        # nothing at all
    sets = [(a_c, a_o), (b_c, b_o), (c_c, c_o), (e_c, e_o)]
    want = [oracle.mesh_batch(c, o, ocache, threads=8) for c, o in sets]
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        dev = []
        for c, o in sets:
            dc, do = m.dev_alloc(c.size), m.dev_alloc(o.size * 8)
            m.upload(dc, c)
            m.upload(do, o)
            dev.append((dc, do, len(c)))
        m.reserve(4096, 700 * 73728, 700 * 400 * 24)
        order = rng.randint(0, 4, size=60)
        for last in range(4):
            for k in list(order) + [last]:
                m.mesh_batch_async(*dev[k])
            assert_batch_equal(m.copy_out(m.batch_finish()), want[last])
        for dc, do, _ in dev:
            m.dev_free(dc)
            m.dev_free(do)


def test_single_chunk_async_batches_mixed_with_larger_ones(ocache):
    """n == 1 takes the fused single-CTA launch (analyse_kernel<., FUSED>); queued back to back with ordinary
    launch pairs, without a finish in between, it must neither disturb their chunk tickets nor overwrite
    hand-over buffers that an earlier launch still reads.  Whatever comes last must be exact."""
    import mimas_b200
    rng = np.random.RandomState(44)
    big_c, big_o = mimas_b200.gen_terrain_host(78, (-8, -8, -2), (16, 16, 3))          # 768 chunks
    cb_c, cb_o = checkerboard_chunks(3)
    ones = [(np.full((1, 4096), 5, np.uint8), np.array([[16, 32, 48]], np.int64)),
            (cb_c[:1], cb_o[:1]),
            (rng.randint(0, 256, size=(1, 4096)).astype(np.uint8), np.array([[-160, 2 ** 30, 0]], np.int64)),
            (G.cases["layered_terrain"]["chunk"][None], G.cases["layered_terrain"]["offs"][None])]
    sets = ones + [(big_c, big_o), (cb_c, cb_o)]
    want = [oracle.mesh_batch(c, o, ocache, threads=8) for c, o in sets]
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        dev = []
        for c, o in sets:
            dc, do = m.dev_alloc(c.size), m.dev_alloc(o.size * 8)
            m.upload(dc, np.ascontiguousarray(c))
            m.upload(do, np.ascontiguousarray(o))
            dev.append((dc, do, len(c)))
        m.reserve(1024, 3 * 73728 + 768 * 4000, 4096 * 24)
        for trial in range(12):
            seq = list(rng.randint(0, len(sets), size=10))
            seq += [int(rng.randint(0, 4))] * int(rng.randint(1, 4))     # a run of single-chunk batches ...
            last = trial % len(sets)                                     # ... then the batch that is checked
            for k in seq + [last]:
                m.mesh_batch_async(*dev[k])
            assert_batch_equal(m.copy_out(m.batch_finish()), want[last])
        for dc, do, _ in dev:
            m.dev_free(dc)
            m.dev_free(do)


def test_misaligned_device_chunks_are_rejected(mesher):
    """The chunks are fetched with bulk async copies (16-byte aligned source): a misaligned device pointer is
    an argument error, not a device fault that poisons the context."""
    import ctypes as C
    from mimas_b200._lib import lib
    d = mesher.dev_alloc(2 * 4096 + 64)
    try:
        rc = lib().mm_mesh_batch_async(mesher._ctx, C.c_void_p(d + 8), C.c_void_p(d + 4096 + 16), 1)
        assert rc == 1  # MM_ERR_INVALID
        c = G.cases["single_stone_origin"]
        m = mesher.mesh_for_chunk(c["offs"], c["chunk"])      # the context still works
        assert first_diff(m.intransparent.tobytes(), c["intr"]) is None
    finally:
        mesher.dev_free(d)


def test_misaligned_pinned_host_input(mesher, ocache):
    """Pinned host chunks at an odd address: the mapped path copies them instead of reading them in place."""
    buf = mesher.host_alloc(3 * 4096 + 3 * 24 + 64)
    try:
        rng = np.random.RandomState(5)
        chunks = rng.randint(0, 60, size=(3, 4096)).astype(np.uint8)
        offs = np.array([[0, 0, 0], [16, -16, 32], [-2 ** 24, 2 ** 24, 16]], np.int64)
        buf[4:4 + chunks.size] = chunks.reshape(-1)
        o0 = 4 + chunks.size + 4   # 8-byte aligned offset for the origins, chunks at +4
        buf[o0:o0 + 72] = offs.reshape(-1).view(np.uint8)
        hm = mesher.mesh_batch_host_raw(buf.ctypes.data + 4, buf.ctypes.data + o0, 3)
        want = oracle.mesh_batch(chunks, offs, ocache, threads=2)
        import ctypes as C
        got = bytes((C.c_uint8 * (36 * hm.n_intransparent)).from_address(hm.intransparent)) if hm.n_intransparent else b""
        assert first_diff(got, want["intransparent"].tobytes()) is None
    finally:
        mesher.host_free(buf)


def test_full_size_terrain_batch_4096(mesher, ocache):
    """BASELINE.json configs[1] at full size: 4096 terrain chunks, every byte compared."""
    import mimas_b200
    chunks, offs = mimas_b200.gen_terrain_host(78, (-16, -16, -2), (32, 32, 4))
    assert_batch_equal(mesher.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_batch_larger_than_one_launch(mesher, ocache):
    """More chunks than one analyse/expand launch pair takes (4096): the second pair continues the packed
    streams where the first ended (carry), across runs of empty chunks and a partial last tile."""
    import mimas_b200
    chunks, offs = mimas_b200.gen_terrain_host(78, (-16, -16, -2), (32, 32, 4))
    rng = np.random.RandomState(21)
    extra = np.zeros((1500, 4096), np.uint8)
    extra[::7] = np.array([0, 0, 0, 5, 3, 42, 11, 200], np.uint8)[rng.randint(0, 8, size=(len(extra[::7]), 4096))]
    chunks = np.concatenate([chunks[1024:], extra, chunks[:1024]])
    eo = (rng.randint(-500, 500, size=(1500, 3)) * 16).astype(np.int64)
    offs = np.concatenate([offs[1024:], eo, offs[:1024]])
    assert len(chunks) == 5596
    assert_batch_equal(mesher.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, ocache, threads=8))


def test_pipelined_host_path_pieces_overflow_and_sizes(ocache):
    """Host batches of >= 1024 chunks are cut into pieces whose copies and kernels overlap
    (pipelined_batch_host); the bytes are those of the plain path, also when the arenas overflow
    half-way (fall back, grow, run again) and for sizes that are no multiple of the piece."""
    import mimas_b200
    rng = np.random.RandomState(9)
    tc, to = mimas_b200.gen_terrain_host(31, (-8, -8, -2), (16, 16, 5))
    cc, co = checkerboard_chunks(600)
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        # arenas sized by a small sparse batch first, then a dense one that overflows them in a later piece
        sparse = np.zeros((1100, 4096), np.uint8)
        sparse[:, 7] = 5
        assert_batch_equal(m.mesh_batch(to[:1100], sparse), oracle.mesh_batch(sparse, to[:1100], ocache, threads=8))
        chunks = np.concatenate([sparse[:900], cc, tc[:777]])
        offs = np.concatenate([to[:900], co, to[:777]])
        want = oracle.mesh_batch(chunks, offs, ocache, threads=8)
        assert_batch_equal(m.mesh_batch(offs, chunks), want)   # overflow -> plain path
        assert_batch_equal(m.mesh_batch(offs, chunks), want)   # now pipelined
        for n in (1024, 1025, 1537, 2277):
            sel = rng.choice(len(chunks), n, replace=n > len(chunks))
            assert_batch_equal(m.mesh_batch(offs[sel], chunks[sel]),
                               oracle.mesh_batch(chunks[sel], offs[sel], ocache, threads=8))


def test_push_block_matches_oracle(mesher):
    """push_block (lib.rs:183-219) for every blocked-face combination."""
    from mimas_b200.meshgen import BlockTextureIds
    for mask in range(64):
        order = ([0, 0, -1], [0, -1, 0], [-1, 0, 0], [0, 0, 1], [0, 1, 0], [1, 0, 0])
        blocked = lambda o, mask=mask: bool(mask >> order.index(list(o)) & 1)
        got = mesher.push_block((1.5, -2.25, 1e7), BlockTextureIds(3, 4, 5), 0.5, blocked)
        want = oracle.push_block((1.5, -2.25, 1e7), (3, 4, 5), 0.5, mask)
        assert first_diff(got.tobytes(), want.tobytes()) is None


# ---- halo mode: an extension (the reference has no halo, lib.rs:277-280); parity is against the oracle's
# `blocked` with halo slabs, which tests/test_halo.py checks against a numpy face count ---------------------

def test_halo_of_air_is_the_reference_mode(mesher, ocache):
    chunks = np.stack([G.cases[n]["chunk"] for n in G.names])
    offs = np.stack([G.cases[n]["offs"] for n in G.names])
    got = mesher.mesh_batch(offs, chunks, halo=np.zeros((len(chunks), 6, 256), np.uint8))
    assert got.intransparent.tobytes() == b"".join(G.cases[n]["intr"] for n in G.names)
    assert got.transparent.tobytes() == b"".join(G.cases[n]["trans"] for n in G.names)


def test_halo_random_slabs_match_oracle(mesher, ocache):
    rng = np.random.RandomState(21)
    n = 300
    chunks = rng.randint(0, 256, size=(n, 4096)).astype(np.uint8)
    chunks[rng.rand(n, 4096) < 0.3] = 0
    chunks[:40] = 5  # solid chunks: only what the halo leaves open comes back
    halo = rng.randint(0, 256, size=(n, 6, 256)).astype(np.uint8)
    halo[rng.rand(n, 6, 256) < 0.5] = 0
    halo[:8] = 5
    offs = (rng.randint(-2 ** 30, 2 ** 30, size=(n, 3)) * 16).astype(np.int64)
    got = mesher.mesh_batch(offs, chunks, halo=halo)
    assert_batch_equal(got, oracle.mesh_batch(chunks, offs, ocache, threads=8, halo=halo))
    assert list(got.ranges_i["vertex_count"][:8]) == [0] * 8


def test_halo_terrain_box_gathered_on_device(mesher, ocache):
    """mm_gather_halos on a resident box = the numpy gather; the halo mesh of the box = the oracle's, and it
    has fewer faces than the reference mode (the chunk borders below ground vanish)."""
    import mimas_b200
    from tests.util import halos_for_box
    dims, c0 = (12, 10, 5), (-6, -5, -2)
    n = dims[0] * dims[1] * dims[2]
    chunks, offs = mimas_b200.gen_terrain_host(78, c0, dims)
    halo = halos_for_box(chunks, dims, outside=3)
    d_c, d_o, d_h = mesher.dev_alloc(n * 4096), mesher.dev_alloc(n * 24), mesher.dev_alloc(n * 1536)
    try:
        mesher.upload(d_c, chunks)
        mesher.upload(d_o, offs)
        mesher.gather_halos(d_c, dims, d_h, outside_id=3)
        mesher.sync()
        assert mesher.download(d_h, n * 1536).tobytes() == halo.tobytes()
        want = oracle.mesh_batch(chunks, offs, ocache, threads=8, halo=halo)
        for _ in range(2):
            mesher.mesh_batch_async(d_c, d_o, n, d_halo=d_h)
        got = mesher.copy_out(mesher.batch_finish())
        assert_batch_equal(got, want)
        plain = mesher.copy_out(mesher.mesh_batch_device(d_c, d_o, n))
        assert len(plain.intransparent) > len(got.intransparent)
        # a second halo batch after a plain one (the two analyse kernels alternate on one stream)
        assert_batch_equal(mesher.copy_out(mesher.mesh_batch_device(d_c, d_o, n, d_halo=d_h)), want)
    finally:
        for p in (d_c, d_o, d_h):
            mesher.dev_free(p)


def test_halo_batch_larger_than_one_launch_and_misaligned_halo(mesher, ocache):
    import ctypes as C
    from mimas_b200._lib import lib
    from tests.util import checkerboard_chunks
    chunks, offs = checkerboard_chunks(4500)
    rng = np.random.RandomState(2)
    halo = np.where(rng.rand(4500, 6, 256) < 0.5, 5, 0).astype(np.uint8)
    got = mesher.mesh_batch(offs, chunks, halo=halo)
    assert_batch_equal(got, oracle.mesh_batch(chunks, offs, ocache, threads=8, halo=halo))
    d = mesher.dev_alloc(4096 + 24 + 1536 + 64)
    try:
        rc = lib().mm_mesh_batch_halo_async(mesher._ctx, C.c_void_p(d), C.c_void_p(d + 4096), C.c_void_p(d + 4096 + 32 + 8), 1)
        assert rc == 1  # MM_ERR_INVALID: device halos are read 16 bytes at a time
    finally:
        mesher.dev_free(d)
