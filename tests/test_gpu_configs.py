"""GPU parity tests, through the C ABI, shaped like the BASELINE.json configs that tests/test_gpu_parity.py does
not cover at their own shape, plus the output ring (several batches retrievable while later ones are meshed).
Bit-exact against the CPU oracle (tolerance zero: integer / byte work and exactly rounded f32 adds).

  config 1  single chunk at the spawn position, fused single-CTA launch and the two-kernel pair
  config 3  worst-case checkerboard chunks, >= 512 of them fully byte-compared
  config 4  view radius 8 (1445 chunks) full mesh, then single-block edits each remeshing exactly its chunk --
            the reference's trigger: MapBlockHandle::set -> on_change -> meshgen thread
            (mimas-common/map.rs:128-134, mimas-client/client.rs:224-226,946-950)
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from tests.util import assert_batch_equal, checkerboard_chunks, first_diff, load_golden, oracle_cache, product_cache

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
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


# ---- config 1 ---------------------------------------------------------------------------------------------
SPAWN_CASES = [((3, 2, 1), (48, 32, 16)), ((3, 2, 0), (48, 32, 0))]   # btchn of the default spawn (60, 40, 20), and the chunk below


@pytest.mark.parametrize("cpos,offs", SPAWN_CASES)
def test_config1_single_chunk_fused_launch(cpos, offs, mesher, ocache):
    import mimas_b200
    chunks, o = mimas_b200.gen_terrain_host(78, cpos, (1, 1, 1))
    assert tuple(o[0]) == offs
    want_i, want_t = oracle.mesh_for_chunk(o[0], chunks[0], ocache)
    l0 = mesher.launch_count()
    got = mesher.mesh_for_chunk(o[0], chunks[0])
    assert mesher.launch_count() - l0 == 1          # one chunk = ONE launch: analyse_kernel<., FUSED>
    assert first_diff(got.intransparent.tobytes(), want_i.tobytes()) is None
    assert first_diff(got.transparent.tobytes(), want_t.tobytes()) is None


def test_config1_single_chunk_without_the_fused_launch(ocache):
    """MM_NO_FUSED_SINGLE=1 is read when a context is created: a fresh process takes the two-kernel pair."""
    code = r"""
import sys, numpy as np
sys.path.insert(0, %r)
import mimas_b200, oracle
from tests.util import load_golden, oracle_cache, product_cache, first_diff
g = load_golden(); oc = oracle_cache(g)
with mimas_b200.Mesher(0, product_cache(g)) as m:
    for cpos in ((3, 2, 1), (3, 2, 0)):
        chunks, o = mimas_b200.gen_terrain_host(78, cpos, (1, 1, 1))
        wi, wt = oracle.mesh_for_chunk(o[0], chunks[0], oc)
        l0 = m.launch_count()
        got = m.mesh_for_chunk(o[0], chunks[0])
        assert m.launch_count() - l0 == 2, m.launch_count() - l0
        assert first_diff(got.intransparent.tobytes(), wi.tobytes()) is None
        assert first_diff(got.transparent.tobytes(), wt.tobytes()) is None
    for name in ("layered_terrain", "random_dense_ids", "checkerboard_odd", "one_sapling", "all_air"):
        c = g.cases[name]
        got = m.mesh_for_chunk(c["offs"], c["chunk"])
        assert first_diff(got.intransparent.tobytes(), c["intr"]) is None
        assert first_diff(got.transparent.tobytes(), c["trans"]) is None
print("pair ok")
""" % ROOT
    env = dict(os.environ, MM_NO_FUSED_SINGLE="1")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "pair ok" in r.stdout, r.stdout + r.stderr


# ---- config 3 ---------------------------------------------------------------------------------------------
def test_config3_checkerboard_576_chunks_every_byte(mesher, ocache):
    """576 worst-case chunks (12288 quads each, 1.53 GB of vertices): more quads than a 32-bit byte offset holds,
    every byte compared."""
    chunks, offs = checkerboard_chunks(576)
    got = mesher.mesh_batch(offs, chunks)
    assert list(got.ranges_i["vertex_count"]) == [73728] * 576
    assert_batch_equal(got, oracle.mesh_batch(chunks, offs, ocache, threads=os.cpu_count() or 1))


# ---- config 4 ---------------------------------------------------------------------------------------------
def test_config4_view_radius8_then_edits(mesher, ocache):
    import mimas_b200
    r = 8
    dims = (2 * r + 1, 2 * r + 1, 5)                      # z radius 2 (mimas-common/config.rs:39)
    chunks, offs = mimas_b200.gen_terrain_host(78, (-r, -r, -2), dims)
    assert len(chunks) == 1445
    threads = os.cpu_count() or 1
    assert_batch_equal(mesher.mesh_batch(offs, chunks), oracle.mesh_batch(chunks, offs, ocache, threads=threads))
    # 256 single-block edits: dig (-> air) where there is a block, place stone where there is none
    # (client.rs:946-950 / 1023-1024); each edit remeshes exactly its own chunk, in full, no neighbour
    rng = np.random.RandomState(44)
    surface = [i for i in range(len(chunks)) if 0 < np.count_nonzero(chunks[i]) < 4096]
    ec = np.concatenate([rng.choice(surface, 160, replace=False), rng.choice(len(chunks), 96, replace=False)])
    cells = rng.randint(0, 4096, size=len(ec))
    world = chunks.copy()
    edited_ids = []
    for j, (ci, cell) in enumerate(zip(ec, cells)):
        world[ci, cell] = 0 if world[ci, cell] != 0 else 5
        edited_ids.append(int(ci))
        if j < 48:   # one mesh_for_chunk call per edit, as the reference's meshgen thread does (client.rs:259-267)
            wi, wt = oracle.mesh_for_chunk(offs[ci], world[ci], ocache)
            got = mesher.mesh_for_chunk(offs[ci], world[ci])
            assert first_diff(got.intransparent.tobytes(), wi.tobytes()) is None, f"edit {j}"
            assert first_diff(got.transparent.tobytes(), wt.tobytes()) is None, f"edit {j}"
    # the same edits remeshed as host batches of 16 and 256 (mm_mesh_batch_host)
    for bs in (16, 256):
        for s in range(0, len(ec), bs):
            ids = edited_ids[s:s + bs]
            c, o = np.ascontiguousarray(world[ids]), np.ascontiguousarray(offs[ids])
            assert_batch_equal(mesher.mesh_batch(o, c), oracle.mesh_batch(c, o, ocache, threads=threads))
    # and the whole edited region once more
    assert_batch_equal(mesher.mesh_batch(offs, world), oracle.mesh_batch(world, offs, ocache, threads=threads))


# ---- output ring ------------------------------------------------------------------------------------------
def _ring_inputs(m):
    import mimas_b200
    rng = np.random.RandomState(77)
    sets = []
    for k in range(10):
        kind = k % 5
        if kind == 0:
            c, o = mimas_b200.gen_terrain_host(78, (-8 + 3 * k, -8, -2), (12, 8, 4))
        elif kind == 1:
            c = rng.randint(0, 256, size=(90 + 7 * k, 4096)).astype(np.uint8)
            o = (rng.randint(-1000, 1000, size=(len(c), 3)) * 16).astype(np.int64)
        elif kind == 2:
            c, o = checkerboard_chunks(6 + k)
        elif kind == 3:
            c, o = np.zeros((500, 4096), np.uint8), np.zeros((500, 3), np.int64)
            c[::50] = 42
        else:
            c = G.cases["layered_terrain"]["chunk"][None].copy()
            o = G.cases["layered_terrain"]["offs"][None].copy()
        dc, do = m.dev_alloc(c.size), m.dev_alloc(o.size * 8)
        m.upload(dc, np.ascontiguousarray(c))
        m.upload(do, np.ascontiguousarray(o))
        sets.append((c, o, dc, do))
    return sets


def test_ring_of_three_streams_thirty_batches_all_retrievable(ocache):
    """30 different batches through a ring of 3 slots: each is fetched (mm_batch_wait + copy) while the two behind
    it are already queued, and ALL 30 are byte-compared -- the streamed rate is consumable."""
    import mimas_b200
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        sets = _ring_inputs(m)
        want = [oracle.mesh_batch(c, o, ocache, threads=8) for c, o, _, _ in sets]
        m.ring_configure(3)
        order = [(7 * i + i // 3) % len(sets) for i in range(30)]
        tickets = []
        for i, k in enumerate(order):
            c, o, dc, do = sets[k]
            tickets.append(m.submit(dc, do, len(c)))
            if i >= 2:   # two batches stay in flight behind the one that is read
                j = i - 2
                assert_batch_equal(m.copy_out_ticket(tickets[j]), want[order[j]])
        for j in (28, 29):
            assert_batch_equal(m.copy_out_ticket(tickets[j]), want[order[j]])
        assert tickets == list(range(tickets[0], tickets[0] + 30))
        # a ticket whose slot has been reused is gone, and says so
        from mimas_b200._lib import MMError
        with pytest.raises(MMError) as e:
            m.batch_wait(tickets[0])
        assert e.value.code == 6  # MM_ERR_NO_BATCH
        # small arenas: every slot overflows on first use, grows, runs again
        m.ring_configure(2)
        for k in (2, 0, 5, 7):
            c, o, dc, do = sets[k]
            t = m.submit(dc, do, len(c))
            assert_batch_equal(m.copy_out_ticket(t), want[k])
        for _, _, dc, do in sets:
            m.dev_free(dc)
            m.dev_free(do)


def test_many_batches_by_one_call_and_draw_of_an_older_ticket(ocache):
    """mm_mesh_batches_submit enqueues k batches x repeat in one foreign call; mm_draw_ticket shades an older
    batch on the consumer stream while newer ones are in the ring."""
    import mimas_b200
    from mimas_b200.render import SHADED_DTYPE, Camera
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        all_sets = _ring_inputs(m)
        sets = all_sets[:4]
        want = [oracle.mesh_batch(c, o, ocache, threads=8) for c, o, _, _ in sets]
        m.ring_configure(4)
        m.reserve(600, 60 * 73728, 500 * 24 * 10)
        first = m.submit_many([s[2] for s in sets], [s[3] for s in sets], [len(s[0]) for s in sets], repeat=3)
        last4 = [first + 8 + k for k in range(4)]      # the last round is what the ring holds
        cam = Camera(pos=(3.0, -2.0, 12.0), pitch=-20.0, yaw=40.0, viewing_range=200.0)
        c0, o0 = sets[0][0], sets[0][1]
        d = m.draw_ticket(last4[0], cam)
        vm, pm = cam.get_matrix(), cam.get_perspective()
        drawn = oracle.cull_chunks(o0, vm, pm, cam.pos, cam.viewing_range)
        assert d.n_drawn == len(drawn) and 0 < d.n_drawn
        mesh = m.copy_out_ticket(last4[0])
        vi = np.concatenate([mesh.chunk(int(c)).intransparent for c in drawn])
        vt = np.concatenate([mesh.chunk(int(c)).transparent for c in drawn])
        wantv = oracle.vertex_stage(np.concatenate([vi, vt]), vm, pm)
        d = m.draw_ticket(last4[0], cam)
        got = m.download(d.d_shaded, 48 * (len(vi) + len(vt))).view(SHADED_DTYPE)
        assert got.tobytes() == wantv.tobytes()
        for k in range(4):
            assert_batch_equal(m.copy_out_ticket(last4[k]), want[k])
        for _, _, dc, do in all_sets:
            m.dev_free(dc)
            m.dev_free(do)


def test_ring_poll_empty_batches_and_halo_tickets(ocache):
    """mm_batch_poll, zero-chunk batches between real ones, and a halo batch through the ring."""
    import time
    import mimas_b200
    from tests.util import halos_for_box
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        dims = (6, 5, 4)
        n = dims[0] * dims[1] * dims[2]
        chunks, offs = mimas_b200.gen_terrain_host(78, (-3, -2, -2), dims)
        halo = halos_for_box(chunks, dims, outside=0)
        dc, do, dh = m.dev_alloc(n * 4096), m.dev_alloc(n * 24), m.dev_alloc(n * 1536)
        m.upload(dc, chunks); m.upload(do, offs); m.upload(dh, halo)
        m.ring_configure(4)
        t_plain = m.submit(dc, do, n)
        t_empty = m.submit(dc, do, 0)
        t_halo = m.submit(dc, do, n, d_halo=dh)
        t_last = m.submit(dc, do, n)
        deadline = time.time() + 30
        while not m.batch_poll(t_last):                      # the newest batch: polled through the stream
            assert time.time() < deadline
        assert m.batch_poll(t_plain) and m.batch_poll(t_empty) and m.batch_poll(t_halo)
        want = oracle.mesh_batch(chunks, offs, ocache, threads=8)
        assert_batch_equal(m.copy_out_ticket(t_plain), want)
        assert_batch_equal(m.copy_out_ticket(t_last), want)
        assert_batch_equal(m.copy_out_ticket(t_halo), oracle.mesh_batch(chunks, offs, ocache, threads=8, halo=halo))
        e = m.batch_wait(t_empty)
        assert e.n_chunks == 0 and e.n_intransparent == 0 and e.n_transparent == 0
        for p in (dc, do, dh):
            m.dev_free(p)


def test_single_chunk_host_calls_between_ring_batches_leave_them_alone(ocache):
    """The reference's call pattern (one chunk, host in / host out: the mapped path) while ring batches are in flight:
    neither disturbs the other."""
    import mimas_b200
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        chunks, offs = mimas_b200.gen_terrain_host(78, (-4, -4, -2), (8, 8, 4))
        n = len(chunks)
        want = oracle.mesh_batch(chunks, offs, ocache, threads=8)
        dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
        m.upload(dc, chunks); m.upload(do, offs)
        m.ring_configure(2)
        t1 = m.submit(dc, do, n)
        t2 = m.submit(dc, do, n)
        for name in ("layered_terrain", "one_sapling", "checkerboard_odd", "all_air"):
            c = G.cases[name]
            got = m.mesh_for_chunk(c["offs"], c["chunk"])
            assert first_diff(got.intransparent.tobytes(), c["intr"]) is None
            assert first_diff(got.transparent.tobytes(), c["trans"]) is None
        small = m.mesh_batch(offs[:40], chunks[:40])                    # 40 chunks: still the mapped path
        assert_batch_equal(small, oracle.mesh_batch(chunks[:40], offs[:40], ocache, threads=4))
        assert_batch_equal(m.copy_out_ticket(t1), want)
        assert_batch_equal(m.copy_out_ticket(t2), want)
        m.dev_free(dc); m.dev_free(do)


def test_error_behaviour_of_the_c_abi():
    """The reference's mesher is infallible (no Result, lib.rs:231-232); the C ABI answers every misuse with a status
    code and a message, never with a crash or a silently wrong mesh: no table, bad arguments, no batch, bad ring size,
    a table row of an unknown kind -- and the context keeps working afterwards."""
    import ctypes as C
    import mimas_b200
    from mimas_b200._lib import MMError, lib
    l = lib()
    with mimas_b200.Mesher(0) as m:                       # no table yet
        chunk = np.zeros(4096, np.uint8)
        with pytest.raises(MMError) as e:
            m.mesh_for_chunk((0, 0, 0), chunk)
        assert e.value.code == 5                          # MM_ERR_NO_TABLE
        assert b"mm_set_block_table" in l.mm_last_error_string(m._ctx)
        with pytest.raises(MMError) as e:
            m.batch_finish()
        assert e.value.code == 6                          # MM_ERR_NO_BATCH
        with pytest.raises(MMError) as e:
            m.ring_configure(9)
        assert e.value.code == 1                          # MM_ERR_INVALID
        bad = np.zeros(256, dtype=mimas_b200.meshgen.BLOCK_ENTRY_DTYPE)
        bad["kind"][7] = 3
        with pytest.raises(MMError) as e:
            m.set_table(bad)
        assert e.value.code == 1
        assert l.mm_mesh_batch(m._ctx, None, None, None, 5, None) == 1
        assert l.mm_batch_wait(m._ctx, 0, None) == 1
        assert l.mm_ctx_create(99, C.byref(C.c_void_p())) == 1   # no such device
        m.set_cache(product_cache(G))                     # and now it works
        c = G.cases["layered_terrain"]
        got = m.mesh_for_chunk(c["offs"], c["chunk"])
        assert first_diff(got.intransparent.tobytes(), c["intr"]) is None
        with pytest.raises(MMError) as e:
            m.batch_wait(12345)                           # a ticket that was never issued
        assert e.value.code == 6
