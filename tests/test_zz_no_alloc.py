"""mm_mesh_batch_noalloc (no counterpart in the reference): the mesher
with its hand-over scratch in a static device array, every output buffer the caller's, nothing allocated by the call.
Checked through the C ABI against the existing entry point mm_mesh_batch on the same inputs, tolerance zero (byte
compare, as every mesher test here), on two calls in a row that share the scratch array.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from tests.util import first_diff, load_golden, product_cache

pytestmark = pytest.mark.gpu
MAX_CHUNKS = 256  # MM_NOALLOC_MAX_CHUNKS


@pytest.fixture(scope="module")
def mesher():
    import mimas_b200
    m = mimas_b200.Mesher(0, product_cache(load_golden()))
    yield m
    m.close()


def _resident(m, c0, dims):
    import mimas_b200
    chunks, offs = mimas_b200.gen_terrain_host(78, c0, dims)
    n = len(chunks)
    dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
    m.upload(dc, chunks)
    m.upload(do, np.ascontiguousarray(offs, dtype=np.int64))
    return dc, do, n


def _reference(m, dc, do, n):
    """the existing entry point: mm_mesh_batch + mm_copy_out"""
    return m.copy_out(m.mesh_batch_device(dc, do, n))


class _Outputs:
    """caller-owned device buffers for one call, allocated BEFORE the call that must not allocate"""

    def __init__(self, m, n, verts_i, verts_t):
        self.m, self.n, self.vi, self.vt = m, n, verts_i, verts_t
        self.d_i, self.d_t = m.dev_alloc(max(verts_i, 1) * 36), m.dev_alloc(max(verts_t, 1) * 36)
        self.d_ri, self.d_rt, self.d_tot = m.dev_alloc(n * 16), m.dev_alloc(n * 16), m.dev_alloc(16)
        for p, b in ((self.d_i, max(verts_i, 1) * 36), (self.d_t, max(verts_t, 1) * 36)):
            m.upload(p, np.full(b, 0xA5, dtype=np.uint8))  # stale bytes the call has to overwrite

    def call(self, dc, do, cap_i=None, cap_t=None):
        self.m.mesh_batch_noalloc(dc, do, self.n, self.d_i, self.vi if cap_i is None else cap_i, self.d_t,
                                  self.vt if cap_t is None else cap_t, self.d_ri, self.d_rt, self.d_tot)

    def fetch(self):
        m = self.m
        tot = m.download(self.d_tot, 16).view("<u8")
        return (m.download(self.d_i, self.vi * 36), m.download(self.d_t, self.vt * 36), m.download(self.d_ri, self.n * 16),
                m.download(self.d_rt, self.n * 16), int(tot[0]), int(tot[1]))

    def free(self):
        for p in (self.d_i, self.d_t, self.d_ri, self.d_rt, self.d_tot):
            self.m.dev_free(p)


def _same(got, want):
    gi, gt, gri, grt, qi, qt = got
    assert qi * 6 == len(want.intransparent) and qt * 6 == len(want.transparent)
    assert first_diff(gi.tobytes(), want.intransparent.tobytes()) is None
    assert first_diff(gt.tobytes(), want.transparent.tobytes()) is None
    assert gri.tobytes() == want.ranges_i.tobytes() and grt.tobytes() == want.ranges_t.tobytes()   # range records


def test_two_calls_in_a_row_match_mm_mesh_batch(mesher):
    m = mesher
    a = _resident(m, (-16, 0, -2), (8, 8, 4))     # 256 chunks = the maximum: surface, plants, caves, air
    b = _resident(m, (24, 24, -1), (6, 6, 3))     # 108 chunks elsewhere: the scratch still holds call one's content
    want_a, want_b = _reference(m, *a), _reference(m, *b)
    assert a[2] == MAX_CHUNKS and len(want_a.intransparent) > 0 and len(want_a.transparent) > 0 and len(want_b.transparent) > 0
    oa = _Outputs(m, a[2], len(want_a.intransparent), len(want_a.transparent))
    ob = _Outputs(m, b[2], len(want_b.intransparent), len(want_b.transparent))
    l0 = m.launch_count()
    oa.call(a[0], a[1])   # two calls back to back on the context's stream, no synchronisation between them
    ob.call(b[0], b[1])
    m.sync()
    assert m.launch_count() - l0 == 4   # 2 x (analysis + expansion): the CUDA kernels ran, nothing else
    _same(oa.fetch(), want_a)
    _same(ob.fetch(), want_b)
    # and once more over the same buffers, the other way round, after the existing entry point used the stream
    _reference(m, *b)
    ob.call(b[0], b[1])
    oa.call(a[0], a[1])
    m.sync()
    _same(ob.fetch(), want_b)
    _same(oa.fetch(), want_a)
    for o in (oa, ob):
        o.free()
    for r in (a, b):
        m.dev_free(r[0]); m.dev_free(r[1])


def test_limits_and_small_buffers(mesher):
    from mimas_b200._lib import MMError
    m = mesher
    r = _resident(m, (8, 8, -1), (4, 4, 2))
    want = _reference(m, *r)
    assert len(want.intransparent) > 6 and len(want.transparent) > 0
    o = _Outputs(m, r[2], len(want.intransparent), len(want.transparent))
    with pytest.raises(MMError, match="MM_NOALLOC_MAX_CHUNKS"):   # beyond the static scratch: a clear error, nothing launched
        m.mesh_batch_noalloc(r[0], r[1], MAX_CHUNKS + 1, o.d_i, o.vi, o.d_t, o.vt, o.d_ri, o.d_rt, o.d_tot)
    with pytest.raises(MMError):
        m.mesh_batch_noalloc(r[0] + 4, r[1], r[2], o.d_i, o.vi, o.d_t, o.vt, o.d_ri, o.d_rt, o.d_tot)  # misaligned chunks
    # a stream that does not fit its buffer is left alone; totals and range records still say what it needs
    o.call(r[0], r[1], cap_i=len(want.intransparent) - 6)
    m.sync()
    gi, gt, gri, grt, qi, qt = o.fetch()
    assert qi * 6 == len(want.intransparent) and qt * 6 == len(want.transparent)
    assert set(np.unique(gi)) == {0xA5}
    assert first_diff(gt.tobytes(), want.transparent.tobytes()) is None
    assert gri.tobytes() == want.ranges_i.tobytes()
    # n == 0: totals are zeroed, nothing else happens
    m.mesh_batch_noalloc(0, 0, 0, o.d_i, o.vi, o.d_t, o.vt, o.d_ri, o.d_rt, o.d_tot)
    m.sync()
    assert not m.download(o.d_tot, 16).any()
    o.call(r[0], r[1])
    m.sync()
    _same(o.fetch(), want)
    o.free()
    m.dev_free(r[0]); m.dev_free(r[1])
