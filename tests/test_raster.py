"""The raster oracle (oracle/raster_oracle.c: clipping, rasterisation, depth test, fragment-shader.glsl:14-24)
against hand-derived cases and an independent numpy ray cast (CPU), and the CUDA rasteriser
(csrc/mm_raster.cu) against that oracle, bit for bit (-m gpu).  Parity to the reference's GL output is
unpinned: see the oracle's header."""
import numpy as np

import oracle
from mimas_b200.render import Camera


def unpack(img):
    return np.stack([(img >> s) & 0xFF for s in (0, 8, 16, 24)], axis=-1).astype(np.uint8)


def make_shaded(tris, lamb=1.0, layer=0, vpos=None, uv=None):
    """tris: [n, 3, 4] clip coordinates."""
    tris = np.asarray(tris, dtype=np.float32)
    s = np.zeros(tris.shape[0] * 3, dtype=oracle.SHADED_DTYPE)
    s["clip"] = tris.reshape(-1, 4)
    s["vlamb"] = lamb
    s["vtex_ind"] = layer
    s["vposition"] = (0, 0, 0, 1) if vpos is None else np.asarray(vpos, np.float32).reshape(-1, 4)
    if uv is not None:
        s["vtex_pos"] = np.asarray(uv, np.float32).reshape(-1, 2)
    return s


def flat_tex(*colors, size=16):
    return np.stack([np.broadcast_to(np.asarray(c, np.float32), (size, size, 4)) for c in colors]).copy()


CLEAR = (13, 3, 153, 0)  # (0.05, 0.01, 0.6, 0.0) * 255 rounded


def cw_quad(x0, y0, x1, y1, z=0.0, w=1.0):
    """Two clockwise (front-facing, client.rs:667) triangles covering [x0,x1] x [y0,y1] in NDC."""
    a, b, c, d = (x0, y0, z, w), (x1, y0, z, w), (x1, y1, z, w), (x0, y1, z, w)
    return [[a, c, b], [a, d, c]]


def test_clear_colour_and_full_screen_quad():
    tex = flat_tex((1.0, 0.5, 0.25, 1.0))
    img = unpack(oracle.raster_frame(make_shaded(np.zeros((0, 3, 4))), 8, 4, 40.0, 60.0, tex))
    assert (img == CLEAR).all()
    img = unpack(oracle.raster_frame(make_shaded(cw_quad(-1, -1, 1, 1), lamb=0.8), 8, 4, 40.0, 60.0, tex))
    # vposition = (0,0,0,1): length 1, fog = clamp((1 - 60) / 40) = 0 -> colour = 0.8 * texel, alpha = 1
    assert (img == (204, 102, 51, 255)).all()


def test_backface_culled_and_shared_edges_covered_once():
    tex = flat_tex((1, 1, 1, 1))
    ccw = [[t[0], t[2], t[1]] for t in cw_quad(-1, -1, 1, 1)]
    assert (unpack(oracle.raster_frame(make_shaded(ccw), 16, 16, 40, 60, tex)) == CLEAR).all()
    # a fan of triangles around an off-centre point: every pixel is covered, none is left to the clear colour;
    # with one layer per triangle and a depth that favours nobody, coverage is a partition
    rng = np.random.RandomState(0)
    cx, cy = 0.137, -0.291
    ring = [(-1, -1), (0.3, -1), (1, -1), (1, 0.2), (1, 1), (-0.4, 1), (-1, 1), (-1, 0.1)]
    tris = []
    for i in range(len(ring)):
        p, q = ring[i], ring[(i + 1) % len(ring)]
        tris.append([(cx, cy, 0, 1), (q[0], q[1], 0, 1), (p[0], p[1], 0, 1)])  # clockwise
    s = make_shaded(tris)
    s["vtex_ind"] = np.repeat(np.arange(len(tris)), 3)
    tex = flat_tex(*[(rng.rand(), rng.rand(), rng.rand(), 1.0) for _ in tris])
    img, keys = oracle.raster_frame(s, 64, 48, 40, 60, tex, want_keys=True)
    assert (unpack(img)[..., 3] == 255).all()
    # each triangle alone covers a set of pixels; the sets are disjoint and their union is the screen
    total = np.zeros((48, 64), int)
    for i in range(len(tris)):
        one = unpack(oracle.raster_frame(s[3 * i:3 * i + 3], 64, 48, 40, 60, tex))
        total += (one[..., 3] == 255)
    assert (total == 1).all()


def test_depth_test_alpha_test_and_fog():
    tex = flat_tex((1, 0, 0, 1), (0, 1, 0, 1), (0, 0, 1, 0.4))
    near = make_shaded(cw_quad(-1, -1, 0, 1, z=-0.5), layer=0)
    far = make_shaded(cw_quad(-1, -1, 1, 1, z=0.5), layer=1)
    holes = make_shaded(cw_quad(-1, -1, 1, 1, z=-0.9), layer=2)  # alpha 0.4 < 0.5: discarded everywhere
    for order in ([near, far, holes], [holes, far, near]):
        img = unpack(oracle.raster_frame(np.concatenate(order), 8, 2, 40, 60, tex))
        assert (img[:, :4] == (255, 0, 0, 255)).all() and (img[:, 4:] == (0, 255, 0, 255)).all()
    # fog: |vposition| = sqrt(70^2 + 1) -> factor (70.007 - 60) / 40 = 0.2502; mix(green, grey, f)
    fogged = make_shaded(cw_quad(-1, -1, 1, 1), layer=1, vpos=[(0, 0, -70, 1)] * 6)
    img = unpack(oracle.raster_frame(fogged, 4, 4, 40, 60, tex))
    f = (np.sqrt(4901.0) - 60) / 40
    want = np.round(np.array([0.5 * f, 1 - f + 0.5 * f, 0.5 * f, 1.0]) * 255)
    assert (np.abs(img.astype(int) - want) <= 1).all()


def test_perspective_ground_plane_against_ray_cast():
    """A big textured ground quad seen through the reference's camera: per pixel, the texel the oracle sampled
    equals the one a ray cast through the pixel centre hits (perspective-correct interpolation, near-plane
    clipping: the quad passes under and behind the camera)."""
    cam = Camera(pos=(3.3, -2.1, 7.7), pitch=35.0, yaw=25.0)  # pitch > 0 looks down (client.rs:1382-1386)
    V = np.asarray(cam.get_matrix(), np.float64).reshape(4, 4).T   # column-major -> math convention
    P = np.asarray(cam.get_perspective(), np.float64).reshape(4, 4).T
    L = 40.0
    corners = np.array([(-L, -L, 0, 1), (L, -L, 0, 1), (L, L, 0, 1), (-L, L, 0, 1)], np.float64)
    uvs = (corners[:, :2] + L)  # one texel repeat per block, like a merged quad (u, v run to 2L)
    ts = 4
    rng = np.random.RandomState(3)
    tex = rng.rand(1, ts, ts, 4).astype(np.float32)
    tex[..., 3] = 1.0
    W, H = 96, 64
    best = None
    for order in ([0, 2, 1, 0, 3, 2], [0, 1, 2, 0, 2, 3]):  # one of the windings faces the camera
        idx = np.array(order)
        vp = (V @ corners[idx].T).T
        s = make_shaded((P @ vp.T).T.reshape(2, 3, 4), vpos=vp, uv=uvs[idx])
        img, keys = oracle.raster_frame(s, W, H, 1e9, 1e9, tex, want_keys=True)
        if (keys != np.uint64(0xFFFFFFFFFFFFFFFF)).any():
            best = img
    assert best is not None
    got = unpack(best)
    # ray cast: pixel centre -> NDC -> world ray -> z = 0 plane
    inv = np.linalg.inv(P @ V)
    checked = bad = 0
    for y in range(H):
        for x in range(W):
            nx, ny = (x + 0.5) / W * 2 - 1, (y + 0.5) / H * 2 - 1
            a, b = inv @ np.array([nx, ny, -1, 1.0]), inv @ np.array([nx, ny, 1, 1.0])
            a, b = a[:3] / a[3], b[:3] / b[3]
            d = b - a
            if abs(d[2]) < 1e-12:
                continue
            t = -a[2] / d[2]
            hit = a + t * d
            if t <= 0 or t >= 1 or abs(hit[0]) >= L or abs(hit[1]) >= L:
                continue  # sky, beyond the far plane, or off the quad
            u, v = hit[0] + L, hit[1] + L
            fu, fv = (u % 1.0) * ts, (v % 1.0) * ts
            if min(fu % 1.0, 1 - fu % 1.0, fv % 1.0, 1 - fv % 1.0) < 0.02:
                continue  # too close to a texel border for float32 vs float64 to agree
            texel = tex[0, int(fv), int(fu)]
            want = np.round(np.clip(texel, 0, 1) * 255)
            bad += not (np.abs(got[y, x].astype(int) - want) <= 1).all()
            checked += 1
    # float32 interpolation vs the float64 ray: near the horizon a texel is a small fraction of a pixel
    assert checked > 3000 and bad <= 3, (checked, bad)


# ---- GPU: the CUDA rasteriser (csrc/mm_raster.cu) is bit-exact with the oracle -------------------------------
import pytest  # noqa: E402


@pytest.fixture(scope="module")
def mesher():
    import mimas_b200
    from tests.util import load_golden, product_cache
    m = mimas_b200.Mesher(0, product_cache(load_golden()))
    yield m
    m.close()


def gpu_raster(m, shaded, w, h, fog_near, fog_far, tex):
    m.set_textures(tex)
    shaded = np.ascontiguousarray(shaded, dtype=oracle.SHADED_DTYPE)
    d = m.dev_alloc(max(48 * len(shaded), 48))
    try:
        if len(shaded):
            m.upload(d, shaded.view(np.uint8))
        return m.raster_frame(d, len(shaded), w, h, fog_near, fog_far)
    finally:
        m.dev_free(d)


@pytest.mark.gpu
def test_gpu_hand_cases_equal_oracle(mesher):
    tex = flat_tex((1, 0, 0, 1), (0, 1, 0, 1), (0, 0, 1, 0.4))
    scenes = [make_shaded(np.zeros((0, 3, 4))),
              make_shaded(cw_quad(-1, -1, 1, 1), lamb=0.8),
              np.concatenate([make_shaded(cw_quad(-1, -1, 0, 1, z=-0.5), layer=0), make_shaded(cw_quad(-1, -1, 1, 1, z=0.5), layer=1),
                              make_shaded(cw_quad(-1, -1, 1, 1, z=-0.9), layer=2)]),
              make_shaded(cw_quad(-1, -1, 1, 1), layer=1, vpos=[(0, 0, -70, 1)] * 6)]
    for s in scenes:
        for (w, h) in ((8, 4), (64, 48), (333, 17)):
            want = oracle.raster_frame(s, w, h, 40.0, 60.0, tex)
            got = gpu_raster(mesher, s, w, h, 40.0, 60.0, tex)
            assert (got == want).all()


@pytest.mark.gpu
def test_gpu_random_triangles_equal_oracle(mesher):
    """Random clip-space triangles of every kind: behind the camera, crossing the near / far / guard-band planes,
    sub-pixel, screen-filling, degenerate, with random attributes and alpha holes."""
    rng = np.random.RandomState(5)
    n = 4000
    tris = rng.randn(n, 3, 4).astype(np.float32)
    tris[:, :, 3] = np.abs(tris[:, :, 3]) + 0.05
    tris[: n // 4, :, 3] = rng.randn(n // 4, 3) * 0.7                 # w of either sign: near-plane clipping
    tris[n // 4: n // 2] *= rng.rand(n // 4, 1, 1).astype(np.float32) * 30  # huge: guard band, queue of big triangles
    small = slice(n // 2, 3 * n // 4)
    tris[small, 1:, :2] = tris[small, :1, :2] + rng.randn(n // 4, 2, 2).astype(np.float32) * 0.01  # tiny
    tris[-50:, 2] = tris[-50:, 1]                                                                  # degenerate
    s = make_shaded(tris)
    s["vlamb"] = rng.rand(3 * n) * 1.4
    s["vtex_pos"] = rng.randn(3 * n, 2) * 3
    s["vtex_ind"] = rng.randint(0, 9, 3 * n)                                                        # beyond the last layer too
    s["vposition"] = rng.randn(3 * n, 4) * 30
    tex = rng.rand(6, 16, 16, 4).astype(np.float32)
    tex[..., 3] = (rng.rand(6, 16, 16) < 0.8)
    for (w, h, fn, ff) in ((160, 90, 40.0, 60.0), (97, 131, 5.0, 10.0)):
        want = oracle.raster_frame(s, w, h, fn, ff, tex)
        got = gpu_raster(mesher, s, w, h, fn, ff, tex)
        bad = np.argwhere(got != want)
        assert len(bad) == 0, (len(bad), bad[:5], got[tuple(bad[0])], want[tuple(bad[0])])


@pytest.mark.gpu
def test_gpu_frame_of_a_meshed_region_equals_oracle(mesher):
    """The whole headless frame: mesh a terrain region, cull + vertex-shade on the device, rasterise; against the
    oracle chain (CPU mesher -> culling -> vertex stage -> raster oracle) fed with the same camera."""
    import mimas_b200
    from mimas_b200.render import synthetic_textures
    from tests.util import load_golden, oracle_cache
    chunks, offs = mimas_b200.gen_terrain_host(78, (-3, -3, -1), (6, 6, 3))
    tex = synthetic_textures(80, crossed=range(30, 44))
    mesher.set_textures(tex)
    n = len(chunks)
    d_c, d_o = mesher.dev_alloc(n * 4096), mesher.dev_alloc(n * 24)
    try:
        mesher.upload(d_c, chunks)
        mesher.upload(d_o, offs)
        mesher.mesh_batch_device(d_c, d_o, n)
        oc = oracle_cache(load_golden())
        want_mesh = oracle.mesh_batch(chunks, offs, oc, threads=8)
        for cam in (Camera(pos=(5.0, 3.0, 30.0), pitch=40.0, yaw=30.0, viewing_range=200.0),
                    Camera(pos=(-20.0, 10.0, 22.0), pitch=10.0, yaw=-60.0, viewing_range=60.0)):
            got = mesher.render_frame(cam, 200, 150, 40.0, 60.0)
            vis = oracle.cull_chunks(offs, cam.get_matrix(), cam.get_perspective(), cam.pos, cam.viewing_range)
            parts = [want_mesh["intransparent"][want_mesh["first_i"][c]:want_mesh["first_i"][c] + want_mesh["count_i"][c]] for c in vis]
            parts += [want_mesh["transparent"][want_mesh["first_t"][c]:want_mesh["first_t"][c] + want_mesh["count_t"][c]] for c in vis]
            verts = np.concatenate(parts) if parts else np.zeros(0, oracle.VERTEX_DTYPE)
            shaded = oracle.vertex_stage(verts, cam.get_matrix(), cam.get_perspective())
            want = oracle.raster_frame(shaded, 200, 150, 40.0, 60.0, tex)
            assert (got == want).all(), int((got != want).sum())
            assert (unpack(got)[..., 3] > 0).mean() > 0.2   # the frame is not empty
    finally:
        mesher.dev_free(d_c)
        mesher.dev_free(d_o)


def test_png_writer_and_synthetic_textures(tmp_path):
    import struct
    import zlib
    from mimas_b200.render import synthetic_textures, write_png
    tex = synthetic_textures(10, crossed=(3,))
    assert tex.shape == (10, 16, 16, 4) and tex.dtype == np.float32 and tex.min() >= 0 and tex.max() <= 1
    assert set(np.unique(tex[3, :, :, 3])) <= {0.0, 1.0} and (tex[2, :, :, 3] == 1).all()
    img = (np.arange(6, dtype=np.uint32).reshape(2, 3) * 0x010203 + 0xFF000000)
    path = tmp_path / "t.png"
    write_png(str(path), img)
    data = path.read_bytes()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    w, h, depth, ctype = struct.unpack(">IIBB", data[16:26])
    assert (w, h, depth, ctype) == (3, 2, 8, 2)
    at = data.index(b"IDAT")
    n = struct.unpack(">I", data[at - 4:at])[0]
    raw = zlib.decompress(data[at + 4:at + 4 + n])
    rows = [raw[i * 10 + 1:(i + 1) * 10] for i in range(2)]
    # top row of the file = row 1 of the image (GL's origin is at the bottom)
    assert rows[0] == bytes([3 * 3, 3 * 2, 3 * 1, 4 * 3, 4 * 2, 4 * 1, 5 * 3, 5 * 2, 5 * 1])
    assert rows[1][:3] == bytes([0, 0, 0])
