"""Headless consumer of the vertex stream: chunk culling (client.rs:679-696 with the
frustum_query sphere test) and the vertex shader (vertex-shader.glsl:18-27).
CPU: the restatement (oracle/render_oracle.c) against hand-derived answers.  GPU: the CUDA kernels are
bit-exact with the restatement (both round every product and sum separately; tolerance zero).  Against real
GL the tolerance would be GLSL's (no ordering / FMA guarantees) -- unavailable on this box; parity to the
reference is unpinned for this row (no tests or vectors in the reference, frustum_query absent)."""
import numpy as np
import pytest

import oracle
from mimas_b200.render import SHADED_DTYPE, Camera
from tests.util import load_golden, product_cache

G = load_golden()
IDENT = np.eye(4, dtype=np.float32)


def test_oracle_vertex_stage_hand_values():
    v = np.zeros(3, dtype=oracle.VERTEX_DTYPE)
    v["position"] = [(1, 2, 3), (0, 0, 0), (-4, 5, 6)]
    v["normal"] = [(1, 1, 0), (0, -1, -1), (0, 0, 1)]     # the mesher's sign-vector "normals" 
    v["tex_pos"] = [(0, 1), (2, 3), (4, 5)]
    v["tex_ind"] = [6, 7, 65535]
    vm = IDENT.copy()
    vm[3, :3] = (10, 20, 30)                               # column 3 = translation (column-major: m[col][row])
    pm = np.diag([2, 3, 4, 1]).astype(np.float32)
    s = oracle.vertex_stage(v, vm, pm)
    assert s["vposition"].tolist() == [[11, 22, 33, 1], [10, 20, 30, 1], [6, 25, 36, 1]]
    assert s["clip"].tolist() == [[22, 66, 132, 1], [20, 60, 120, 1], [12, 75, 144, 1]]
    l = np.float32(1) / np.sqrt(np.float32(2))
    # (1,1,0)/sqrt2: light a = -l*ny < 0.2 -> 0.2; light b = l*ny = 0.5
    ny = np.float32(1) / np.sqrt(np.float32(2))
    assert s["vlamb"][0] == np.float32(0.2) + l * ny
    assert s["vlamb"][1] == (l * ny + l * ny) + np.float32(0.2)        # (0,-1,-1): facing light a fully
    assert s["vlamb"][2] == np.float32(0.2) + l * np.float32(1)
    assert s["vtex_pos"].tolist() == [[0, 1], [2, 3], [4, 5]] and s["vtex_ind"].tolist() == [6, 7, 65535]


def test_oracle_frustum_and_culling_hand_cases():
    cam = Camera(pos=(8, 8, 8), pitch=0, yaw=0, viewing_range=100.0)     # looking along +x
    vm, pm = cam.get_matrix(), cam.get_perspective()
    pl = oracle.frustum_planes(vm, pm)
    assert np.allclose(np.linalg.norm(pl[:, :3], axis=1), 1, atol=1e-6)    # normalised planes
    offs = np.array([[64, 0, 0],      # ahead: drawn
                     [-96, 0, 0],     # behind the camera: frustum-culled (sphere radius 27.7 does not reach)
                     [0, 0, 0],       # the camera's own chunk: drawn (sphere around it)
                     [160, 0, 0],     # ahead but origin 152 blocks away > viewing range 100
                     [64, 200, 0],    # far to the side: outside the 90 degree frustum? no: beyond viewing range
                     [48, 64, 0]],    # 45 degrees off-axis: on the frustum edge, sphere intersects
                    dtype=np.int64)
    drawn = oracle.cull_chunks(offs, vm, pm, cam.pos, cam.viewing_range)
    assert drawn.tolist() == [0, 2, 5]
    assert oracle.cull_chunks(offs, vm, pm, cam.pos, 1000.0).tolist() == [0, 2, 3, 5]
    # looking the other way flips ahead / behind
    back = Camera(pos=(8, 8, 8), pitch=0, yaw=180, viewing_range=1000.0)
    assert oracle.cull_chunks(offs, back.get_matrix(), back.get_perspective(), back.pos, 1000.0).tolist() == [1, 2]


@pytest.fixture(scope="module")
def mesher():
    import mimas_b200
    m = mimas_b200.Mesher(0, product_cache(G))
    yield m
    m.close()


@pytest.mark.gpu
def test_cuda_culling_matches_restatement(mesher):
    rng = np.random.RandomState(9)
    offs = (rng.randint(-40, 40, size=(5000, 3)) * 16).astype(np.int64)
    offs[:4] = [[2 ** 40, 0, 0], [0, -2 ** 40, 16], [0, 0, 0], [48, 32, 16]]
    for cam in (Camera(), Camera(pos=(-100.5, 33.25, 7.0), pitch=-35.0, yaw=123.0, viewing_range=300.0),
                Camera(pos=(0, 0, 0), pitch=89.0, yaw=-179.0, aspect_ratio=2.5, viewing_range=1e9)):
        want = oracle.cull_chunks(offs, cam.get_matrix(), cam.get_perspective(), cam.pos, cam.viewing_range)
        got = mesher.cull_chunks(offs, cam)
        assert 0 < len(want) < len(offs)
        assert got.tolist() == want.tolist()


@pytest.mark.gpu
def test_cuda_culling_large_batch(mesher):
    """60000 chunk origins: 1.4 MB of offsets and 240 KB of draw list, i.e. numpy buffers that live in mmap'ed
    memory above 4 GiB -- every pointer must cross the ctypes boundary at full width (_lib.py argtypes)."""
    rng = np.random.RandomState(10)
    offs = (rng.randint(-60, 60, size=(60000, 3)) * 16).astype(np.int64)
    cam = Camera(pos=(5.5, -3.25, 9.0), pitch=-10.0, yaw=77.0, viewing_range=600.0)
    want = oracle.cull_chunks(offs, cam.get_matrix(), cam.get_perspective(), cam.pos, cam.viewing_range)
    got = mesher.cull_chunks(offs, cam)
    assert 1000 < len(want) < len(offs)
    assert got.tolist() == want.tolist()


@pytest.mark.gpu
def test_cuda_draw_batch_matches_restatement(mesher):
    """One frame: cull the chunks of a meshed terrain region, shade the visible ones; every output byte equals
    culling + vertex stage of the restatement on the oracle-checked meshes."""
    import mimas_b200
    dims, c0 = (12, 12, 4), (-6, -6, -2)
    n = dims[0] * dims[1] * dims[2]
    dc, do = mesher.dev_alloc(n * 4096), mesher.dev_alloc(n * 24)
    try:
        mesher.gen_terrain_device(78, c0, dims, dc, do)
        out = mesher.mesh_batch_device(dc, do, n)
        mesh = mesher.copy_out(out)
        offs = mesher.download(do, n * 24).view(np.int64).reshape(n, 3)
        for cam in (Camera(pos=(3.0, -2.0, 12.0), pitch=-20.0, yaw=40.0, viewing_range=90.0),
                    Camera(pos=(60.0, 40.0, 20.0), pitch=10.0, yaw=-150.0, viewing_range=128.0)):
            d = mesher.draw_batch(cam)
            vm, pm = cam.get_matrix(), cam.get_perspective()
            drawn = oracle.cull_chunks(offs, vm, pm, cam.pos, cam.viewing_range)
            assert d.n_chunks == n and d.n_drawn == len(drawn) and 0 < d.n_drawn < n
            assert mesher.download(d.d_draw_list, 4 * d.n_drawn).view(np.uint32).tolist() == drawn.tolist()
            vi = np.concatenate([mesh.chunk(int(c)).intransparent for c in drawn])
            vt = np.concatenate([mesh.chunk(int(c)).transparent for c in drawn])
            assert d.n_shaded_i == len(vi) and d.n_shaded_t == len(vt)
            want = oracle.vertex_stage(np.concatenate([vi, vt]), vm, pm)
            got = mesher.download(d.d_shaded, 48 * (len(vi) + len(vt))).view(SHADED_DTYPE)
            assert got.tobytes() == want.tobytes()
        # plain vertex stage over the whole stream
        d_out = mesher.dev_alloc(48 * out.n_intransparent)
        cam = Camera(pos=(1, 2, 3), pitch=5, yaw=6)
        mesher.vertex_stage(out.d_intransparent, out.n_intransparent, cam, d_out)
        mesher.sync()
        got = mesher.download(d_out, 48 * out.n_intransparent).view(SHADED_DTYPE)
        assert got.tobytes() == oracle.vertex_stage(mesh.intransparent, cam.get_matrix(), cam.get_perspective()).tobytes()
        mesher.dev_free(d_out)
    finally:
        mesher.dev_free(dc)
        mesher.dev_free(do)
