"""TEST INFRASTRUCTURE -- ctypes front-end of the C parity oracle (oracle/meshgen_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package.  The product package (mimas_b200) never does.

PARITY UNPINNED: see the header of meshgen_oracle.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmeshgen_oracle.so")

# #[repr(C)] Vertex, SURVEY.md §8b / mimas-meshgen/lib.rs:18-24
VERTEX_DTYPE = np.dtype([
    ("tex_ind", "<u2"), ("pad0", "<u2"), ("tex_pos", "<f4", (2,)),
    ("position", "<f4", (3,)), ("normal", "<f4", (3,)),
])
assert VERTEX_DTYPE.itemsize == 36


class Bti(C.Structure):
    _fields_ = [("id_sides", C.c_uint16), ("id_top", C.c_uint16), ("id_bottom", C.c_uint16)]


class Cache(C.Structure):
    _fields_ = [
        ("fallback_id", Bti), ("len", C.c_uint32), ("style", C.c_uint8 * 256),
        ("blocky", Bti * 256), ("crossed", C.c_uint16 * 256),
    ]


def build(force: bool = False) -> str:
    """Compile the oracle with its Makefile (gcc).  Building the checker is not using it."""
    src = os.path.join(HERE, "meshgen_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-s", "-B"], check=True)
    return LIB_PATH


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    build()
    lib = C.CDLL(LIB_PATH)
    vpp = C.POINTER(C.c_void_p)
    u64p = C.POINTER(C.c_uint64)
    lib.mo_mesh_for_chunk.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Cache), vpp, u64p, vpp, u64p]
    lib.mo_mesh_for_chunk.restype = None
    lib.mo_free.argtypes = [C.c_void_p]
    lib.mo_push_block.argtypes = [C.c_float, C.c_float, C.c_float, C.POINTER(Bti), C.c_float,
                                  C.c_uint32, vpp, u64p]
    lib.mo_count_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(Cache),
                                   C.c_void_p, C.c_void_p]
    lib.mo_mesh_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(Cache),
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_uint32, C.POINTER(C.c_double)]
    lib.mo_mesh_batch.restype = C.c_uint64
    lib.mo_sizeof_vertex.restype = C.c_uint32
    lib.mo_sizeof_cache.restype = C.c_uint32
    assert lib.mo_sizeof_vertex() == 36
    assert lib.mo_sizeof_cache() == C.sizeof(Cache)
    _lib = lib
    return lib


def cache_from_entries(entries, n_known: int, fallback) -> Cache:
    """entries: rows (kind, sides, top, bottom, crossed) for ids < n_known
    (kind 0 none / 1 blocky / 2 crossed); fallback = (sides, top, bottom)."""
    c = Cache()
    c.fallback_id = Bti(*fallback)
    c.len = n_known
    for i in range(n_known):
        kind, s, t, b, x = entries[i]
        c.style[i] = kind
        c.blocky[i] = Bti(s, t, b)
        c.crossed[i] = x
    return c


def cache_from_texture_id_cache(tic) -> Cache:
    """tic: oracle.block_table.TextureIdCache"""
    return cache_from_entries(tic.to_entries(), len(tic.styles), tic.fallback)


def _take(ptr, n):
    if n == 0:
        out = np.zeros(0, dtype=VERTEX_DTYPE)
    else:
        buf = (C.c_uint8 * (n * 36)).from_address(ptr.value)
        out = np.frombuffer(buf, dtype=VERTEX_DTYPE).copy()
    if ptr.value:
        load().mo_free(ptr)
    return out


def mesh_for_chunk(offs, chunk, cache: Cache):
    """mimas_meshgen::mesh_for_chunk (lib.rs:231-232): returns (intransparent, transparent)."""
    lib = load()
    offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(3)
    chunk = np.ascontiguousarray(chunk, dtype=np.uint8).reshape(4096)
    pi, pt = C.c_void_p(), C.c_void_p()
    ni, nt = C.c_uint64(), C.c_uint64()
    lib.mo_mesh_for_chunk(offs.ctypes.data, chunk.ctypes.data, C.byref(cache),
                          C.byref(pi), C.byref(ni), C.byref(pt), C.byref(nt))
    return _take(pi, ni.value), _take(pt, nt.value)


def push_block(pos, ids, siz: float, blocked_mask: int):
    lib = load()
    p, n = C.c_void_p(), C.c_uint64()
    b = Bti(*ids)
    lib.mo_push_block(pos[0], pos[1], pos[2], C.byref(b), siz, blocked_mask, C.byref(p), C.byref(n))
    return _take(p, n.value)


def _halo_ptr(halo, n):
    """halo: None (reference semantics) or [n, 6, 256] block ids (optional extension, see blocked() in the C file)."""
    if halo is None:
        return None, None
    halo = np.ascontiguousarray(halo, dtype=np.uint8).reshape(n, 1536)
    return halo, C.c_void_p(halo.ctypes.data)


def count_batch(chunks, offs, cache: Cache, halo=None):
    lib = load()
    chunks = np.ascontiguousarray(chunks, dtype=np.uint8).reshape(-1, 4096)
    n = chunks.shape[0]
    offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(n, 3)
    ci = np.zeros(n, dtype=np.uint32)
    ct = np.zeros(n, dtype=np.uint32)
    keep, hp = _halo_ptr(halo, n)
    lib.mo_count_batch_halo(C.c_void_p(chunks.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_uint64(n), C.byref(cache),
                            C.c_void_p(ci.ctypes.data), C.c_void_p(ct.ctypes.data), hp)
    return ci, ct


def mesh_batch(chunks, offs, cache: Cache, threads: int = 1, discard: bool = False, halo=None):
    """Packed batch result in chunk order: dict(intransparent, transparent, count_i, count_t,
    first_i, first_t, seconds).  discard=True builds every mesh but keeps nothing (the CPU
    baseline's timing mode)."""
    lib = load()
    chunks = np.ascontiguousarray(chunks, dtype=np.uint8).reshape(-1, 4096)
    n = chunks.shape[0]
    offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(n, 3)
    secs = C.c_double()
    keep, hp = _halo_ptr(halo, n)
    lib.mo_mesh_batch_halo.restype = C.c_uint64
    if discard:
        verts = lib.mo_mesh_batch_halo(C.c_void_p(chunks.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_uint64(n), C.byref(cache),
                                       None, None, None, None, C.c_uint32(threads), C.byref(secs), hp)
        return dict(vertices=int(verts), seconds=secs.value)
    ci, ct = count_batch(chunks, offs, cache, halo)
    fi = np.zeros(n, dtype=np.uint64)
    ft = np.zeros(n, dtype=np.uint64)
    if n:
        fi[1:] = np.cumsum(ci[:-1], dtype=np.uint64)
        ft[1:] = np.cumsum(ct[:-1], dtype=np.uint64)
    vi = np.zeros(int(ci.sum(dtype=np.uint64)), dtype=VERTEX_DTYPE)
    vt = np.zeros(int(ct.sum(dtype=np.uint64)), dtype=VERTEX_DTYPE)
    lib.mo_mesh_batch_halo(C.c_void_p(chunks.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_uint64(n), C.byref(cache),
                           C.c_void_p(vi.ctypes.data), C.c_void_p(fi.ctypes.data), C.c_void_p(vt.ctypes.data),
                           C.c_void_p(ft.ctypes.data), C.c_uint32(threads), C.byref(secs), hp)
    return dict(intransparent=vi, transparent=vt, count_i=ci, count_t=ct,
                first_i=fi, first_t=ft, seconds=secs.value)


# ---- render oracle (oracle/render_oracle.c): culling + vertex shader restatement ----------------------
class RoCamera(C.Structure):
    _fields_ = [("vmatrix", C.c_float * 16), ("pmatrix", C.c_float * 16), ("pos", C.c_float * 3),
                ("viewing_range", C.c_float)]


SHADED_DTYPE = np.dtype([("clip", "<f4", (4,)), ("vposition", "<f4", (4,)), ("vlamb", "<f4"),
                         ("vtex_pos", "<f4", (2,)), ("vtex_ind", "<u4")])
_rlib = None


def _render():
    global _rlib
    if _rlib is None:
        subprocess.run(["make", "-C", HERE, "-s"], check=True)
        _rlib = C.CDLL(os.path.join(HERE, "librender_oracle.so"))
        _rlib.ro_cull_chunks.restype = C.c_uint32
    return _rlib


def _ro_camera(vmatrix, pmatrix, pos, viewing_range) -> RoCamera:
    cam = RoCamera()
    cam.vmatrix[:] = [float(v) for v in np.asarray(vmatrix, dtype=np.float32).reshape(-1)]
    cam.pmatrix[:] = [float(v) for v in np.asarray(pmatrix, dtype=np.float32).reshape(-1)]
    cam.pos[:] = [float(v) for v in pos]
    cam.viewing_range = float(viewing_range)
    return cam


def cull_chunks(offs, vmatrix, pmatrix, pos, viewing_range) -> np.ndarray:
    offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(-1, 3)
    draw = np.zeros(max(len(offs), 1), dtype=np.uint32)
    cam = _ro_camera(vmatrix, pmatrix, pos, viewing_range)
    n = _render().ro_cull_chunks(offs.ctypes.data_as(C.c_void_p), C.c_uint32(len(offs)), C.byref(cam), draw.ctypes.data_as(C.c_void_p))
    return draw[:n].copy()


def vertex_stage(verts, vmatrix, pmatrix) -> np.ndarray:
    verts = np.ascontiguousarray(verts, dtype=VERTEX_DTYPE)
    out = np.zeros(len(verts), dtype=SHADED_DTYPE)
    cam = _ro_camera(vmatrix, pmatrix, (0, 0, 0), 0)
    _render().ro_vertex_stage(verts.ctypes.data_as(C.c_void_p), C.c_uint64(len(verts)), C.byref(cam), out.ctypes.data_as(C.c_void_p))
    return out


def frustum_planes(vmatrix, pmatrix) -> np.ndarray:
    pl = np.zeros((6, 4), dtype=np.float32)
    v = np.ascontiguousarray(vmatrix, dtype=np.float32).reshape(-1)
    p = np.ascontiguousarray(pmatrix, dtype=np.float32).reshape(-1)
    _render().ro_frustum(v.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p), pl.ctypes.data_as(C.c_void_p))
    return pl


# ---- raster oracle (oracle/raster_oracle.c): clipping, rasterisation, depth test, fragment shader --------
class RaFrame(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("fog_near", C.c_float), ("fog_far", C.c_float),
                ("clear", C.c_float * 4)]


CLEAR_COLOR = (0.05, 0.01, 0.6, 0.0)  # client.rs:675


def raster_frame(shaded, width, height, fog_near, fog_far, tex, clear=CLEAR_COLOR, want_keys=False):
    """shaded: SHADED_DTYPE triangle list; tex: [layers, size, size, 4] float32.  Returns the image as
    uint32 [height, width] (R | G<<8 | B<<16 | A<<24, row 0 at the bottom) and optionally the
    depth24 << 32 | colour keys."""
    shaded = np.ascontiguousarray(shaded, dtype=SHADED_DTYPE)
    tex = np.ascontiguousarray(tex, dtype=np.float32)
    assert tex.ndim == 4 and tex.shape[1] == tex.shape[2] and tex.shape[3] == 4
    f = RaFrame(width, height, fog_near, fog_far)
    f.clear[:] = [float(c) for c in clear]
    img = np.zeros((height, width), dtype=np.uint32)
    keys = np.zeros((height, width), dtype=np.uint64)
    rc = _render().ra_raster_frame(shaded.ctypes.data_as(C.c_void_p), C.c_uint64(len(shaded)), C.byref(f),
                                   tex.ctypes.data_as(C.c_void_p), C.c_uint32(tex.shape[0]), C.c_uint32(tex.shape[1]),
                                   img.ctypes.data_as(C.c_void_p), keys.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return (img, keys) if want_keys else img
