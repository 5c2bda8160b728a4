"""Host-side mirror of the `mimas-meshgen` crate API on top of libmimas_meshgen_cuda.so.

Same names and argument meaning as the reference (mimas-meshgen/lib.rs):
  TextureId, BlockTextureIds (:44-71), MeshDrawStyle (:28-42), TextureIdObtainer (:81-84),
  TextureIdCache.from_hdl / get_bl_tex_ids / get_mesh_draw_style / get_inv_texture_id (:86-118),
  push_block (:183-219), ChunkMesh (:226-229), mesh_for_chunk (:231-232).
The mesher itself runs only on the GPU: importing this module without the CUDA library, or
creating a Mesher without a CUDA device, raises -- there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple, Union

import numpy as np

from . import game_params as gp
from ._lib import MMError, check, lib

CHUNKSIZE = 16  # mimas-common/map.rs:11

# #[repr(C)] Vertex (lib.rs:18-24): include/mimas_meshgen_cuda.h mm_vertex
VERTEX_DTYPE = np.dtype([("tex_ind", "<u2"), ("pad0", "<u2"), ("tex_pos", "<f4", (2,)),
                         ("position", "<f4", (3,)), ("normal", "<f4", (3,))])
RANGE_DTYPE = np.dtype([("first_vertex", "<u8"), ("vertex_count", "<u4"), ("reserved", "<u4")])
BLOCK_ENTRY_DTYPE = np.dtype([("id_sides", "<u2"), ("id_top", "<u2"), ("id_bottom", "<u2"), ("id_crossed", "<u2"),
                              ("kind", "u1"), ("reserved", "u1", (3,))])
CUBE_DTYPE = np.dtype([("pos", "<f4", (3,)), ("siz", "<f4"), ("id_sides", "<u2"), ("id_top", "<u2"),
                       ("id_bottom", "<u2"), ("blocked_mask", "<u2")])
assert VERTEX_DTYPE.itemsize == 36 and RANGE_DTYPE.itemsize == 16
assert BLOCK_ENTRY_DTYPE.itemsize == 12 and CUBE_DTYPE.itemsize == 24

KIND_NONE, KIND_BLOCKY, KIND_CROSSED = 0, 1, 2

TextureId = int  # #[repr(transparent)] struct TextureId(pub u16), lib.rs:14-16


@dataclass(frozen=True)
class BlockTextureIds:
    id_sides: TextureId
    id_top: TextureId
    id_bottom: TextureId

    @staticmethod
    def uniform(id: TextureId) -> "BlockTextureIds":
        return BlockTextureIds(id, id, id)

    @staticmethod
    def new_tb(id_top_bottom: TextureId, id_sides: TextureId) -> "BlockTextureIds":
        return BlockTextureIds(id_sides, id_top_bottom, id_top_bottom)

    @staticmethod
    def new(id_sides: TextureId, id_top: TextureId, id_bottom: TextureId) -> "BlockTextureIds":
        return BlockTextureIds(id_sides, id_top, id_bottom)


@dataclass(frozen=True)
class Blocky:
    ids: BlockTextureIds


@dataclass(frozen=True)
class CrossedStyle:
    id: TextureId


MeshDrawStyle = Union[Blocky, CrossedStyle]


class TextureIdObtainer:
    """trait TextureIdObtainer, lib.rs:81-84."""

    def mesh_draw_st(self, ds: gp.DrawStyle) -> MeshDrawStyle:
        raise NotImplementedError

    def inv_texture_id(self, path: str) -> TextureId:
        raise NotImplementedError


class HeadlessAssets(TextureIdObtainer):
    """Texture-id allocation of mimas-client/assets.rs:140-201 without loading any pixels:
    every add_asset hands out the next index (assets.rs:140-144)."""

    def __init__(self):
        self.count = 0

    def _add_asset(self) -> TextureId:
        i = self.count
        if i > 0xFFFF:
            raise OverflowError("TextureId is a u16")
        self.count += 1
        return i

    def mesh_draw_st(self, ds: gp.DrawStyle) -> MeshDrawStyle:
        if isinstance(ds, gp.Colored):                      # assets.rs:148-152
            full = self._add_asset()
            half = self._add_asset()                        # colorh(): darker sides
            return Blocky(BlockTextureIds.new_tb(full, half))
        if isinstance(ds, gp.Crossed):                      # assets.rs:153-158
            return CrossedStyle(self._add_asset())
        if isinstance(ds, gp.Texture):                      # assets.rs:159-164
            return Blocky(BlockTextureIds.uniform(self._add_asset()))
        if isinstance(ds, gp.TextureSidesTop):              # assets.rs:165-173
            s = self._add_asset()
            tb = self._add_asset()
            return Blocky(BlockTextureIds.new_tb(tb, s))
        if isinstance(ds, gp.TextureSidesTopBottom):        # assets.rs:174-186
            s = self._add_asset()
            t = self._add_asset()
            b = self._add_asset()
            return Blocky(BlockTextureIds.new(s, t, b))
        raise TypeError(f"not a DrawStyle: {ds!r}")

    def inv_texture_id(self, path: str) -> TextureId:       # assets.rs:189-194
        return self._add_asset()


class TextureIdCache:
    """lib.rs:73-118.  `to_table()` flattens it into the 256-row device block table."""

    def __init__(self, fallback_id: BlockTextureIds, mesh_draw_styles: Sequence[Optional[MeshDrawStyle]],
                 inv_texture_ids: Sequence[Optional[TextureId]] = ()):
        self.fallback_id = fallback_id
        self.mesh_draw_styles = list(mesh_draw_styles)
        self.block_texture_ids = [s.ids if isinstance(s, Blocky) else None for s in self.mesh_draw_styles]
        self.inv_texture_ids = list(inv_texture_ids)
        if len(self.mesh_draw_styles) > 256:
            raise ValueError("MapBlock ids are u8")

    @classmethod
    def from_hdl(cls, hdl: gp.GameParams, ob: TextureIdObtainer) -> "TextureIdCache":
        fb = ob.mesh_draw_st(gp.DEFAULT_DRAW_STYLE)  # Colored([0,0,0,1]), lib.rs:88
        assert isinstance(fb, Blocky)
        styles = [ob.mesh_draw_st(p.draw_style) if p.draw_style is not None else None for p in hdl.block_params]
        inv = [ob.inv_texture_id(p.inv_texture) if p.inv_texture is not None else None for p in hdl.block_params]
        return cls(fb.ids, styles, inv)

    def get_bl_tex_ids(self, bl: int) -> Optional[BlockTextureIds]:
        return self.block_texture_ids[bl] if bl < len(self.block_texture_ids) else self.fallback_id

    def get_mesh_draw_style(self, bl: int) -> Optional[MeshDrawStyle]:
        return self.mesh_draw_styles[bl] if bl < len(self.mesh_draw_styles) else None

    def get_inv_texture_id(self, bl: int) -> Optional[TextureId]:
        return self.inv_texture_ids[bl] if bl < len(self.inv_texture_ids) else None

    def to_table(self) -> np.ndarray:
        known = np.zeros(len(self.mesh_draw_styles), dtype=BLOCK_ENTRY_DTYPE)
        for i, s in enumerate(self.mesh_draw_styles):
            if isinstance(s, Blocky):
                known[i] = (s.ids.id_sides, s.ids.id_top, s.ids.id_bottom, 0, KIND_BLOCKY, (0, 0, 0))
            elif isinstance(s, CrossedStyle):
                known[i] = (0, 0, 0, s.id, KIND_CROSSED, (0, 0, 0))
        fb = (C.c_uint16 * 3)(self.fallback_id.id_sides, self.fallback_id.id_top, self.fallback_id.id_bottom)
        out = np.zeros(256, dtype=BLOCK_ENTRY_DTYPE)
        check(lib().mm_table_from_cache(fb, known.ctypes.data, len(known), out.ctypes.data), None)
        return out

    @classmethod
    def from_table_rows(cls, rows: Sequence[Tuple[int, int, int, int, int]], fallback: Tuple[int, int, int]):
        """rows of (kind, sides, top, bottom, crossed) for the known ids; fallback (sides, top, bottom)."""
        styles: List[Optional[MeshDrawStyle]] = []
        for kind, s, t, b, x in rows:
            styles.append(Blocky(BlockTextureIds(s, t, b)) if kind == KIND_BLOCKY
                          else CrossedStyle(x) if kind == KIND_CROSSED else None)
        return cls(BlockTextureIds(*fallback), styles)


@dataclass
class ChunkMesh:
    """lib.rs:226-229: two vertex vectors (structured numpy arrays of VERTEX_DTYPE)."""
    intransparent: np.ndarray
    transparent: np.ndarray


@dataclass
class BatchMesh:
    """Packed result of a batch: chunk i owns intransparent[ranges_i[i].first_vertex :][: vertex_count]."""
    intransparent: np.ndarray
    transparent: np.ndarray
    ranges_i: np.ndarray
    ranges_t: np.ndarray

    def chunk(self, i: int) -> ChunkMesh:
        a, b = self.ranges_i[i], self.ranges_t[i]
        return ChunkMesh(self.intransparent[int(a["first_vertex"]):int(a["first_vertex"]) + int(a["vertex_count"])],
                         self.transparent[int(b["first_vertex"]):int(b["first_vertex"]) + int(b["vertex_count"])])


class _BatchOut(C.Structure):
    _fields_ = [("d_intransparent", C.c_void_p), ("d_transparent", C.c_void_p), ("d_ranges_i", C.c_void_p),
                ("d_ranges_t", C.c_void_p), ("n_intransparent", C.c_uint64), ("n_transparent", C.c_uint64),
                ("n_chunks", C.c_uint32), ("launches", C.c_uint32), ("kernel_ms", C.c_float),
                ("reserved", C.c_uint32)]


class _HostMesh(C.Structure):
    _fields_ = [("intransparent", C.c_void_p), ("transparent", C.c_void_p), ("ranges_i", C.c_void_p),
                ("ranges_t", C.c_void_p), ("n_intransparent", C.c_uint64), ("n_transparent", C.c_uint64),
                ("n_chunks", C.c_uint32), ("path", C.c_uint32)]


HOST_PATH_NAMES = {1: "mapped", 2: "pipelined", 3: "plain"}  # MM_PATH_*


def _copy_from(ptr: Optional[int], count: int, dtype: np.dtype) -> np.ndarray:
    if not count:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_uint8 * (count * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()


class Mesher:
    """One mm_ctx: the state the reference's meshgen thread owns (mimas-client/client.rs:258-267)."""

    def __init__(self, device: int = 0, cache: Optional[TextureIdCache] = None):
        self._lib = lib()
        self._ctx = C.c_void_p()
        rc = self._lib.mm_ctx_create(device, C.byref(self._ctx))
        if rc != 0:
            self._ctx = C.c_void_p()
            raise MMError(rc, "mm_ctx_create failed (no CUDA device? this package has no CPU fallback)")
        self.device = device
        self._cache = None
        if cache is not None:
            self.set_cache(cache)

    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            self._lib.mm_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc: int):
        check(rc, self._ctx)

    # -- block table ---------------------------------------------------------------------
    def set_cache(self, cache: TextureIdCache):
        self.set_table(cache.to_table())
        self._cache = cache

    def set_table(self, table: np.ndarray):
        table = np.ascontiguousarray(table, dtype=BLOCK_ENTRY_DTYPE)
        if table.shape != (256,):
            raise ValueError("block table needs 256 rows")
        self._ck(self._lib.mm_set_block_table(self._ctx, table.ctypes.data))
        self._cache = None

    # -- crate API ------------------------------------------------------------------------
    def mesh_for_chunk(self, offs, chunk, cache: Optional[TextureIdCache] = None) -> ChunkMesh:
        """mesh_for_chunk(offs :Vector3<isize>, chunk :&MapChunkData, cache :&TextureIdCache), lib.rs:231."""
        if cache is not None and cache is not self._cache:
            self.set_cache(cache)
        offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(3)
        chunk = np.ascontiguousarray(chunk, dtype=np.uint8).reshape(4096)
        pi, pt = C.c_void_p(), C.c_void_p()
        ni, nt = C.c_uint64(), C.c_uint64()
        self._ck(self._lib.mm_mesh_for_chunk(self._ctx, offs.ctypes.data, chunk.ctypes.data,
                                             C.byref(pi), C.byref(ni), C.byref(pt), C.byref(nt)))
        return ChunkMesh(_copy_from(pi.value, ni.value, VERTEX_DTYPE), _copy_from(pt.value, nt.value, VERTEX_DTYPE))

    def mesh_batch(self, offs, chunks, halo=None) -> BatchMesh:
        """Host in, host out: n chunks meshed in one call (batch form of mesh_for_chunk).
        `halo` ([n, 6, 256] block ids, see mm_mesh_batch in the header) switches on the halo extension: faces
        against blocky cells of the neighbouring chunks are culled.  None = the reference's semantics."""
        chunks = np.ascontiguousarray(chunks, dtype=np.uint8).reshape(-1, 4096)
        n = chunks.shape[0]
        offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(n, 3)
        if halo is not None:
            halo = np.ascontiguousarray(halo, dtype=np.uint8).reshape(n, 1536)
            out = _BatchOut()
            self._ck(self._lib.mm_mesh_batch(self._ctx, C.c_void_p(chunks.ctypes.data), C.c_void_p(offs.ctypes.data),
                                             C.c_void_p(halo.ctypes.data), n, C.byref(out)))
            return self.copy_out(out)
        hm = _HostMesh()
        self._ck(self._lib.mm_mesh_batch_host(self._ctx, chunks.ctypes.data, offs.ctypes.data, n, C.byref(hm)))
        return BatchMesh(_copy_from(hm.intransparent, hm.n_intransparent, VERTEX_DTYPE),
                         _copy_from(hm.transparent, hm.n_transparent, VERTEX_DTYPE),
                         _copy_from(hm.ranges_i, n, RANGE_DTYPE), _copy_from(hm.ranges_t, n, RANGE_DTYPE))

    def push_block(self, pos, block_ids: BlockTextureIds, siz: float, blocked) -> np.ndarray:
        """push_block(r, [x,y,z], block_ids, siz, blocked), lib.rs:184: returns the vertices that would
        be appended to `r`.  `blocked` is the closure: called with [dx,dy,dz] in source order."""
        order = ([0, 0, -1], [0, -1, 0], [-1, 0, 0], [0, 0, 1], [0, 1, 0], [1, 0, 0])
        mask = 0
        for i, o in enumerate(order):
            if blocked(o):
                mask |= 1 << i
        cube = np.zeros(1, dtype=CUBE_DTYPE)
        cube[0] = (tuple(np.float32(v) for v in pos), np.float32(siz), block_ids.id_sides, block_ids.id_top,
                   block_ids.id_bottom, mask)
        return self.push_blocks(cube)

    def push_blocks(self, cubes: np.ndarray) -> np.ndarray:
        cubes = np.ascontiguousarray(cubes, dtype=CUBE_DTYPE)
        out = np.zeros(36 * len(cubes), dtype=VERTEX_DTYPE)
        n_out = C.c_uint64()
        self._ck(self._lib.mm_push_blocks(self._ctx, cubes.ctypes.data, len(cubes), out.ctypes.data, C.byref(n_out)))
        return out[:n_out.value].copy()

    # -- device-resident batch path (benchmarks, server-side batch submission) ---------------
    def set_stream(self, cuda_stream: int):
        self._ck(self._lib.mm_ctx_set_stream(self._ctx, C.c_void_p(cuda_stream)))

    def reserve(self, max_chunks: int, intransparent_vertices: int, transparent_vertices: int):
        self._ck(self._lib.mm_reserve(self._ctx, max_chunks, intransparent_vertices, transparent_vertices))

    def mesh_batch_device(self, d_chunks: int, d_offs: int, n: int, d_halo: Optional[int] = None) -> _BatchOut:
        """Synchronous mesh of device-resident inputs (raw device pointers)."""
        out = _BatchOut()
        self._ck(self._lib.mm_mesh_batch(self._ctx, C.c_void_p(d_chunks), C.c_void_p(d_offs),
                                         C.c_void_p(d_halo) if d_halo else None, n, C.byref(out)))
        return out

    def mesh_batch_async(self, d_chunks: int, d_offs: int, n: int, d_halo: Optional[int] = None):
        if d_halo:
            self._ck(self._lib.mm_mesh_batch_halo_async(self._ctx, C.c_void_p(d_chunks), C.c_void_p(d_offs),
                                                        C.c_void_p(d_halo), n))
        else:
            self._ck(self._lib.mm_mesh_batch_async(self._ctx, C.c_void_p(d_chunks), C.c_void_p(d_offs), n))

    def gather_halos(self, d_chunks: int, dims, d_halo: int, outside_id: int = 0):
        """Halo slabs (n * 1536 bytes at d_halo) of a device-resident box of dims = (nx, ny, nz) chunks in
        gen_terrain order; `outside_id` is what lies beyond the box."""
        self._ck(self._lib.mm_gather_halos(self._ctx, C.c_void_p(d_chunks), dims[0], dims[1], dims[2], outside_id,
                                           C.c_void_p(d_halo)))

    # -- output ring: several batches retrievable while later ones are meshed (mm_ring_configure) ----------
    def ring_configure(self, slots: int):
        self._ck(self._lib.mm_ring_configure(self._ctx, slots))

    def submit(self, d_chunks: int, d_offs: int, n: int, d_halo: Optional[int] = None) -> int:
        """mm_mesh_batch_submit: enqueue one batch into the next ring slot; returns its ticket."""
        t = C.c_uint64()
        self._ck(self._lib.mm_mesh_batch_submit(self._ctx, C.c_void_p(d_chunks), C.c_void_p(d_offs),
                                                C.c_void_p(d_halo) if d_halo else None, n, C.byref(t)))
        return t.value

    @staticmethod
    def prepare_many(d_chunks: Sequence[int], d_offs: Sequence[int], n: Sequence[int]):
        """The argument arrays of mm_mesh_batches_submit, built once (outside a timed region: marshalling a dozen
        pointers in Python costs more than the call)."""
        k = len(n)
        return ((C.c_void_p * k)(*d_chunks), (C.c_void_p * k)(*d_offs), (C.c_uint32 * k)(*n), k)

    def submit_many(self, d_chunks=None, d_offs=None, n=None, repeat: int = 1, prepared=None) -> int:
        """mm_mesh_batches_submit: k batches (x repeat, round robin) enqueued by ONE foreign call; returns the
        first ticket (the others follow consecutively).  `prepared` = the result of prepare_many."""
        pc, po, pn, k = prepared if prepared is not None else self.prepare_many(d_chunks, d_offs, n)
        t = C.c_uint64()
        self._ck(self._lib.mm_mesh_batches_submit(self._ctx, pc, po, pn, k, repeat, C.byref(t)))
        return t.value

    def batch_wait(self, ticket: int) -> _BatchOut:
        out = _BatchOut()
        self._ck(self._lib.mm_batch_wait(self._ctx, ticket, C.byref(out)))
        return out

    def batch_poll(self, ticket: int) -> bool:
        d = C.c_int()
        self._ck(self._lib.mm_batch_poll(self._ctx, ticket, C.byref(d)))
        return bool(d.value)

    def copy_out_ticket(self, ticket: int) -> BatchMesh:
        out = self.batch_wait(ticket)
        vi = np.zeros(out.n_intransparent, dtype=VERTEX_DTYPE)
        vt = np.zeros(out.n_transparent, dtype=VERTEX_DTYPE)
        ri = np.zeros(out.n_chunks, dtype=RANGE_DTYPE)
        rt = np.zeros(out.n_chunks, dtype=RANGE_DTYPE)
        self._ck(self._lib.mm_copy_out_ticket(self._ctx, ticket, vi.ctypes.data, vt.ctypes.data, ri.ctypes.data,
                                              rt.ctypes.data))
        return BatchMesh(vi, vt, ri, rt)

    def draw_ticket(self, ticket: int, cam):
        """One frame over the batch of `ticket` on the consumer stream (later batches keep meshing)."""
        from .render import DrawOut
        out = DrawOut()
        cc = cam.to_c() if hasattr(cam, "to_c") else cam   # a Camera, or the mm_camera it was turned into
        self._ck(self._lib.mm_draw_ticket(self._ctx, ticket, C.byref(cc), C.byref(out)))
        return out

    def draw_ticket_async(self, ticket: int, cam):
        cc = cam.to_c() if hasattr(cam, "to_c") else cam   # a Camera, or the mm_camera it was turned into
        self._ck(self._lib.mm_draw_ticket_async(self._ctx, ticket, C.byref(cc)))

    def draw_collect(self):
        from .render import DrawOut
        out = DrawOut()
        self._ck(self._lib.mm_draw_collect(self._ctx, C.byref(out)))
        return out

    def batch_finish(self) -> _BatchOut:
        out = _BatchOut()
        self._ck(self._lib.mm_batch_finish(self._ctx, C.byref(out)))
        return out

    def copy_out(self, out: _BatchOut) -> BatchMesh:
        vi = np.zeros(out.n_intransparent, dtype=VERTEX_DTYPE)
        vt = np.zeros(out.n_transparent, dtype=VERTEX_DTYPE)
        ri = np.zeros(out.n_chunks, dtype=RANGE_DTYPE)
        rt = np.zeros(out.n_chunks, dtype=RANGE_DTYPE)
        self._ck(self._lib.mm_copy_out(self._ctx, vi.ctypes.data, vt.ctypes.data, ri.ctypes.data, rt.ctypes.data))
        return BatchMesh(vi, vt, ri, rt)

    def gen_terrain_device(self, seed: int, c0, dims, d_chunks: int, d_offs: int):
        self._ck(self._lib.mm_gen_terrain_device(self._ctx, seed, c0[0], c0[1], c0[2], dims[0], dims[1], dims[2],
                                                 C.c_void_p(d_chunks), C.c_void_p(d_offs)))

    def gen_mapgen_device(self, seed: int, c0, dims, d_chunks: int, d_offs: int, params=None, d_tree_counts: int = 0):
        """mimas-server mapgen phase one (mapgen.rs:64-245) for a box of chunks, on the device (mm_gen_mapgen_device).
        `params`: mimas_b200.mapgen.MapgenParams (default: the default game's ores and plants)."""
        from . import mapgen as mg
        cp = mg.to_c(params if params is not None else mg.default_mapgen_params())
        self._ck(self._lib.mm_gen_mapgen_device(self._ctx, seed, C.byref(cp), c0[0], c0[1], c0[2], dims[0], dims[1], dims[2],
                                                C.c_void_p(d_chunks), C.c_void_p(d_offs) if d_offs else None,
                                                C.c_void_p(d_tree_counts) if d_tree_counts else None))

    def dev_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._ck(self._lib.mm_dev_alloc(self._ctx, nbytes, C.byref(p)))
        return p.value

    def dev_free(self, ptr: int):
        self._ck(self._lib.mm_dev_free(self._ctx, C.c_void_p(ptr)))

    def upload(self, d_ptr: int, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        self._ck(self._lib.mm_dev_upload(self._ctx, C.c_void_p(d_ptr), arr.ctypes.data, arr.nbytes))

    def download(self, d_ptr: int, nbytes: int) -> np.ndarray:
        out = np.zeros(nbytes, dtype=np.uint8)
        self._ck(self._lib.mm_dev_download(self._ctx, out.ctypes.data, C.c_void_p(d_ptr), nbytes))
        return out

    def host_alloc(self, nbytes: int) -> np.ndarray:
        """Page-locked host buffer as a uint8 array (free with host_free(arr))."""
        p = C.c_void_p()
        self._ck(self._lib.mm_host_alloc(self._ctx, nbytes, C.byref(p)))
        return np.frombuffer((C.c_uint8 * max(nbytes, 1)).from_address(p.value), dtype=np.uint8)[:nbytes]

    def host_free(self, arr: np.ndarray):
        self._ck(self._lib.mm_host_free(self._ctx, C.c_void_p(arr.ctypes.data)))

    def mesh_batch_host_raw(self, chunks_ptr: int, offs_ptr: int, n: int) -> _HostMesh:
        """mm_mesh_batch_host on raw host pointers; the result points into library-owned pinned memory."""
        hm = _HostMesh()
        self._ck(self._lib.mm_mesh_batch_host(self._ctx, C.c_void_p(chunks_ptr), C.c_void_p(offs_ptr), n, C.byref(hm)))
        return hm

    # -- headless draw: culling + vertex shader on the device (render.py, csrc/mm_render.cu) ---------
    def cull_chunks(self, offs, cam) -> np.ndarray:
        """client.rs:679-696: indices (ascending) of the chunks a frame would draw."""
        offs = np.ascontiguousarray(offs, dtype=np.int64).reshape(-1, 3)
        draw = np.zeros(max(len(offs), 1), dtype=np.uint32)
        n = C.c_uint32()
        cc = cam.to_c() if hasattr(cam, "to_c") else cam   # a Camera, or the mm_camera it was turned into
        self._ck(self._lib.mm_cull_chunks(self._ctx, offs.ctypes.data, len(offs), C.byref(cc), draw.ctypes.data, C.byref(n)))
        return draw[:n.value].copy()

    def vertex_stage(self, d_in: int, n: int, cam, d_out: int):
        cc = cam.to_c() if hasattr(cam, "to_c") else cam   # a Camera, or the mm_camera it was turned into
        self._ck(self._lib.mm_vertex_stage(self._ctx, C.c_void_p(d_in), C.c_uint64(n), C.byref(cc), C.c_void_p(d_out)))

    def draw_batch(self, cam):
        """One frame over the last finished device batch: returns render.DrawOut (device pointers)."""
        from .render import DrawOut
        out = DrawOut()
        cc = cam.to_c() if hasattr(cam, "to_c") else cam   # a Camera, or the mm_camera it was turned into
        self._ck(self._lib.mm_draw_batch(self._ctx, C.byref(cc), C.byref(out)))
        return out

    # -- headless rasteriser + fragment shader (csrc/mm_raster.cu) ---------------------------------------
    def set_textures(self, tex):
        """The texture array (assets.rs:194-211): [layers, size, size, 4] float32, layer = TextureId."""
        tex = np.ascontiguousarray(tex, dtype=np.float32)
        if tex.ndim != 4 or tex.shape[1] != tex.shape[2] or tex.shape[3] != 4:
            raise ValueError("textures must be [layers, size, size, 4]")
        self._ck(self._lib.mm_set_textures(self._ctx, tex.ctypes.data, tex.shape[0], tex.shape[1]))

    def raster_frame(self, d_verts: int, n_vertices: int, width: int, height: int, fog_near: float = 40.0,
                     fog_far: float = 60.0, clear=(0.05, 0.01, 0.6, 0.0), want_ms: bool = False):
        """Shaded triangle list in device memory -> uint32 [height, width] image (R | G<<8 | B<<16 | A<<24,
        row 0 at the bottom as in GL): clipping, culling, rasterisation, depth test, fragment-shader.glsl."""
        from .render import CFrame
        f = CFrame(width, height, fog_near, fog_far)
        f.clear[:] = [float(c) for c in clear]
        d_img = self.dev_alloc(4 * width * height)
        try:
            ms = C.c_float()
            self._ck(self._lib.mm_raster_frame(self._ctx, C.c_void_p(d_verts), C.c_uint64(n_vertices), C.byref(f),
                                               C.c_void_p(d_img), C.byref(ms)))
            img = self.download(d_img, 4 * width * height).view(np.uint32).reshape(height, width).copy()
        finally:
            self.dev_free(d_img)
        return (img, ms.value) if want_ms else img

    def render_frame(self, cam, width: int, height: int, fog_near: float = 40.0, fog_far: float = 60.0, want_ms=False):
        """A whole headless frame over the last finished device batch: cull, vertex shader, rasterise."""
        d = self.draw_batch(cam)
        return self.raster_frame(d.d_shaded or 0, int(d.n_shaded_i + d.n_shaded_t), width, height, fog_near, fog_far,
                                 want_ms=want_ms)

    def debug_trace(self, pairs: int) -> np.ndarray:
        """Device timeline (ns) of the launch pairs since the previous call, shape (k, 2, 4): see
        mm_debug_trace in the header.  Re-arms the ring with `pairs` records (0 = off)."""
        armed = getattr(self, "_trace_pairs", 0)
        have = np.zeros((max(armed, 1), 8), dtype=np.uint64)
        n = C.c_uint64()
        self._ck(self._lib.mm_debug_trace(self._ctx, pairs, have.ctypes.data, C.byref(n)))
        self._trace_pairs = pairs
        return have[:min(int(n.value), armed)].reshape(-1, 2, 4)

    def launch_count(self) -> int:
        v = C.c_uint64()
        self._ck(self._lib.mm_launch_count(self._ctx, C.byref(v)))
        return v.value

    def sync(self):
        self._ck(self._lib.mm_ctx_sync(self._ctx))

    def mesh_batch_noalloc(self, d_chunks: int, d_offs: int, n: int, d_out_i: int, cap_i_vertices: int, d_out_t: int,
                           cap_t_vertices: int, d_ranges_i: int, d_ranges_t: int, d_totals: int):
        """mm_mesh_batch_noalloc: every buffer is the caller's, nothing is allocated, n <= 256; asynchronous."""
        self._ck(self._lib.mm_mesh_batch_noalloc(self._ctx, C.c_void_p(d_chunks), C.c_void_p(d_offs), n, C.c_void_p(d_out_i),
                                                 cap_i_vertices, C.c_void_p(d_out_t), cap_t_vertices, C.c_void_p(d_ranges_i),
                                                 C.c_void_p(d_ranges_t), C.c_void_p(d_totals)))

    def device_info(self) -> Tuple[int, int, int]:
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        self._ck(self._lib.mm_device_info(self._ctx, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value


def default_cache() -> TextureIdCache:
    """Block table of the default game (mimas-server/game-params.toml), shipped as data/default_block_table.json
    (derived by data/make_default_table.py) so that no reference checkout is needed at run time."""
    import json
    import os
    t = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "default_block_table.json")))
    return TextureIdCache.from_table_rows([tuple(r) for r in t["entries"]], tuple(t["fallback"]))


_default: Optional[Mesher] = None


def mesh_for_chunk(offs, chunk, cache: TextureIdCache) -> ChunkMesh:
    """Module-level drop-in with the crate's signature (lib.rs:231-232); uses one process-wide
    Mesher on device 0, like the single meshgen thread of the reference client."""
    global _default
    if _default is None:
        _default = Mesher(0)
    return _default.mesh_for_chunk(offs, chunk, cache)


_host_lib = None


def gen_terrain_host(seed: int, c0, dims) -> Tuple[np.ndarray, np.ndarray]:
    """Synthetic mapgen look-alike on the host (same bytes as the device generator).  Served by the host-only
    libmimas_terrain_host.so, so that CPU-side callers (the bench's CPU arms) never map the CUDA library."""
    global _host_lib
    if _host_lib is None:
        import os
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libmimas_terrain_host.so")
        if not os.path.exists(path):
            raise ImportError(f"{path} is missing: build it with `make -C mimas_b200/csrc`")
        h = C.CDLL(path)
        h.mm_gen_terrain_host.argtypes = [C.c_uint64, C.c_int64, C.c_int64, C.c_int64, C.c_uint32, C.c_uint32,
                                          C.c_uint32, C.c_void_p, C.c_void_p]
        h.mm_gen_terrain_host.restype = C.c_int
        _host_lib = h
    n = int(dims[0]) * int(dims[1]) * int(dims[2])
    chunks = np.zeros((n, 4096), dtype=np.uint8)
    offs = np.zeros((n, 3), dtype=np.int64)
    check(_host_lib.mm_gen_terrain_host(seed, c0[0], c0[1], c0[2], dims[0], dims[1], dims[2],
                                        chunks.ctypes.data, offs.ctypes.data), None)
    return chunks, offs
