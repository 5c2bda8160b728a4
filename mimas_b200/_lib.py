"""Loader of libmimas_meshgen_cuda.so (built in-tree by mimas_b200/csrc/Makefile).

Fails loudly when the library is missing: the mesher has no Python or CPU implementation.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# MM_LIB_PATH: a differently tuned build of the same library (scripts/variants.sh); never a different implementation
LIB_PATH = os.environ.get("MM_LIB_PATH") or os.path.join(HERE, "libmimas_meshgen_cuda.so")

STATUS_NAMES = {0: "MM_OK", 1: "MM_ERR_INVALID", 2: "MM_ERR_NO_DEVICE", 3: "MM_ERR_CUDA", 4: "MM_ERR_OOM",
                5: "MM_ERR_NO_TABLE", 6: "MM_ERR_NO_BATCH", 7: "MM_ERR_UNSUPPORTED"}

# every symbol include/mimas_meshgen_cuda.h declares
SYMBOLS = (
    "mm_abi_version", "mm_ctx_create", "mm_ctx_destroy", "mm_last_error_string", "mm_ctx_set_stream",
    "mm_set_block_table", "mm_table_from_cache", "mm_reserve", "mm_mesh_batch", "mm_mesh_batch_async",
    "mm_batch_finish", "mm_copy_out", "mm_mesh_batch_host", "mm_mesh_for_chunk", "mm_push_blocks",
    "mm_gen_terrain_host", "mm_gen_terrain_device", "mm_dev_alloc", "mm_dev_free", "mm_dev_upload",
    "mm_dev_download", "mm_ctx_sync", "mm_device_info", "mm_host_alloc", "mm_host_free", "mm_launch_count",
    "mm_debug_trace", "mm_vertex_stage", "mm_cull_chunks", "mm_draw_batch", "mm_mesh_batch_halo_async",
    "mm_gather_halos", "mm_set_textures", "mm_raster_frame", "mm_ring_configure", "mm_mesh_batch_submit",
    "mm_mesh_batches_submit", "mm_batch_wait", "mm_batch_poll", "mm_copy_out_ticket", "mm_draw_ticket",
    "mm_gen_mapgen_device", "mm_draw_ticket_async", "mm_draw_collect", "mm_mesh_batch_noalloc",
)


class MMError(RuntimeError):
    def __init__(self, code: int, detail: str = ""):
        self.code = code
        super().__init__(f"{STATUS_NAMES.get(code, code)}: {detail}")


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C mimas_b200/csrc` "
                          "(or __graft_entry__.build()); there is no fallback implementation")
    l = C.CDLL(LIB_PATH)
    vp, u32, u64, i64 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int64
    l.mm_abi_version.restype = u32
    l.mm_last_error_string.restype = C.c_char_p
    l.mm_last_error_string.argtypes = [vp]
    l.mm_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
    l.mm_ctx_destroy.argtypes = [vp]
    l.mm_ctx_set_stream.argtypes = [vp, vp]
    l.mm_set_block_table.argtypes = [vp, vp]
    l.mm_table_from_cache.argtypes = [vp, vp, u32, vp]
    l.mm_reserve.argtypes = [vp, u32, u64, u64]
    l.mm_mesh_batch.argtypes = [vp, vp, vp, vp, u32, vp]
    l.mm_mesh_batch_async.argtypes = [vp, vp, vp, u32]
    l.mm_mesh_batch_halo_async.argtypes = [vp, vp, vp, vp, u32]
    l.mm_gather_halos.argtypes = [vp, vp, u32, u32, u32, C.c_uint8, vp]
    l.mm_set_textures.argtypes = [vp, vp, u32, u32]
    l.mm_raster_frame.argtypes = [vp, vp, u64, vp, vp, vp]
    l.mm_batch_finish.argtypes = [vp, vp]
    l.mm_copy_out.argtypes = [vp, vp, vp, vp, vp]
    l.mm_mesh_batch_host.argtypes = [vp, vp, vp, u32, vp]
    l.mm_mesh_for_chunk.argtypes = [vp, vp, vp, C.POINTER(vp), C.POINTER(u64), C.POINTER(vp), C.POINTER(u64)]
    l.mm_push_blocks.argtypes = [vp, vp, u32, vp, C.POINTER(u64)]
    l.mm_gen_terrain_host.argtypes = [u64, i64, i64, i64, u32, u32, u32, vp, vp]
    l.mm_gen_terrain_device.argtypes = [vp, u64, i64, i64, i64, u32, u32, u32, vp, vp]
    l.mm_dev_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    l.mm_dev_free.argtypes = [vp, vp]
    l.mm_dev_upload.argtypes = [vp, vp, vp, C.c_size_t]
    l.mm_dev_download.argtypes = [vp, vp, vp, C.c_size_t]
    l.mm_ctx_sync.argtypes = [vp]
    l.mm_mesh_batch_noalloc.argtypes = [vp, vp, vp, u32, vp, u64, vp, u64, vp, vp, vp]
    l.mm_host_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    l.mm_host_free.argtypes = [vp, vp]
    l.mm_launch_count.argtypes = [vp, C.POINTER(u64)]
    l.mm_device_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    l.mm_cull_chunks.argtypes = [vp, vp, u32, vp, vp, C.POINTER(u32)]
    l.mm_debug_trace.argtypes = [vp, u32, vp, C.POINTER(u64)]
    l.mm_vertex_stage.argtypes = [vp, vp, u64, vp, vp]
    l.mm_draw_batch.argtypes = [vp, vp, vp]
    l.mm_ring_configure.argtypes = [vp, u32]
    l.mm_mesh_batch_submit.argtypes = [vp, vp, vp, vp, u32, C.POINTER(u64)]
    l.mm_mesh_batches_submit.argtypes = [vp, vp, vp, vp, u32, u32, C.POINTER(u64)]
    l.mm_batch_wait.argtypes = [vp, u64, vp]
    l.mm_batch_poll.argtypes = [vp, u64, C.POINTER(C.c_int)]
    l.mm_copy_out_ticket.argtypes = [vp, u64, vp, vp, vp, vp]
    l.mm_draw_ticket.argtypes = [vp, u64, vp, vp]
    l.mm_draw_ticket_async.argtypes = [vp, u64, vp]
    l.mm_draw_collect.argtypes = [vp, vp]
    l.mm_gen_mapgen_device.argtypes = [vp, u64, vp, i64, i64, i64, u32, u32, u32, vp, vp, vp]
    l.mm_abi_version.argtypes = []
    for s in SYMBOLS:
        f = getattr(l, s)
        # a symbol without argtypes would pass Python ints as 32-bit C ints and truncate every pointer
        if f.argtypes is None:
            raise ImportError(f"_lib.py declares no argument types for {s}")
        if s not in ("mm_abi_version", "mm_last_error_string"):
            f.restype = C.c_int
    if l.mm_abi_version() != 1:
        raise ImportError("libmimas_meshgen_cuda.so: ABI version mismatch")
    _lib = l
    return l


def check(rc: int, ctx) -> None:
    if rc != 0:
        detail = ""
        if ctx is not None and getattr(ctx, "value", None):
            detail = lib().mm_last_error_string(ctx).decode(errors="replace")
        raise MMError(rc, detail)
