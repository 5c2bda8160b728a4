"""TEST INFRASTRUCTURE -- ctypes front-end of oracle/mapgen_oracle.c (restatement of mimas-server/mapgen.rs:64-245).
PARITY UNPINNED (see the header of the C file).  Never imported by the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


class Pcg32(C.Structure):
    _fields_ = [("state", C.c_uint64), ("increment", C.c_uint64)]


def load():
    global _lib
    if _lib is None:
        subprocess.run(["make", "-C", HERE, "-s"], check=True)
        l = C.CDLL(os.path.join(HERE, "libmapgen_oracle.so"))
        l.mg_pcg32_next_u32.restype = C.c_uint32
        l.mg_pcg32_next_u64.restype = C.c_uint64
        l.mg_pcg32_gen_f64.restype = C.c_double
        l.mg_xxh64.restype = C.c_uint64
        l.mg_xxh64.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64]
        l.mg_pos_hash.restype = C.c_uint64
        l.mg_perlin2.restype = C.c_double
        l.mg_perlin2.argtypes = [C.c_void_p, C.c_double, C.c_double]
        l.mg_perlin3.restype = C.c_double
        l.mg_perlin3.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double]
        l.mg_seed_u32.restype = C.c_uint32
        l.mg_seed_u32.argtypes = [C.c_uint64, C.c_char_p]
        l.mg_seed_u64.restype = C.c_uint64
        l.mg_seed_u64.argtypes = [C.c_uint64, C.c_char_p]
        l.mg_sizeof_params.restype = C.c_uint32
        l.mg_gen_region.argtypes = [C.c_uint64, C.c_int64, C.c_int64, C.c_int64, C.c_uint32, C.c_uint32, C.c_uint32,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = l
    return _lib


def pcg32(state: int, stream: int) -> Pcg32:
    p = Pcg32()
    load().mg_pcg32_new(C.byref(p), C.c_uint64(state), C.c_uint64(stream))
    return p


def pos_hash(pos) -> int:
    a = np.asarray(pos, dtype=np.int64)
    return load().mg_pos_hash(a.ctypes.data_as(C.c_void_p))


def perm_table(seed: int) -> np.ndarray:
    v = np.zeros(256, dtype=np.uint8)
    load().mg_perm_table(C.c_uint32(seed), v.ctypes.data_as(C.c_void_p))
    return v


def perlin2(table: np.ndarray, x: float, y: float) -> float:
    return load().mg_perlin2(table.ctypes.data, x, y)


def perlin3(table: np.ndarray, x: float, y: float, z: float) -> float:
    return load().mg_perlin3(table.ctypes.data, x, y, z)


def gen_region(seed: int, c0, dims, c_params):
    """c_params: a ctypes struct with the layout of mm_mapgen_params (mimas_b200.mapgen.to_c builds one; the oracle's
    mg_params has the same layout).  Returns (chunks [n,4096], offs [n,3], tree_counts [n])."""
    l = load()
    assert l.mg_sizeof_params() == C.sizeof(c_params)
    n = int(dims[0]) * int(dims[1]) * int(dims[2])
    chunks = np.zeros((n, 4096), dtype=np.uint8)
    offs = np.zeros((n, 3), dtype=np.int64)
    trees = np.zeros(n, dtype=np.uint32)
    l.mg_gen_region(seed, c0[0], c0[1], c0[2], dims[0], dims[1], dims[2], C.cast(C.byref(c_params), C.c_void_p),
                    chunks.ctypes.data, offs.ctypes.data, trees.ctypes.data)
    return chunks, offs, trees
