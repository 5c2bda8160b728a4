"""Headless consumer of the vertex stream: the camera of mimas-client/client.rs:1382-1400
and ctypes views of mm_camera / mm_shaded_vertex / mm_draw_out.  GL is unavailable on this box: culling
(client.rs:679-696) and the vertex shader (vertex-shader.glsl:18-27) run as CUDA kernels
(csrc/mm_render.cu) and leave a shaded vertex buffer in device memory."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

SHADED_DTYPE = np.dtype([("clip", "<f4", (4,)), ("vposition", "<f4", (4,)), ("vlamb", "<f4"),
                         ("vtex_pos", "<f4", (2,)), ("vtex_ind", "<u4")])
assert SHADED_DTYPE.itemsize == 48


class CCamera(C.Structure):
    _fields_ = [("vmatrix", C.c_float * 16), ("pmatrix", C.c_float * 16), ("pos", C.c_float * 3),
                ("viewing_range", C.c_float)]


class DrawOut(C.Structure):
    _fields_ = [("d_shaded", C.c_void_p), ("d_draw_list", C.c_void_p), ("n_shaded_i", C.c_uint64),
                ("n_shaded_t", C.c_uint64), ("n_drawn", C.c_uint32), ("n_chunks", C.c_uint32),
                ("kernel_ms", C.c_float), ("reserved", C.c_uint32)]


class CFrame(C.Structure):  # mm_frame
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("fog_near", C.c_float), ("fog_far", C.c_float),
                ("clear", C.c_float * 4)]


def synthetic_textures(layers: int, size: int = 16, seed: int = 7, crossed=()) -> np.ndarray:
    """A stand-in texture array [layers, size, size, 4] (the game's PNGs are not shipped here): per layer a
    base colour with per-texel variation; layers in `crossed` get alpha holes like the plant textures
    (fragment-shader.glsl:16-18 discards alpha < 0.5)."""
    rng = np.random.RandomState(seed)
    base = rng.rand(layers, 1, 1, 3).astype(np.float32) * 0.7 + 0.2
    var = (rng.rand(layers, size, size, 1).astype(np.float32) - 0.5) * 0.25
    tex = np.ones((layers, size, size, 4), dtype=np.float32)
    tex[..., :3] = np.clip(base + var, 0.0, 1.0)
    for l in crossed:
        if l < layers:
            tex[l, :, :, 3] = (rng.rand(size, size) < 0.45).astype(np.float32)
    return tex


def write_png(path: str, img: np.ndarray):
    """uint32 [height, width] image from raster_frame (row 0 at the bottom) -> 8-bit RGB PNG, top row first
    (stdlib only: zlib + the three mandatory chunks)."""
    import struct
    import zlib
    rgb = np.stack([(img >> s) & 0xFF for s in (0, 8, 16)], axis=-1).astype(np.uint8)[::-1]
    h, w = img.shape
    raw = np.concatenate([np.zeros((h, 1), np.uint8), rgb.reshape(h, w * 3)], axis=1).tobytes()  # filter byte 0 per row

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def _f32(x):
    return np.float32(x)


class Camera:
    """client.rs:1300-1400: position, pitch / yaw in degrees, aspect ratio; `viewing_range` is
    config.viewing_range (mimas-common/config.rs).  The two matrices are nalgebra's `look_at_rh` and
    `new_perspective` (third-party arithmetic, restated here in f32 from their published formulas; the
    kernels take the matrices as given, so any host may supply its own)."""

    def __init__(self, pos=(60.0, 40.0, 20.0), pitch=0.0, yaw=0.0, aspect_ratio=1024.0 / 768.0, viewing_range=128.0):
        self.pos = np.array(pos, dtype=np.float32)
        self.pitch, self.yaw = float(pitch), float(yaw)
        self.aspect_ratio = float(aspect_ratio)
        self.viewing_range = float(viewing_range)

    def direction(self) -> np.ndarray:  # client.rs:1382-1386
        pitch, yaw = _f32(math.radians(-self.pitch)), _f32(math.radians(-self.yaw))
        return np.array([np.cos(pitch) * np.cos(yaw), np.cos(pitch) * np.sin(yaw), np.sin(pitch)], dtype=np.float32)

    def get_matrix(self) -> np.ndarray:
        """Matrix4::look_at_rh(eye = pos, target = pos + direction, up = z), client.rs:1388-1393; [col][row]."""
        eye = self.pos
        f = self.direction()
        f = f / np.float32(np.linalg.norm(f))
        up = np.array([0, 0, 1], dtype=np.float32)
        s = np.cross(f, up).astype(np.float32)
        s = s / np.float32(np.linalg.norm(s))
        u = np.cross(s, f).astype(np.float32)
        m = np.zeros((4, 4), dtype=np.float32)  # row-major here, transposed on return
        m[0, :3], m[1, :3], m[2, :3] = s, u, -f
        m[0, 3], m[1, 3], m[2, 3] = -np.dot(s, eye), -np.dot(u, eye), np.dot(f, eye)
        m[3, 3] = 1
        return np.ascontiguousarray(m.T)

    def get_perspective(self) -> np.ndarray:
        """Matrix4::new_perspective(aspect, fov = 90 deg, znear = 0.1, zfar = 1024), client.rs:1395-1400."""
        fov, zfar, znear = _f32(math.radians(90.0)), _f32(1024.0), _f32(0.1)
        t = _f32(1.0) / np.tan(fov / _f32(2.0))
        m = np.zeros((4, 4), dtype=np.float32)
        m[0, 0] = t / _f32(self.aspect_ratio)
        m[1, 1] = t
        m[2, 2] = (zfar + znear) / (znear - zfar)
        m[2, 3] = _f32(2.0) * zfar * znear / (znear - zfar)
        m[3, 2] = -1
        return np.ascontiguousarray(m.T)

    def to_c(self) -> CCamera:
        """mm_camera of this view.  Pass the result itself to the draw calls of a hot loop: building the two
        matrices in numpy costs far more than a frame's enqueue."""
        c = CCamera()
        c.vmatrix[:] = [float(v) for v in self.get_matrix().reshape(-1)]
        c.pmatrix[:] = [float(v) for v in self.get_perspective().reshape(-1)]
        c.pos[:] = [float(v) for v in self.pos]
        c.viewing_range = self.viewing_range
        return c
