"""TEST INFRASTRUCTURE -- second, independent restatement of mimas-meshgen/lib.rs in pure
Python (numpy.float32 scalars for every arithmetic step; slow, small cases only).

Purpose: no Rust toolchain exists here to arbitrate, so the C oracle (meshgen_oracle.c)
is cross-checked bit-for-bit against this transliteration, which keeps the reference's
own shape: closures, a generic Walker, and macros written as functions.
PARITY UNPINNED (see meshgen_oracle.c).  Never imported by the product package.
"""
from __future__ import annotations

import numpy as np

f32 = np.float32
CHUNKSIZE = 16  # mimas-common/map.rs:11
SQRT_2 = f32(1.41421356237309504880168872420969808)  # std::f32::consts::SQRT_2


def zsig(v):  # lib.rs:123-131
    if v > 0.0:
        return f32(1.0)
    if v < 0.0:
        return f32(-1.0)
    return f32(0.0)


def sign(x, y, z):  # lib.rs:133-137
    return [zsig(x), zsig(y), zsig(z)]


def V(tex_ind, tex_pos, position, normal):  # lib.rs:18-24
    return (int(tex_ind), [f32(a) for a in tex_pos], [f32(a) for a in position], [f32(a) for a in normal])


Z = f32(0.0)


def rpush_face(r, p, d, tex_ind):  # lib.rs:139-149
    (x, y, z), (xsd, ysd, yd, zd) = p, d
    r.append(V(tex_ind, [Z, Z], [x, y, z], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [xsd + ysd, Z], [x + xsd, y + ysd, z], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [Z, yd + zd], [x, y + yd, z + zd], sign(xsd, ysd + yd, zd)))

    r.append(V(tex_ind, [xsd + ysd, Z], [x + xsd, y + ysd, z], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [xsd + ysd, yd + zd], [x + xsd, y + yd + ysd, z + zd], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [Z, yd + zd], [x, y + yd, z + zd], sign(xsd, ysd + yd, zd)))


def rpush_face_rev(r, p, d, tex_ind):  # lib.rs:150-160
    (x, y, z), (xsd, ysd, yd, zd) = p, d
    r.append(V(tex_ind, [Z, yd + zd], [x, y + yd, z + zd], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [xsd + ysd, Z], [x + xsd, y + ysd, z], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [Z, Z], [x, y, z], sign(-xsd, -ysd - yd, -zd)))

    r.append(V(tex_ind, [Z, yd + zd], [x, y + yd, z + zd], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [xsd + ysd, yd + zd], [x + xsd, y + yd + ysd, z + zd], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [xsd + ysd, Z], [x + xsd, y + ysd, z], sign(-xsd, -ysd - yd, -zd)))


def rpush_face_bidi(r, p, d, td, tex_ind):  # lib.rs:162-181
    (x, y, z), (xsd, ysd, yd, zd), (xstd, ystd, ytd, ztd) = p, d, td
    r.append(V(tex_ind, [Z, Z], [x, y, z], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [xstd + ystd, Z], [x + xsd, y + ysd, z], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [Z, ytd + ztd], [x, y + yd, z + zd], sign(xsd, ysd + yd, zd)))

    r.append(V(tex_ind, [xstd + ystd, Z], [x + xsd, y + ysd, z], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [xstd + ystd, ytd + ztd], [x + xsd, y + yd + ysd, z + zd], sign(xsd, ysd + yd, zd)))
    r.append(V(tex_ind, [Z, ytd + ztd], [x, y + yd, z + zd], sign(xsd, ysd + yd, zd)))
    # Reverse face
    r.append(V(tex_ind, [Z, ytd + ztd], [x, y + yd, z + zd], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [xstd + ystd, Z], [x + xsd, y + ysd, z], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [Z, Z], [x, y, z], sign(-xsd, -ysd - yd, -zd)))

    r.append(V(tex_ind, [Z, ytd + ztd], [x, y + yd, z + zd], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [xstd + ystd, ytd + ztd], [x + xsd, y + yd + ysd, z + zd], sign(-xsd, -ysd - yd, -zd)))
    r.append(V(tex_ind, [xstd + ystd, Z], [x + xsd, y + ysd, z], sign(-xsd, -ysd - yd, -zd)))


def push_block(r, pos, block_ids, siz, blocked):  # lib.rs:183-219
    x, y, z = (f32(a) for a in pos)
    siz = f32(siz)
    s, t, b = block_ids
    if not blocked([0, 0, -1]):
        rpush_face(r, (x, y, z), (siz, Z, siz, Z), b)
    if not blocked([0, -1, 0]):
        rpush_face_rev(r, (x, y, z), (siz, Z, Z, siz), s)
    if not blocked([-1, 0, 0]):
        rpush_face(r, (x, y, z), (Z, siz, Z, siz), s)
    if not blocked([0, 0, 1]):
        rpush_face_rev(r, (x, y, z + siz), (siz, Z, siz, Z), t)
    if not blocked([0, 1, 0]):
        rpush_face(r, (x, y + siz, z), (siz, Z, Z, siz), s)
    if not blocked([1, 0, 0]):
        rpush_face_rev(r, (x + siz, y, z), (Z, siz, Z, siz), s)


class Walker:  # lib.rs:235-272
    def __init__(self):
        self.last = None

    def next(self, v, item, emit):
        if item is None and self.last is not None:
            last_v, l_item = self.last
            vlen = v - last_v
            emit(l_item, last_v, vlen)
            self.last = None
        elif item is not None and self.last is not None:
            last_v, l_item = self.last
            if item != l_item:
                vlen = v - last_v
                emit(l_item, last_v, vlen)
                self.last = (v, item)
        elif item is not None and self.last is None:
            self.last = (v, item)


class PyCache:
    """TextureIdCache (lib.rs:73-118) over the flattened entry rows."""

    def __init__(self, entries, n_known, fallback):
        self.entries, self.n, self.fallback = entries, n_known, tuple(fallback)

    def get_bl_tex_ids(self, bl):  # lib.rs:105-109
        if bl < self.n:
            kind, s, t, b, _ = self.entries[bl]
            return (s, t, b) if kind == 1 else None
        return self.fallback

    def get_crossed(self, bl):  # lib.rs:110-113 (+ the `if let Crossed` at :421)
        if bl < self.n and self.entries[bl][0] == 2:
            return self.entries[bl][4]
        return None


def mesh_for_chunk(offs, chunk, cache: PyCache):  # lib.rs:231-446
    offs = [int(a) for a in offs]
    chunk = np.asarray(chunk, dtype=np.uint8).reshape(4096)
    r = []

    def get_blk(p):  # mimas-common/map.rs:83-86
        return int(chunk[p[0] * CHUNKSIZE * CHUNKSIZE + p[1] * CHUNKSIZE + p[2]])

    def blocked(o, pos):  # lib.rs:273-283
        p = [pos[0] + o[0], pos[1] + o[1], pos[2] + o[2]]
        if any(v < 0 or v >= CHUNKSIZE for v in p):
            return False
        return cache.get_bl_tex_ids(get_blk(p)) is not None

    def get_tex_ind(pos, offsets):  # lib.rs:284-293
        texture_ids = cache.get_bl_tex_ids(get_blk(pos))
        if texture_ids is not None and blocked(offsets, pos):
            return None
        return texture_ids

    def walk_for_all_blocks(f, face, offsets, g):  # lib.rs:294-313
        for c1 in range(CHUNKSIZE):
            for c2 in range(CHUNKSIZE):
                walker = Walker()
                for cinner in range(CHUNKSIZE):
                    rel_pos = f(c1, c2, cinner)
                    texture_ids = get_tex_ind(rel_pos, offsets)
                    tex_ind = face(texture_ids) if texture_ids is not None else None
                    g(walker, tex_ind, rel_pos)
                g(walker, None, f(c1, c2, CHUNKSIZE))

    siz = f32(1.0)
    add = lambda rel: [offs[0] + rel[0], offs[1] + rel[1], offs[2] + rel[2]]
    sides, top, bottom = (lambda b: b[0]), (lambda b: b[1]), (lambda b: b[2])

    def g1(walker, color, rel_pos):  # lib.rs:322-328
        pos = add(rel_pos)
        x, z = f32(pos[0]), f32(pos[2])
        walker.next(f32(pos[1]), color, lambda tx, last_y, ylen:
                    rpush_face(r, (x, last_y, z), (siz, Z, ylen, Z), tx))
    walk_for_all_blocks(lambda c1, c2, ci: [c1, ci, c2], bottom, [0, 0, -1], g1)

    def g2(walker, color, rel_pos):  # lib.rs:338-344
        pos = add(rel_pos)
        y, z = f32(pos[1]), f32(pos[2])
        walker.next(f32(pos[0]), color, lambda tx, last_x, xlen:
                    rpush_face_rev(r, (last_x, y, z), (xlen, Z, Z, siz), tx))
    walk_for_all_blocks(lambda c1, c2, ci: [ci, c1, c2], sides, [0, -1, 0], g2)

    def g3(walker, color, rel_pos):  # lib.rs:354-360
        pos = add(rel_pos)
        x, z = f32(pos[0]), f32(pos[2])
        walker.next(f32(pos[1]), color, lambda tx, last_y, ylen:
                    rpush_face(r, (x, last_y, z), (Z, ylen, Z, siz), tx))
    walk_for_all_blocks(lambda c1, c2, ci: [c1, ci, c2], sides, [-1, 0, 0], g3)

    def g4(walker, color, rel_pos):  # lib.rs:370-376
        pos = add(rel_pos)
        x, z = f32(pos[0]), f32(pos[2])
        walker.next(f32(pos[1]), color, lambda tx, last_y, ylen:
                    rpush_face_rev(r, (x, last_y, z + siz), (siz, Z, ylen, Z), tx))
    walk_for_all_blocks(lambda c1, c2, ci: [c1, ci, c2], top, [0, 0, 1], g4)

    def g5(walker, color, rel_pos):  # lib.rs:386-392
        pos = add(rel_pos)
        y, z = f32(pos[1]), f32(pos[2])
        walker.next(f32(pos[0]), color, lambda tx, last_x, xlen:
                    rpush_face(r, (last_x, y + siz, z), (xlen, Z, Z, siz), tx))
    walk_for_all_blocks(lambda c1, c2, ci: [ci, c1, c2], sides, [0, 1, 0], g5)

    def g6(walker, color, rel_pos):  # lib.rs:402-408
        pos = add(rel_pos)
        x, z = f32(pos[0]), f32(pos[2])
        walker.next(f32(pos[1]), color, lambda tx, last_y, ylen:
                    rpush_face_rev(r, (x + siz, last_y, z), (Z, ylen, Z, siz), tx))
    walk_for_all_blocks(lambda c1, c2, ci: [c1, ci, c2], sides, [1, 0, 0], g6)

    rt = []
    for x in range(CHUNKSIZE):  # lib.rs:415-440
        for y in range(CHUNKSIZE):
            for z in range(CHUNKSIZE):
                tx = cache.get_crossed(get_blk([x, y, z]))
                if tx is None:
                    continue
                pos = add([x, y, z])
                sq = siz * SQRT_2 * f32(0.5)
                sqh = sq * f32(0.5)
                fx, fy, fz = f32(pos[0]) + f32(0.5) * siz, f32(pos[1]) + f32(0.5) * siz, f32(pos[2])
                tsiz = siz * f32(0.5)
                rpush_face_bidi(rt, (fx - sqh, fy - sqh, fz), (sq, sq, Z, siz * f32(0.95)),
                                (tsiz, tsiz, Z, siz * f32(0.95)), tx)
                rpush_face_bidi(rt, (fx + sqh, fy - sqh, fz), (-sq, sq, Z, siz * f32(0.95)),
                                (tsiz, tsiz, Z, siz * f32(0.95)), tx)
    return r, rt


def to_structured(verts, dtype):
    out = np.zeros(len(verts), dtype=dtype)
    if verts:
        out["tex_ind"] = np.array([v[0] for v in verts], dtype=np.uint16)
        out["tex_pos"] = np.array([v[1] for v in verts], dtype=np.float32)
        out["position"] = np.array([v[2] for v in verts], dtype=np.float32)
        out["normal"] = np.array([v[3] for v in verts], dtype=np.float32)
    return out
