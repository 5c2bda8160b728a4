"""Known-answer vectors for the chunk mesher, derived INDEPENDENTLY of oracle/.

The reference ships no golden vectors for mimas-meshgen, so these are
derived by hand-style reasoning from the macros, written down as a table of quad corners
per face direction (SURVEY.md §8a "pass table") and a run finder built on
itertools.groupby -- a different formulation from the Walker state machine that the
oracles restate.  Does not import oracle/ and does not read /root/reference.

What each pass emits for a run with start a0 (world coord, f32), length L = f32(end)-f32(a0)
and row world coords (x, y, z) -- derived from rpush_face!/rpush_face_rev!
(mimas-meshgen/lib.rs:139-160) with the arguments of lib.rs:326,342,358,374,390,406:

  pass  P0            P1              P2              P4                uvP1  uvP2  normal      order
  1 z-  (x,a0,z)      (x+1,a0,z)      (x,a0+L,z)      (x+1,a0+L,z)      (1,0) (0,L) ( 1, 1, 0)  fwd
  2 y-  (a0,y,z)      (a0+L,y,z)      (a0,y,z+1)      (a0+L,y,z+1)      (L,0) (0,1) (-1, 0,-1)  rev
  3 x-  (x,a0,z)      (x,a0+L,z)      (x,a0,z+1)      (x,a0+L,z+1)      (L,0) (0,1) ( 0, 1, 1)  fwd
  4 z+  (x,a0,z+1)    (x+1,a0,z+1)    (x,a0+L,z+1)    (x+1,a0+L,z+1)    (1,0) (0,L) (-1,-1, 0)  rev
  5 y+  (a0,y+1,z)    (a0+L,y+1,z)    (a0,y+1,z+1)    (a0+L,y+1,z+1)    (L,0) (0,1) ( 1, 0, 1)  fwd
  6 x+  (x+1,a0,z)    (x+1,a0+L,z)    (x+1,a0,z+1)    (x+1,a0+L,z+1)    (L,0) (0,1) ( 0,-1,-1)  rev
  fwd = P0,P1,P2,P1,P4,P2   rev = P2,P1,P0,P2,P4,P1   uvP0 = (0,0), uvP4 = (uvP1.u, uvP2.v)
  (normal components are sign(L) where the table shows +-1 next to an L-extent; L == 0
   only happens beyond 2^24 and is handled below.)

    python tests/golden/make_kats.py        # rewrites tests/golden/kats.npz
"""
import itertools
import json
import os

import numpy as np

f32 = np.float32
HERE = os.path.dirname(os.path.abspath(__file__))
VERTEX_DTYPE = np.dtype([("tex_ind", "<u2"), ("pad0", "<u2"), ("tex_pos", "<f4", (2,)),
                         ("position", "<f4", (3,)), ("normal", "<f4", (3,))])

TABLE = json.load(open(os.path.join(HERE, "default_block_table.json")))


def table_rows(n_known=None):
    rows = [tuple(r) for r in TABLE["entries"]]
    fb = tuple(TABLE["fallback"])
    full = rows + [(1, fb[0], fb[1], fb[2], 0)] * (256 - len(rows))  # unknown id -> blocky fallback
    return full


def sgn(v):
    return f32(1.0) if v > 0 else f32(-1.0) if v < 0 else f32(0.0)


def quad(p, corners, uvs, normal, tex, rev):
    order = [2, 1, 0, 2, 3, 1] if rev else [0, 1, 2, 1, 3, 2]
    out = np.zeros(6, dtype=VERTEX_DTYPE)
    for i, k in enumerate(order):
        out[i]["tex_ind"] = tex
        out[i]["tex_pos"] = uvs[k]
        out[i]["position"] = corners[k]
        out[i]["normal"] = normal
    return out


def mesh(chunk, offs, rows):
    """Independent mesher: per pass, per row, group consecutive equal keys."""
    g = chunk.reshape(16, 16, 16)  # [x][y][z], index x*256+y*16+z
    blocky = np.array([r[0] == 1 for r in rows])
    tex = {"s": np.array([r[1] for r in rows]), "t": np.array([r[2] for r in rows]),
           "b": np.array([r[3] for r in rows])}
    B = blocky[g]
    ox, oy, oz = (int(v) for v in offs)
    verts = []
    # (inner axis, face key, occluder offset, rev)
    passes = [("y", "b", (0, 0, -1), False), ("x", "s", (0, -1, 0), True), ("y", "s", (-1, 0, 0), False),
              ("y", "t", (0, 0, 1), True), ("x", "s", (0, 1, 0), False), ("y", "s", (1, 0, 0), True)]
    for pi, (inner, face, (dx, dy, dz), rev) in enumerate(passes, start=1):
        # occluded[x,y,z] = neighbour inside the chunk and blocky
        occ = np.zeros_like(B)
        src = B[max(dx, 0):16 + min(dx, 0), max(dy, 0):16 + min(dy, 0), max(dz, 0):16 + min(dz, 0)]
        occ[max(-dx, 0):16 + min(-dx, 0), max(-dy, 0):16 + min(-dy, 0), max(-dz, 0):16 + min(-dz, 0)] = src
        vis = B & ~occ
        key = np.where(vis, tex[face][g], -1)
        for c1 in range(16):
            for c2 in range(16):
                row = key[c1, :, c2] if inner == "y" else key[:, c1, c2]
                pos = 0
                for k, grp in itertools.groupby(row.tolist()):
                    n = len(list(grp))
                    if k >= 0:
                        if inner == "y":
                            x, z = f32(ox + c1), f32(oz + c2)
                            a0, a1 = f32(oy + pos), f32(oy + pos + n)
                        else:
                            y, z = f32(oy + c1), f32(oz + c2)
                            a0, a1 = f32(ox + pos), f32(ox + pos + n)
                        L = a1 - a0
                        one, zero = f32(1), f32(0)
                        if pi == 1:
                            c = [(x, a0, z), (x + one, a0, z), (x, a0 + L, z), (x + one, a0 + L, z)]
                            uv = [(0, 0), (1, 0), (0, L), (1, L)]; nrm = (1, sgn(L), 0)
                        elif pi == 2:
                            c = [(a0, y, z), (a0 + L, y, z), (a0, y, z + one), (a0 + L, y, z + one)]
                            uv = [(0, 0), (L, 0), (0, 1), (L, 1)]; nrm = (sgn(-L), 0, -1)
                        elif pi == 3:
                            c = [(x, a0, z), (x, a0 + L, z), (x, a0, z + one), (x, a0 + L, z + one)]
                            uv = [(0, 0), (L, 0), (0, 1), (L, 1)]; nrm = (0, sgn(L), 1)
                        elif pi == 4:
                            zz = z + one
                            c = [(x, a0, zz), (x + one, a0, zz), (x, a0 + L, zz), (x + one, a0 + L, zz)]
                            uv = [(0, 0), (1, 0), (0, L), (1, L)]; nrm = (-1, sgn(-L), 0)
                        elif pi == 5:
                            yy = y + one
                            c = [(a0, yy, z), (a0 + L, yy, z), (a0, yy, z + one), (a0 + L, yy, z + one)]
                            uv = [(0, 0), (L, 0), (0, 1), (L, 1)]; nrm = (sgn(L), 0, 1)
                        else:
                            xx = x + one
                            c = [(xx, a0, z), (xx, a0 + L, z), (xx, a0, z + one), (xx, a0 + L, z + one)]
                            uv = [(0, 0), (L, 0), (0, 1), (L, 1)]; nrm = (0, sgn(-L), -1)
                        verts.append(quad(None, c, uv, nrm, k, rev))
                    pos += n
    intr = np.concatenate(verts) if verts else np.zeros(0, dtype=VERTEX_DTYPE)

    # crossed plants, lib.rs:414-440: linear index order, 4 quads (2 bidi) per plant
    SQ = np.frombuffer(np.uint32(0x3F3504F3).tobytes(), dtype=np.float32)[0]
    SQH = np.frombuffer(np.uint32(0x3EB504F3).tobytes(), dtype=np.float32)[0]
    H = np.frombuffer(np.uint32(0x3F733333).tobytes(), dtype=np.float32)[0]
    tv = []
    for idx in range(4096):
        r = rows[int(chunk[idx])]
        if r[0] != 2:
            continue
        x, y, z = idx >> 8, (idx >> 4) & 15, idx & 15
        X, Y, Zc = f32(ox + x) + f32(0.5), f32(oy + y) + f32(0.5), f32(oz + z)
        for bx, dxs in ((X - SQH, SQ), (X + SQH, -SQ)):
            by = Y - SQH
            c = [(bx, by, Zc), (bx + dxs, by + SQ, Zc), (bx, by, Zc + H), (bx + dxs, by + SQ, Zc + H)]
            uv = [(0, 0), (1, 0), (0, H), (1, H)]
            s = f32(1) if dxs > 0 else f32(-1)
            tv.append(quad(None, c, uv, (s, 1, 1), r[4], False))
            tv.append(quad(None, c, uv, (-s, -1, -1), r[4], True))
    trans = np.concatenate(tv) if tv else np.zeros(0, dtype=VERTEX_DTYPE)
    return intr, trans


def cases():
    AIR, WATER, SAND, GROUND, STONE, GRASSY, SAPLING, ROSE = 0, 1, 2, 3, 5, 11, 42, 51
    out = {}

    def chunk():
        return np.zeros(4096, dtype=np.uint8)

    def put(c, x, y, z, v):
        c[x * 256 + y * 16 + z] = v

    out["all_air"] = (chunk(), (0, 0, 0))
    c = chunk(); put(c, 0, 0, 0, STONE); out["single_stone_origin"] = (c, (0, 0, 0))
    c = chunk(); put(c, 7, 8, 9, STONE); out["single_stone_inner_offs"] = (c, (48, 32, 16))
    out["uniform_stone"] = (np.full(4096, STONE, dtype=np.uint8), (0, 0, 0))
    c = chunk(); put(c, 3, 4, 5, STONE); put(c, 3, 5, 5, SAND); out["two_diff_along_y"] = (c, (16, -32, 0))
    c = chunk(); put(c, 3, 4, 5, STONE); put(c, 3, 5, 5, STONE); out["two_same_along_y"] = (c, (16, -32, 0))
    c = chunk(); put(c, 3, 4, 5, STONE); put(c, 4, 4, 5, STONE); put(c, 5, 4, 5, SAND); out["three_along_x"] = (c, (0, 0, 0))
    idx = np.arange(4096); x, y, z = idx >> 8, (idx >> 4) & 15, idx & 15
    out["checkerboard_even"] = (np.where(((x + y + z) & 1) == 1, STONE, AIR).astype(np.uint8), (0, 0, 0))
    out["checkerboard_odd"] = (np.where(((x + y + z + 1) & 1) == 1, STONE, AIR).astype(np.uint8), (-16, 16, 32))
    c = chunk(); put(c, 1, 1, 1, GRASSY); out["ground_with_grass"] = (c, (0, 0, 0))
    c = chunk(); put(c, 2, 3, 4, SAPLING); out["one_sapling"] = (c, (0, 0, 0))
    c = chunk(); put(c, 2, 3, 4, SAPLING); put(c, 2, 3, 3, GROUND); put(c, 2, 4, 4, STONE); put(c, 15, 15, 15, ROSE)
    out["sapling_does_not_occlude"] = (c, (-48, 160, -16))
    c = chunk(); put(c, 0, 0, 0, 200); put(c, 0, 1, 0, 201); out["unknown_ids_fallback"] = (c, (0, 0, 0))
    out["negative_offs_uniform"] = (np.full(4096, STONE, dtype=np.uint8), (-16, -16, -16))
    c = chunk(); put(c, 0, 0, 0, STONE); put(c, 0, 1, 0, STONE); put(c, 0, 15, 0, WATER)
    out["huge_offs_2p24"] = (c, (1 << 24, (1 << 24) + 16, -(1 << 25)))
    c = np.full(4096, STONE, dtype=np.uint8); out["huge_offs_2p30_uniform"] = (c, (1 << 30, -(1 << 30), 1 << 28))
    # layered terrain-like column stack with an ore vein and a plant on top
    c = chunk().reshape(16, 16, 16)
    c[:, :, :6] = STONE; c[:, :, 6:9] = GROUND; c[:, :, 9] = GRASSY; c[4:7, 5, 2:4] = 9; c[8, 8, 10] = 45
    c[0:3, 0:3, 9:12] = WATER
    out["layered_terrain"] = (c.reshape(4096).copy(), (48, 32, 0))
    rng = np.random.RandomState(1234)
    out["random_dense_ids"] = (rng.choice([0, 1, 2, 3, 5, 9, 11, 42, 48, 17, 200], size=4096).astype(np.uint8), (32, -48, 64))
    return out


def main():
    rows = table_rows()
    blob = {}
    names = []
    for name, (chunk, offs) in cases().items():
        intr, trans = mesh(chunk, offs, rows)
        names.append(name)
        blob[name + "/chunk"] = chunk
        blob[name + "/offs"] = np.array(offs, dtype=np.int64)
        blob[name + "/intr"] = np.frombuffer(intr.tobytes(), dtype=np.uint8)
        blob[name + "/trans"] = np.frombuffer(trans.tobytes(), dtype=np.uint8)
        print(f"{name:28s} intr {len(intr):6d} verts  trans {len(trans):5d} verts")
    blob["names"] = np.array(names)
    blob["rows"] = np.array(rows, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "kats.npz"), **blob)


if __name__ == "__main__":
    main()
