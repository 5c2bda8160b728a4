"""Shared helpers for the test-suite (fixtures, byte comparison)."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

VERTEX_DTYPE = np.dtype([("tex_ind", "<u2"), ("pad0", "<u2"), ("tex_pos", "<f4", (2,)),
                         ("position", "<f4", (3,)), ("normal", "<f4", (3,))])


class Golden:
    def __init__(self):
        self.table = json.load(open(os.path.join(GOLDEN, "default_block_table.json")))
        z = np.load(os.path.join(GOLDEN, "kats.npz"))
        self.names = [str(n) for n in z["names"]]
        self.rows = [tuple(int(v) for v in r) for r in z["rows"]]
        self.cases = {n: dict(chunk=z[n + "/chunk"], offs=z[n + "/offs"],
                              intr=z[n + "/intr"].tobytes(), trans=z[n + "/trans"].tobytes())
                      for n in self.names}

    @property
    def n_known(self):
        return self.table["n_ids"]

    @property
    def fallback(self):
        return tuple(self.table["fallback"])

    @property
    def entries(self):
        return [tuple(r) for r in self.table["entries"]]


def load_golden():
    return Golden()


def first_diff(a: bytes, b: bytes):
    """Human-readable description of the first differing vertex of two vertex streams."""
    if len(a) != len(b):
        return f"length {len(a)} != {len(b)} bytes ({len(a) // 36} vs {len(b) // 36} vertices)"
    va = np.frombuffer(a, dtype=np.uint8).reshape(-1, 36)
    vb = np.frombuffer(b, dtype=np.uint8).reshape(-1, 36)
    bad = np.nonzero((va != vb).any(axis=1))[0]
    if len(bad) == 0:
        return None
    i = int(bad[0])
    sa = np.frombuffer(va[i].tobytes(), dtype=VERTEX_DTYPE)[0]
    sb = np.frombuffer(vb[i].tobytes(), dtype=VERTEX_DTYPE)[0]
    return f"{len(bad)} vertices differ; first at {i} (quad {i // 6}, corner {i % 6}): {sa} != {sb}"
