"""The C++ host mirror of the crate API (include/mimas_meshgen.hpp): TextureIdCache::from_hdl, HeadlessAssets,
mesh_for_chunk, mesh_batch, push_block -- through tests/cxx/host_mirror_test.  The reference's host language is
Rust, which this image cannot build; C++ is the compiled stand-in above the same C ABI."""
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle
from tests.util import load_golden, oracle_cache

HERE = os.path.dirname(os.path.abspath(__file__))
EXE = os.path.join(HERE, "cxx", "host_mirror_test")
G = load_golden()


@pytest.fixture(scope="module")
def exe():
    root = os.path.dirname(HERE)
    subprocess.run(["make", "-C", os.path.join(root, "mimas_b200", "csrc"), "-s"], check=True, capture_output=True)
    subprocess.run(["make", "-C", os.path.join(HERE, "cxx"), "-s"], check=True)
    return EXE


def test_cxx_host_logic_and_loud_failure_without_device(exe):
    """Id allocation order (lib.rs:87-104, assets.rs:145-201), table flattening, and: no CPU fallback."""
    r = subprocess.run([exe, "selfcheck"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "selfcheck ok" in r.stdout


@pytest.mark.gpu
def test_cxx_mesh_for_chunk_matches_vectors_and_oracle(exe, tmp_path):
    import mimas_b200
    rng = np.random.RandomState(17)
    chunks = [G.cases[n]["chunk"] for n in G.names]
    offs = [G.cases[n]["offs"] for n in G.names]
    want_i = [G.cases[n]["intr"] for n in G.names]
    want_t = [G.cases[n]["trans"] for n in G.names]
    tc, to = mimas_b200.gen_terrain_host(78, (-2, -2, -1), (4, 4, 2))
    rc = rng.randint(0, 256, size=(6, 4096)).astype(np.uint8)
    ro = rng.randint(-2 ** 40, 2 ** 40, size=(6, 3)).astype(np.int64)
    ec, eo = np.concatenate([tc, rc]), np.concatenate([to, ro])
    w = oracle.mesh_batch(ec, eo, oracle_cache(G), threads=4)
    vi, vt = w["intransparent"].tobytes(), w["transparent"].tobytes()
    for k in range(len(ec)):
        chunks.append(ec[k])
        offs.append(eo[k])
        want_i.append(vi[36 * int(w["first_i"][k]):36 * (int(w["first_i"][k]) + int(w["count_i"][k]))])
        want_t.append(vt[36 * int(w["first_t"][k]):36 * (int(w["first_t"][k]) + int(w["count_t"][k]))])
    path = tmp_path / "cases.bin"
    with open(path, "wb") as f:
        rows = G.entries[:G.n_known]
        f.write(struct.pack("<II4H", 0x54434D4D, len(rows), *G.fallback, 0))
        for r in rows:
            f.write(struct.pack("<5H", *r))
        f.write(struct.pack("<I", len(chunks)))
        for c, o, a, b in zip(chunks, offs, want_i, want_t):
            f.write(np.asarray(o, dtype=np.int64).tobytes())
            f.write(np.asarray(c, dtype=np.uint8).tobytes())
            f.write(struct.pack("<QQ", len(a) // 36, len(b) // 36))
            f.write(a)
            f.write(b)
    r = subprocess.run([exe, "cases", str(path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "cases ok" in r.stdout
