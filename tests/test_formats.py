"""World-database and wire formats around the mesher (mimas_b200/formats.py): hand-built
byte vectors from the reference's serialisers, round trips, error behaviour, and -- on the GPU -- a world file
meshed end to end against the oracle."""
import gzip
import struct

import numpy as np
import pytest

import oracle
from mimas_b200 import formats as F
from tests.util import assert_batch_equal, load_golden, oracle_cache, product_cache

G = load_golden()


def test_name_id_map_known_bytes_and_roundtrip():
    # serialize_name_id_map (map_storage.rs:204-219): version 0, u16 BE count, (u8 len, bytes)*
    data = b"\x00\x00\x02" + b"\x0bdefault:air" + b"\x0ddefault:water"
    assert F.decode_name_id_map(data) == ["default:air", "default:water"]
    assert F.encode_name_id_map(["default:air", "default:water"]) == data
    assert F.decode_name_id_map(F.encode_name_id_map(F.BUILTIN_NAMES)) == F.BUILTIN_NAMES  # ensure_name_id_roundtrip, :249-255
    assert F.decode_name_id_map(b"\x00\x00\x01\x0cdefault::air") == ["default:air"]     # '::' compatibility, :241-244
    with pytest.raises(F.FormatError):
        F.decode_name_id_map(b"\x01\x00\x00")                                            # version too recent
    with pytest.raises(F.FormatError):
        F.decode_name_id_map(b"\x00\x00\xff")                                            # 255 ids: reserved
    with pytest.raises(F.FormatError):
        F.decode_name_id_map(b"\x00\x00\x01\x05ab")                                      # truncated


def test_chunk_blob_versions_metadata_and_errors():
    rng = np.random.RandomState(1)
    ids = rng.randint(0, 11, size=4096).astype(np.uint8)
    blob = F.encode_chunk_blob(ids)
    assert blob[0] == 1 and blob[1:3] == b"\x1f\x8b"                                     # version byte, then gzip magic
    assert (F.decode_chunk_blob(blob, 11) == ids).all()
    v0 = b"\x00" + gzip.compress(ids.tobytes())                                          # version 0: ids only
    assert (F.decode_chunk_blob(v0) == ids).all()
    # version 1 with one metadata entry (pos, entry count 1, kind 0 + inventory bytes): the ids still come out
    meta = struct.pack(">H", 1) + bytes([3, 4, 5, 1, 0]) + b"\x00" * 12
    with_meta = b"\x01" + gzip.compress(ids.tobytes() + meta)
    assert (F.decode_chunk_blob(with_meta) == ids).all()
    with pytest.raises(F.FormatError, match="invalid block number"):
        F.decode_chunk_blob(blob, 5)
    with pytest.raises(F.FormatError, match="Unsupported map chunk version"):
        F.decode_chunk_blob(b"\x02" + blob[1:])
    with pytest.raises(F.FormatError):
        F.decode_chunk_blob(b"\x01" + gzip.compress(ids.tobytes()[:100]))
    with pytest.raises(F.FormatError):
        F.decode_chunk_blob(b"\x01garbage")


def test_wire_framing_and_chunk_updated_layout():
    ids = (np.arange(4096) % 7).astype(np.uint8)
    body = F.encode_chunk_updated((-16, 32, 48), ids)
    # bincode 1.x: u32 LE variant 10, three i64 LE, 4096 ids, u64 LE metadata length 0
    assert body[:4] == b"\x0a\x00\x00\x00"
    assert body[4:12] == struct.pack("<q", -16) and body[12:20] == struct.pack("<q", 32)
    assert len(body) == 4 + 24 + 4096 + 8 and body[-8:] == b"\x00" * 8
    offs, got = F.decode_chunk_updated(body)
    assert list(offs) == [-16, 32, 48] and (got == ids).all()
    chat = struct.pack("<I", 11) + struct.pack("<Q", 2) + b"hi"                         # Chat(String): not a chunk
    assert F.decode_chunk_updated(chat) is None
    stream = F.frame(chat) + F.frame(body) + F.frame(F.encode_chunk_updated((-16, 32, 48), ids ^ 1)) + F.frame(chat)
    assert stream[:8] == struct.pack(">Q", len(chat))                                   # 8-byte BE length, generic_net.rs:201
    o, c = F.chunks_from_stream(stream)
    assert o.shape == (1, 3) and (c[0] == ids ^ 1).all()                                 # the later update wins
    with pytest.raises(F.FormatError):
        list(F.unframe(stream[:-1]))
    with pytest.raises(F.FormatError):
        F.decode_chunk_updated(body[:100])


def _make_world(path, chunks, offs, names=None):
    w = F.World(str(path), create=True)
    if names is not None:
        w.set_names(names)
    for o, c in zip(offs, chunks):
        w.store_chunk(o, c)
    w.commit()
    w.close()


def test_world_file_schema_and_reading(tmp_path):
    import sqlite3
    import mimas_b200
    chunks, offs = mimas_b200.gen_terrain_host(78, (-2, -1, -1), (3, 2, 2))
    chunks = np.minimum(chunks, 10)  # only built-in names
    _make_world(tmp_path / "w.sqlite", chunks, offs)
    conn = sqlite3.connect(str(tmp_path / "w.sqlite"))
    assert conn.execute("PRAGMA application_id;").fetchone()[0] & 0xFFFFFFFF == 0x84EEAE3C
    assert conn.execute("PRAGMA user_version;").fetchone()[0] == 2
    assert conn.execute("SELECT x, y, z FROM chunks ORDER BY z, y, x").fetchone() == (-2, -1, -1)   # chunk coordinates
    conn.close()
    w = F.World(str(tmp_path / "w.sqlite"))
    assert w.names() == F.BUILTIN_NAMES and w.chunk_count() == 12
    got = list(w.batches(batch=5))
    assert [len(b[1]) for b in got] == [5, 5, 2]
    o, c = np.concatenate([b[0] for b in got]), np.concatenate([b[1] for b in got])
    assert o.tobytes() == offs.tobytes() and c.tobytes() == chunks.tobytes()
    w.close()
    # a database of another application is refused (expect_user_ver, map_storage.rs:76-92)
    other = sqlite3.connect(str(tmp_path / "other.sqlite"))
    other.execute("CREATE TABLE chunks (x INTEGER)")
    other.commit()
    other.close()
    with pytest.raises(F.FormatError):
        F.World(str(tmp_path / "other.sqlite"))


@pytest.mark.gpu
def test_world_file_meshed_end_to_end(tmp_path):
    """world file -> decode -> CUDA mesher == oracle on the same chunks; same through the wire format."""
    import mimas_b200
    chunks, offs = mimas_b200.gen_terrain_host(78, (-4, -4, -2), (8, 8, 4))
    names = [f"default:b{i}" for i in range(56)]
    _make_world(tmp_path / "w.sqlite", chunks, offs, names)
    w = F.World(str(tmp_path / "w.sqlite"))
    oc = oracle_cache(G)
    with mimas_b200.Mesher(0, product_cache(G)) as m:
        for o, c in w.batches(batch=100):
            assert_batch_equal(m.mesh_batch(o, c), oracle.mesh_batch(c, o, oc, threads=4))
        stream = b"".join(F.frame(F.encode_chunk_updated(o, c)) for o, c in zip(offs[:40], chunks[:40]))
        o, c = F.chunks_from_stream(stream)
        assert_batch_equal(m.mesh_batch(o, c), oracle.mesh_batch(c, o, oc, threads=4))
    w.close()
