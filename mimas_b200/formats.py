"""Data formats either side of the mesher: how chunks reach `mesh_for_chunk` in the
reference -- from the world database and from the wire -- decoded on the host into the `n x 4096` byte
batches the CUDA path takes.  Only the block ids matter to the mesher (lib.rs:231-446 never looks at the
metadata), so metadata is validated for framing where cheap and otherwise skipped.

* world database: sqlite, application_id 0x84eeae3c, user_version <= 2 (mimas-server/map_storage.rs:24-29,
  76-92); table `chunks(x, y, z, content)` keyed by CHUNK coordinates (`pos / CHUNKSIZE`, :259,275);
  `content` = version byte (<= 1) + gzip stream of 4096 block ids in index order x*256+y*16+z, then for
  version 1 a big-endian u16 metadata count and the entries (:127-202); `kvstore['name_id_map']` = version
  byte 0, big-endian u16 count, then (u8 length, utf-8 name) per id (:204-247).
* wire: every message is an 8-byte big-endian length followed by a bincode 1.x body
  (mimas-common/generic_net.rs:122,201-206, Cargo.toml:19).  `ServerToClientMsg::ChunkUpdated(Vector3<isize>,
  MapChunkData)` is variant 10 of the enum (protocol.rs:37-53): u32 LE tag, three i64 LE (nalgebra's
  fixed-size storage serialises as a plain tuple), 4096 id bytes (serde-big-array: a tuple, no length),
  then the metadata map (u64 LE entry count, ...) (map.rs:44-62).  The position is in BLOCKS (the chunk
  origin, a multiple of 16: mapgen.rs:369, client.rs:263).
"""
from __future__ import annotations

import gzip
import sqlite3
import struct
from typing import Iterable, Iterator, List, Optional, Sequence, Tuple

import numpy as np

CHUNKSIZE = 16                       # mimas-common/map.rs:11
CHUNK_CELLS = CHUNKSIZE ** 3
MIMAS_SQLITE_APP_ID = 0x84EEAE3C     # map_storage.rs:27 (stored as i32)
USER_VERSION = 2                     # map_storage.rs:29
LEN_BYTES = 8                        # generic_net.rs:122
CHUNK_UPDATED_TAG = 10               # position of ChunkUpdated in ServerToClientMsg, protocol.rs:37-53
BUILTIN_NAMES = ["default:air", "default:water", "default:sand", "default:ground", "default:wood", "default:stone",
                 "default:leaves", "default:tree", "default:cactus", "default:coal", "default:iron_ore"]  # game_params.rs:239-253


class FormatError(ValueError):
    pass


# ---- chunk blob (map_storage.rs:127-202) ---------------------------------------------------------------
def decode_chunk_blob(blob: bytes, first_invalid_id: Optional[int] = None) -> np.ndarray:
    """`deserialize_mapchunk_data`: the 4096 block ids of a stored chunk.  `first_invalid_id` = number of
    names in the world's name-id map: ids at or beyond it are an error, as in `mb_from_id` (:169-172)."""
    if len(blob) < 1:
        raise FormatError("empty chunk blob")
    version = blob[0]
    if version > 1:
        raise FormatError(f"Unsupported map chunk version {version}")  # :153-157
    try:
        buf = gzip.decompress(blob[1:])
    except (OSError, EOFError) as e:
        raise FormatError(f"chunk blob is not a gzip stream: {e}") from None
    if len(buf) < CHUNK_CELLS:
        raise FormatError(f"chunk blob holds {len(buf)} bytes, a chunk needs {CHUNK_CELLS}")
    ids = np.frombuffer(buf, dtype=np.uint8, count=CHUNK_CELLS).copy()
    if first_invalid_id is not None and int(ids.max()) >= first_invalid_id:
        raise FormatError("invalid block number")
    if version > 0:  # metadata: count must be there; the entries are not the mesher's business
        if len(buf) < CHUNK_CELLS + 2:
            raise FormatError("version 1 chunk blob without a metadata count")
    return ids


def encode_chunk_blob(ids: np.ndarray) -> bytes:
    """`serialize_mapchunk_data` for a chunk without metadata: version 1, gzip(ids + u16 BE 0) (:127-149)."""
    ids = np.ascontiguousarray(ids, dtype=np.uint8).reshape(CHUNK_CELLS)
    return b"\x01" + gzip.compress(ids.tobytes() + struct.pack(">H", 0), compresslevel=1, mtime=0)


# ---- name-id map (map_storage.rs:204-247) ----------------------------------------------------------------
def decode_name_id_map(data: bytes) -> List[str]:
    if len(data) < 3:
        raise FormatError("name id map too short")
    if data[0] != 0:
        raise FormatError(f"Unsupported name id map version {data[0]}")
    (count,) = struct.unpack_from(">H", data, 1)
    if count >= 255:
        raise FormatError(f"Too many ids stored in name id map {count}")  # 255 is reserved, :229-233
    names, at = [], 3
    for _ in range(count):
        if at >= len(data):
            raise FormatError("name id map truncated")
        n = data[at]
        raw = data[at + 1:at + 1 + n]
        if len(raw) != n:
            raise FormatError("name id map truncated")
        name = raw.decode("utf-8").replace("::", ":")  # backwards compatibility, :241-244
        if name.count(":") != 1 or not all(name.split(":")):
            raise FormatError(f"bad block name {name!r}")
        names.append(name)
        at += 1 + n
    return names


def encode_name_id_map(names: Sequence[str]) -> bytes:
    out = bytearray(b"\x00" + struct.pack(">H", len(names)))
    for n in names:
        b = n.encode("utf-8")
        if len(b) >= 255:
            raise FormatError("name too long")
        out += bytes([len(b)]) + b
    return bytes(out)


# ---- world database (map_storage.rs:36-99, 256-290, 346-362) -----------------------------------------------
class World:
    """Read access to a mimas world file the way `SqliteStorageBackend` opens it, plus the writer half the
    tests need to build one (same schema, pragmas and blob format)."""

    def __init__(self, path: str, create: bool = False):
        self.conn = sqlite3.connect(path)
        if create:
            app_id = MIMAS_SQLITE_APP_ID - (1 << 32)  # as i32
            self.conn.execute(f"PRAGMA application_id = {app_id};")
            self.conn.execute(f"PRAGMA user_version = {USER_VERSION};")
            self.conn.execute("CREATE TABLE IF NOT EXISTS kvstore (kkey VARCHAR(16) PRIMARY KEY, content BLOB);")
            self.conn.execute("CREATE TABLE IF NOT EXISTS chunks (x INTEGER, y INTEGER, z INTEGER, content BLOB, "
                              "PRIMARY KEY(x, y, z))")
            self.conn.execute("CREATE TABLE IF NOT EXISTS player_kvstore (id_src INTEGER, id INTEGER, kkey VARCHAR(16), "
                              "content BLOB, PRIMARY KEY(id_src, id, kkey))")
            self.conn.commit()
        else:
            (app_id,) = self.conn.execute("PRAGMA application_id;").fetchone()
            if (app_id & 0xFFFFFFFF) != MIMAS_SQLITE_APP_ID:
                raise FormatError(f"expected app id {MIMAS_SQLITE_APP_ID:#x} but was {app_id & 0xFFFFFFFF:#x}")
            (uv,) = self.conn.execute("PRAGMA user_version;").fetchone()
            if uv > USER_VERSION:
                raise FormatError(f"user_version of database {uv} newer than maximum supported {USER_VERSION}")

    def close(self):
        self.conn.close()

    def names(self) -> List[str]:
        """`load_name_id_map`: the stored list, or the built-in one for worlds that have none (:346-354)."""
        row = self.conn.execute("SELECT content FROM kvstore WHERE kkey=?", ("name_id_map",)).fetchone()
        if row is None or row[0] is None:
            return list(BUILTIN_NAMES)
        v = row[0]
        return decode_name_id_map(v.encode("utf-8") if isinstance(v, str) else bytes(v))

    def set_names(self, names: Sequence[str]):
        self.conn.execute("INSERT OR REPLACE INTO kvstore (kkey, content) VALUES (?, ?);",
                          ("name_id_map", encode_name_id_map(names)))
        self.conn.commit()

    def store_chunk(self, pos_blocks: Sequence[int], ids: np.ndarray):
        """`store_chunk`: the key is the chunk coordinate pos / CHUNKSIZE (:257-264)."""
        x, y, z = (int(v) // CHUNKSIZE for v in pos_blocks)
        self.conn.execute("INSERT OR REPLACE INTO chunks (x, y, z, content) VALUES (?, ?, ?, ?);",
                          (x, y, z, encode_chunk_blob(ids)))

    def commit(self):
        self.conn.commit()

    def chunk_count(self) -> int:
        return int(self.conn.execute("SELECT COUNT(*) FROM chunks").fetchone()[0])

    def batches(self, batch: int = 4096) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
        """All stored chunks as (offs [k,3] int64 in BLOCKS, ids [k,4096] uint8) batches for `mesh_batch`."""
        first_invalid = len(self.names())
        offs, ids = [], []
        for x, y, z, content in self.conn.execute("SELECT x, y, z, content FROM chunks ORDER BY z, y, x"):
            offs.append((x * CHUNKSIZE, y * CHUNKSIZE, z * CHUNKSIZE))
            ids.append(decode_chunk_blob(bytes(content), first_invalid))
            if len(ids) == batch:
                yield np.array(offs, dtype=np.int64), np.stack(ids)
                offs, ids = [], []
        if ids:
            yield np.array(offs, dtype=np.int64), np.stack(ids)


# ---- wire (generic_net.rs:122,201-240; protocol.rs:37-53) -----------------------------------------------------
def frame(body: bytes) -> bytes:
    return struct.pack(">Q", len(body)) + body


def unframe(stream: bytes) -> Iterator[bytes]:
    """Message bodies of a byte stream of length-prefixed messages; a trailing partial message is an error."""
    at = 0
    while at < len(stream):
        if at + LEN_BYTES > len(stream):
            raise FormatError("truncated length prefix")
        (n,) = struct.unpack_from(">Q", stream, at)
        at += LEN_BYTES
        if at + n > len(stream):
            raise FormatError("truncated message")
        yield stream[at:at + n]
        at += n


def encode_chunk_updated(pos_blocks: Sequence[int], ids: np.ndarray) -> bytes:
    """bincode body of `ServerToClientMsg::ChunkUpdated(pos, chunk)` for a chunk without metadata."""
    ids = np.ascontiguousarray(ids, dtype=np.uint8).reshape(CHUNK_CELLS)
    return struct.pack("<I3q", CHUNK_UPDATED_TAG, *(int(v) for v in pos_blocks)) + ids.tobytes() + struct.pack("<Q", 0)


def decode_chunk_updated(body: bytes) -> Optional[Tuple[np.ndarray, np.ndarray]]:
    """(offs [3] int64, ids [4096] uint8) if the message is a ChunkUpdated, else None."""
    if len(body) < 4:
        raise FormatError("message without a variant tag")
    (tag,) = struct.unpack_from("<I", body, 0)
    if tag != CHUNK_UPDATED_TAG:
        return None
    need = 4 + 24 + CHUNK_CELLS + 8
    if len(body) < need:
        raise FormatError(f"ChunkUpdated of {len(body)} bytes, at least {need} expected")
    offs = np.array(struct.unpack_from("<3q", body, 4), dtype=np.int64)
    ids = np.frombuffer(body, dtype=np.uint8, count=CHUNK_CELLS, offset=28).copy()
    return offs, ids


def chunks_from_stream(stream: bytes) -> Tuple[np.ndarray, np.ndarray]:
    """Every ChunkUpdated of a captured server->client byte stream as one batch (later updates of the same
    chunk replace earlier ones, as the client's map does: client.rs:420-428)."""
    latest = {}
    for body in unframe(stream):
        r = decode_chunk_updated(body)
        if r is not None:
            latest[tuple(int(v) for v in r[0])] = r[1]
    if not latest:
        return np.zeros((0, 3), np.int64), np.zeros((0, CHUNK_CELLS), np.uint8)
    keys = list(latest)
    return np.array(keys, dtype=np.int64), np.stack([latest[k] for k in keys])
