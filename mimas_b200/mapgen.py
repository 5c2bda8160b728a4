"""Host side of the device mapgen (csrc/mm_mapgen.cu): the `[[mapgen.ore]]` / `[[mapgen.plant]]` tables and the block
roles of game-params.toml, as the reference's loader reads them:
  * `Ore`, `Plant`, `MapgenParams`          mimas-common/game_params.rs:134-154
  * parsing of the two tables                mimas-common/game_params.rs:648-718
  * `BlockRoles::new` (names -> ids)         mimas-common/game_params.rs:197-218
The default game's tables ship as data/default_mapgen_params.json (derived from the reference's
mimas-server/game-params.toml by data/make_default_mapgen.py), so no reference checkout is needed at run time."""
from __future__ import annotations

import ctypes as C
import json
import os
import tomllib
from dataclasses import dataclass, field
from typing import List

from .game_params import NameIdMap

MAX_ORES, MAX_PLANTS = 8, 16


@dataclass
class Ore:                      # game_params.rs:134-143
    block: int
    noise_seed: bytes
    pcg_seed: bytes
    freq: float
    pcg_chance: float
    limit_a: float
    limit_b: float
    limit_boundary: int


@dataclass
class Plant:                    # game_params.rs:145-149
    block: int
    pcg_seed: bytes
    pcg_limit: float


@dataclass
class MapgenParams:
    roles: dict                 # air, water, sand, ground, ground_top, stone -> block id
    ores: List[Ore] = field(default_factory=list)
    plants: List[Plant] = field(default_factory=list)

    def to_json(self) -> dict:
        return {"roles": self.roles,
                "ores": [dict(o.__dict__, noise_seed=o.noise_seed.decode(), pcg_seed=o.pcg_seed.decode()) for o in self.ores],
                "plants": [dict(p.__dict__, pcg_seed=p.pcg_seed.decode()) for p in self.plants]}

    @classmethod
    def from_json(cls, d: dict) -> "MapgenParams":
        return cls(dict(d["roles"]),
                   [Ore(**dict(o, noise_seed=o["noise_seed"].encode(), pcg_seed=o["pcg_seed"].encode())) for o in d["ores"]],
                   [Plant(**dict(p, pcg_seed=p["pcg_seed"].encode())) for p in d["plants"]])


ROLE_NAMES = {"air": "default:air", "water": "default:water", "sand": "default:sand", "ground": "default:ground",
              "ground_top": "default:ground_with_grass", "stone": "default:stone"}


def _seed8(s: str, what: str) -> bytes:
    b = s.encode()
    if len(b) != 8:
        raise ValueError(f"{what} needs to be 8 bytes long but has length {len(b)}")   # game_params.rs:659-661
    return b


def load_mapgen_params_from_toml(text: str) -> MapgenParams:
    doc = tomllib.loads(text)
    ids = NameIdMap()
    for b in doc.get("block", []):          # ids in file order, as the block loader assigns them
        ids.get_or_extend(b["name"])
    roles = {}
    for role, name in ROLE_NAMES.items():
        if name not in ids.name_to_id:
            raise ValueError(f"Couldn't find id for builtin role '{name}'")
        roles[role] = ids.name_to_id[name]
    mp = MapgenParams(roles)
    mg = doc.get("mapgen", {})
    for o in mg.get("ore", []):
        if o["name"] not in ids.name_to_id:
            raise ValueError("invalid name")
        mp.ores.append(Ore(ids.name_to_id[o["name"]], _seed8(o["noise_seed"], "noise_seed"), _seed8(o["pcg_seed"], "pcg_seed"),
                           float(o["freq"]), float(o["pcg_limit"]), float(o["limit_a"]), float(o["limit_b"]),
                           int(o["limit_boundary"])))
    for p in mg.get("plant", []):
        if p["name"] not in ids.name_to_id:
            raise ValueError("invalid name")
        mp.plants.append(Plant(ids.name_to_id[p["name"]], _seed8(p["pcg_seed"], "pcg_seed"), float(p["pcg_limit"])))
    if len(mp.ores) > MAX_ORES or len(mp.plants) > MAX_PLANTS:
        raise ValueError("more ores / plants than the device generator takes")
    return mp


def default_mapgen_params() -> MapgenParams:
    here = os.path.dirname(os.path.abspath(__file__))
    return MapgenParams.from_json(json.load(open(os.path.join(here, "data", "default_mapgen_params.json"))))


# ---- mm_mapgen_params (include/mimas_meshgen_cuda.h) ---------------------------------------------------------------
class COre(C.Structure):
    _fields_ = [("block", C.c_uint8), ("noise_seed", C.c_uint8 * 8), ("pcg_seed", C.c_uint8 * 8), ("freq", C.c_double),
                ("pcg_chance", C.c_double), ("limit_a", C.c_double), ("limit_b", C.c_double), ("limit_boundary", C.c_int64)]


class CPlant(C.Structure):
    _fields_ = [("block", C.c_uint8), ("pcg_seed", C.c_uint8 * 8), ("pcg_limit", C.c_double)]


class CMapgenParams(C.Structure):
    _fields_ = [("air", C.c_uint8), ("water", C.c_uint8), ("sand", C.c_uint8), ("ground", C.c_uint8),
                ("ground_top", C.c_uint8), ("stone", C.c_uint8), ("n_ores", C.c_uint32), ("n_plants", C.c_uint32),
                ("ores", COre * MAX_ORES), ("plants", CPlant * MAX_PLANTS)]


def to_c(mp: MapgenParams) -> CMapgenParams:
    r = mp.roles
    c = CMapgenParams(r["air"], r["water"], r["sand"], r["ground"], r["ground_top"], r["stone"], len(mp.ores), len(mp.plants))
    for i, o in enumerate(mp.ores):
        c.ores[i] = COre(o.block, (C.c_uint8 * 8)(*o.noise_seed), (C.c_uint8 * 8)(*o.pcg_seed), o.freq, o.pcg_chance,
                         o.limit_a, o.limit_b, o.limit_boundary)
    for i, p in enumerate(mp.plants):
        c.plants[i] = CPlant(p.block, (C.c_uint8 * 8)(*p.pcg_seed), p.pcg_limit)
    return c
