"""TEST INFRASTRUCTURE (parity oracle) -- restates how the reference turns game-params.toml
into the mesher's `TextureIdCache`.  Not imported by the product package.

Follows (paths relative to the reference checkout):
  * mimas-common/game_params.rs:238-253   NameIdMap::builtin_name_list (ids 0..10)
  * mimas-common/game_params.rs:255-289   NameIdMap::from_name_list / get_or_extend
  * mimas-common/game_params.rs:478-487   first pass: ids in [[block]] order
  * mimas-common/game_params.rs:494-495   block_params.resize_with(first_invalid_id, Default)
  * mimas-common/game_params.rs:178-195   BlockParams::default(): draw_style = Colored black
  * mimas-common/game_params.rs:499-545   color / texture / crossed -> DrawStyle
  * mimas-common/game_params.rs:546-556   inv_texture (explicit, or the crossed texture)
  * mimas-meshgen/lib.rs:87-104           TextureIdCache::from_hdl call order
  * mimas-client/assets.rs:140-201        sequential TextureId allocation per DrawStyle

PARITY UNPINNED in the same sense as meshgen_oracle.c: the reference has no test that pins
the numeric texture ids; they follow from the allocation order restated here.
"""
from __future__ import annotations

import tomllib
from dataclasses import dataclass, field

BUILTIN_NAME_LIST = [  # game_params.rs:240-252
    "default:air", "default:water", "default:sand", "default:ground", "default:wood",
    "default:stone", "default:leaves", "default:tree", "default:cactus", "default:coal",
    "default:iron_ore",
]

# DrawStyle variants, game_params.rs:22-29
COLORED, CROSSED, TEXTURE, SIDES_TOP, SIDES_TOP_BOTTOM = "Colored", "Crossed", "Texture", "TextureSidesTop", "TextureSidesTopBottom"


@dataclass
class BlockParams:
    # BlockParams::default(), game_params.rs:178-195
    draw_style: tuple | None = (COLORED, (0.0, 0.0, 0.0, 1.0))
    inv_texture: str | None = None


@dataclass
class GameParams:
    names: list = field(default_factory=list)          # id -> name
    block_params: list = field(default_factory=list)   # id -> BlockParams


def parse_game_params(toml_text: str) -> GameParams:
    val = tomllib.loads(toml_text)
    names = list(BUILTIN_NAME_LIST)
    name_to_id = {n: i for i, n in enumerate(names)}
    blocks = val.get("block", [])
    for b in blocks:  # first pass, game_params.rs:483-487
        n = b["name"]
        if n not in name_to_id:
            name_to_id[n] = len(names)
            names.append(n)
    params = [BlockParams() for _ in names]  # resize_with(Default), :494-495
    for b in blocks:  # second pass, :499-611
        bid = name_to_id[b["name"]]
        texture = b.get("texture")
        crossed = b.get("crossed")
        color = b.get("color")
        if color is False or color is None:  # `color = false` -> None, :507-515
            color = None
        if color is not None and texture is not None:
            raise ValueError("Both color and texture specified")
        if color is not None:
            ds = (COLORED, tuple(float(c) for c in color))
        elif isinstance(texture, str):
            ds = (CROSSED, texture) if crossed is True else (TEXTURE, texture)
        elif isinstance(texture, list):
            if len(texture) == 2:
                ds = (SIDES_TOP, texture[0], texture[1])
            elif len(texture) == 3:
                ds = (SIDES_TOP_BOTTOM, texture[0], texture[1], texture[2])
            else:
                raise ValueError("false number of textures")
        elif texture is None:
            ds = None
        else:
            raise ValueError("false type")
        inv = b.get("inv_texture")
        if inv is None and ds is not None and ds[0] == CROSSED:  # :550-555
            inv = ds[1]
        params[bid] = BlockParams(draw_style=ds, inv_texture=inv)
    return GameParams(names=names, block_params=params)


@dataclass
class TextureIdCache:
    """mimas-meshgen/lib.rs:73-79, flattened to plain Python values."""
    fallback: tuple                    # (sides, top, bottom)
    styles: list                       # per id: None | ("Blocky", (s,t,b)) | ("Crossed", id)
    inv_texture_ids: list              # per id: None | id

    def to_entries(self):
        """256 rows of (kind, sides, top, bottom, crossed): the device block table.
        kind: 0 = not drawn, 1 = blocky, 2 = crossed.  ids >= len are blocky-fallback
        (lib.rs:105-109) and never crossed (lib.rs:110-113)."""
        rows = []
        for i in range(256):
            if i < len(self.styles):
                st = self.styles[i]
                if st is None:
                    rows.append((0, 0, 0, 0, 0))
                elif st[0] == "Blocky":
                    rows.append((1, st[1][0], st[1][1], st[1][2], 0))
                else:
                    rows.append((2, 0, 0, 0, st[1]))
            else:
                rows.append((1, self.fallback[0], self.fallback[1], self.fallback[2], 0))
        return rows


class HeadlessObtainer:
    """mimas-client/assets.rs:140-201,225-232 with the pixel loading removed: only the
    sequential id allocation (`add_asset` returns assets.len() before the push)."""

    def __init__(self):
        self.next_id = 0

    def _add(self):
        i = self.next_id
        self.next_id += 1
        return i

    def mesh_draw_st(self, ds):
        kind = ds[0]
        if kind == COLORED:  # assets.rs:148-152: id, id_h; new_tb(id, id_h)
            i = self._add()
            i_h = self._add()
            return ("Blocky", (i_h, i, i))  # (sides, top, bottom)
        if kind == CROSSED:  # :153-158
            return ("Crossed", self._add())
        if kind == TEXTURE:  # :159-164
            i = self._add()
            return ("Blocky", (i, i, i))
        if kind == SIDES_TOP:  # :165-173: id_s, id_tb; new_tb(id_tb, id_s)
            i_s = self._add()
            i_tb = self._add()
            return ("Blocky", (i_s, i_tb, i_tb))
        if kind == SIDES_TOP_BOTTOM:  # :174-186
            i_s = self._add()
            i_t = self._add()
            i_b = self._add()
            return ("Blocky", (i_s, i_t, i_b))
        raise ValueError(kind)

    def inv_texture_id(self, _path):  # assets.rs:189-194
        return self._add()


def texture_id_cache_from_params(gp: GameParams, ob=None) -> TextureIdCache:
    """TextureIdCache::from_hdl, mimas-meshgen/lib.rs:87-104 (call order matters)."""
    ob = ob or HeadlessObtainer()
    fallback = ob.mesh_draw_st((COLORED, (0.0, 0.0, 0.0, 1.0)))[1]          # :88
    styles = [ob.mesh_draw_st(p.draw_style) if p.draw_style is not None else None
              for p in gp.block_params]                                      # :89-91
    inv = [ob.inv_texture_id(p.inv_texture) if p.inv_texture is not None else None
           for p in gp.block_params]                                         # :95-97
    return TextureIdCache(fallback=fallback, styles=styles, inv_texture_ids=inv)


def default_cache_from_toml(toml_text: str) -> TextureIdCache:
    return texture_id_cache_from_params(parse_game_params(toml_text))
