"""Block parameters -> draw styles: the host-side feeder of the mesher's block table.

Mirrors the part of the reference's game-params loader the mesher depends on:
  * ``NameIdMap::builtin_name_list``                mimas-common/game_params.rs:238-253
  * id assignment in ``[[block]]`` order             mimas-common/game_params.rs:478-487
  * ``block_params.resize_with(.., Default)``        mimas-common/game_params.rs:494-495
  * ``BlockParams::default()`` (black ``Colored``)   mimas-common/game_params.rs:178-195
  * ``color`` / ``texture`` / ``crossed`` handling   mimas-common/game_params.rs:499-545
  * ``inv_texture`` (explicit or the crossed image)  mimas-common/game_params.rs:546-556
Everything else in game-params.toml (recipes, tools, ores, ...) is game logic and ignored.
"""
from __future__ import annotations

import tomllib
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple, Union

BUILTIN_NAMES: Tuple[str, ...] = (
    "default:air", "default:water", "default:sand", "default:ground", "default:wood", "default:stone",
    "default:leaves", "default:tree", "default:cactus", "default:coal", "default:iron_ore",
)


# DrawStyle, mimas-common/game_params.rs:22-29
@dataclass(frozen=True)
class Colored:
    rgba: Tuple[float, float, float, float]


@dataclass(frozen=True)
class Crossed:
    path: str


@dataclass(frozen=True)
class Texture:
    path: str


@dataclass(frozen=True)
class TextureSidesTop:
    sides: str
    top_bottom: str


@dataclass(frozen=True)
class TextureSidesTopBottom:
    sides: str
    top: str
    bottom: str


DrawStyle = Union[Colored, Crossed, Texture, TextureSidesTop, TextureSidesTopBottom]
DEFAULT_DRAW_STYLE = Colored((0.0, 0.0, 0.0, 1.0))  # impl Default for DrawStyle, game_params.rs:172-176


@dataclass
class BlockParams:
    """Only the two fields TextureIdCache::from_hdl reads (mimas-meshgen/lib.rs:89-97)."""
    draw_style: Optional[DrawStyle] = DEFAULT_DRAW_STYLE
    inv_texture: Optional[str] = None


class NameIdMap:
    """mimas-common/game_params.rs:255-289: names get consecutive u8 ids on first sight."""

    def __init__(self, names: Sequence[str] = BUILTIN_NAMES):
        self.id_to_name: List[str] = []
        self.name_to_id = {}
        for n in names:
            self.get_or_extend(n)

    def get_or_extend(self, name: str) -> int:
        if name not in self.name_to_id:
            if len(self.id_to_name) >= 256:
                raise ValueError("more than 256 block ids")
            self.name_to_id[name] = len(self.id_to_name)
            self.id_to_name.append(name)
        return self.name_to_id[name]

    @property
    def first_invalid_id(self) -> int:
        return len(self.id_to_name)


@dataclass
class GameParams:
    name_id_map: NameIdMap
    block_params: List[BlockParams]


def _draw_style(block: dict) -> Optional[DrawStyle]:
    color = block.get("color")
    if color is False:  # `color = false` means "no colour", game_params.rs:508-510
        color = None
    texture = block.get("texture")
    if color is not None and texture is not None:
        raise ValueError("Both color and texture specified")
    if color is not None:
        if len(color) != 4:
            raise ValueError("color needs four components")
        return Colored(tuple(float(v) for v in color))
    if texture is None:
        return None
    if isinstance(texture, str):
        return Crossed(texture) if block.get("crossed") is True else Texture(texture)
    if isinstance(texture, list):
        if len(texture) == 2:
            return TextureSidesTop(*texture)
        if len(texture) == 3:
            return TextureSidesTopBottom(*texture)
        raise ValueError(f"false number of textures: {len(texture)}")
    raise ValueError("false type")


def load_params_from_toml(text: str) -> GameParams:
    doc = tomllib.loads(text)
    blocks = doc.get("block", [])
    ids = NameIdMap()
    for b in blocks:  # first pass: ids in file order
        ids.get_or_extend(b["name"])
    params = [BlockParams() for _ in range(ids.first_invalid_id)]
    for b in blocks:  # second pass: parameters
        ds = _draw_style(b)
        inv = b.get("inv_texture")
        if inv is None and isinstance(ds, Crossed):
            inv = ds.path
        params[ids.name_to_id[b["name"]]] = BlockParams(draw_style=ds, inv_texture=inv)
    return GameParams(name_id_map=ids, block_params=params)
