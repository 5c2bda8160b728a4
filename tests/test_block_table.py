"""Block table (TextureIdCache) derivation: committed fixture vs the expected answer written
in SURVEY.md §8, and (only where the reference checkout exists) vs a fresh derivation."""
import os

import pytest

from oracle import block_table as bt
from tests.util import load_golden

G = load_golden()
REF_TOML = "/root/reference/mimas-server/game-params.toml"


def test_fixture_matches_survey_expectation():
    e, names = G.entries, G.table["names"]
    assert G.n_known == 56 and G.fallback == (1, 0, 0)
    assert names[:12] == ["default:air", "default:water", "default:sand", "default:ground", "default:wood",
                          "default:stone", "default:leaves", "default:tree", "default:cactus", "default:coal",
                          "default:iron_ore", "default:ground_with_grass"]
    assert e[0] == (0, 0, 0, 0, 0)                      # air: not drawn
    assert e[1] == (1, 2, 2, 2, 0) and e[6] == (1, 7, 7, 7, 0)   # water, leaves: opaque blocky
    assert e[5] == (1, 6, 6, 6, 0)                      # stone
    assert e[7] == (1, 8, 9, 9, 0) and e[8] == (1, 10, 11, 11, 0)  # tree, cactus: sides / top+bottom
    assert e[11] == (1, 14, 15, 16, 0)                  # ground_with_grass: s, t, b
    assert e[23] == (1, 22, 23, 23, 0)                  # chest
    assert [r[0] for r in e[17:23]] == [0] * 6 and [r[0] for r in e[30:42]] == [0] * 12
    assert e[42] == (2, 0, 0, 0, 30)                    # sapling: crossed 30
    assert [r[4] for r in e[43:56]] == list(range(31, 44))
    inv = [i for i in G.table["inv_texture_ids"] if i is not None]
    assert inv == list(range(44, 76))


@pytest.mark.skipif(not os.path.exists(REF_TOML), reason="reference checkout not present (GPU box)")
def test_fixture_is_fresh():
    tic = bt.default_cache_from_toml(open(REF_TOML).read())
    assert [tuple(r) for r in tic.to_entries()[:56]] == G.entries
    assert tuple(tic.fallback) == G.fallback


def test_unnamed_ids_get_default_black_colour():
    """game_params.rs:494-495 + :178-195: builtin ids that the TOML never defines keep
    BlockParams::default() = Colored black, which allocates two ids (assets.rs:148-152)."""
    gp = bt.parse_game_params('[[block]]\nname = "default:air"\ncolor = false\n')
    tic = bt.texture_id_cache_from_params(gp)
    assert tic.fallback == (1, 0, 0)
    assert tic.styles[0] is None
    assert tic.styles[1] == ("Blocky", (3, 2, 2)) and tic.styles[10] == ("Blocky", (21, 20, 20))
