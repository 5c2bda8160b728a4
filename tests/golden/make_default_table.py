"""Regenerates tests/golden/default_block_table.json from the reference's game content.

Runs ONLY in the build container (reads /root/reference/mimas-server/game-params.toml, which
does not exist on the GPU box); the JSON it writes is the committed fixture that travels.
It is derived data (56 rows of numeric ids), not reference source.

    python tests/golden/make_default_table.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.block_table import parse_game_params, texture_id_cache_from_params  # noqa: E402

SRC = "/root/reference/mimas-server/game-params.toml"


def main():
    gp = parse_game_params(open(SRC).read())
    tic = texture_id_cache_from_params(gp)
    out = {
        "source": "mimas-server/game-params.toml @ reference commit 68a46059",
        "derivation": "oracle/block_table.py (game_params.rs:238-253,478-556; lib.rs:87-104; assets.rs:140-201)",
        "n_ids": len(gp.names),
        "fallback": list(tic.fallback),
        "names": gp.names,
        # rows: [kind, sides, top, bottom, crossed]; kind 0 none, 1 blocky, 2 crossed
        "entries": [list(r) for r in tic.to_entries()[:len(gp.names)]],
        "inv_texture_ids": tic.inv_texture_ids,
    }
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "default_block_table.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", len(gp.names), "ids")


if __name__ == "__main__":
    main()
