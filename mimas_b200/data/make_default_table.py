"""Regenerates default_block_table.json from the reference's game-params.toml (build container only):
the block table of the default game, so that the package works without the reference checkout.

    python mimas_b200/data/make_default_table.py /root/reference/mimas-server/game-params.toml
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import mimas_b200  # noqa: E402


def main(path):
    p = mimas_b200.load_params_from_toml(open(path).read())
    tic = mimas_b200.TextureIdCache.from_hdl(p, mimas_b200.HeadlessAssets())
    t = tic.to_table()
    rows = [[int(e["kind"]), int(e["id_sides"]), int(e["id_top"]), int(e["id_bottom"]), int(e["id_crossed"])]
            for e in t[:len(p.block_params)]]
    out = {"source": "mimas-server/game-params.toml @ reference commit 68a46059",
           "derivation": "mimas_b200/data/make_default_table.py", "names": p.name_id_map.id_to_name,
           "fallback": [tic.fallback_id.id_sides, tic.fallback_id.id_top, tic.fallback_id.id_bottom], "entries": rows}
    json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "default_block_table.json"), "w"),
              indent=1)


if __name__ == "__main__":
    main(sys.argv[1])
