"""Regenerates data/default_mapgen_params.json (and its twin under tests/golden/) from the reference's game content.
Runs ONLY in the build container (reads /root/reference/mimas-server/game-params.toml); the JSON is derived data
(block ids, seeds and thresholds of the [[mapgen.ore]] / [[mapgen.plant]] tables), not reference source.

    python mimas_b200/data/make_default_mapgen.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from mimas_b200.mapgen import load_mapgen_params_from_toml  # noqa: E402

SRC = "/root/reference/mimas-server/game-params.toml"

if __name__ == "__main__":
    mp = load_mapgen_params_from_toml(open(SRC).read())
    d = dict(source="mimas-server/game-params.toml @ reference commit 68a46059",
             derivation="mimas_b200/mapgen.py (game_params.rs:134-154, 197-218, 648-718)", **mp.to_json())
    for path in (os.path.join(HERE, "default_mapgen_params.json"), os.path.join(ROOT, "tests", "golden", "default_mapgen_params.json")):
        with open(path, "w") as f:
            json.dump(d, f, indent=1)
    print("wrote", len(mp.ores), "ores,", len(mp.plants), "plants; roles", mp.roles)
