#!/usr/bin/env python
"""A longer run of the three-way differential fuzz than the test suite affords (tests/test_oracle_fuzz.py runs
seeds 0..511 on every CPU test run): seeds [first, first + count) through tests/fuzz_util.three_way on all host
cores; prints one JSON line with the disagreements found.  CPU only, test infrastructure.
    python tests/golden/fuzz_campaign.py 512 20000 > tests/golden/fuzz_campaign.json"""
import json
import os
import sys
import time
from concurrent.futures import ProcessPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import fuzz_util  # noqa: E402

if __name__ == "__main__":
    first, count = int(sys.argv[1]), int(sys.argv[2])
    t0 = time.time()
    with ProcessPoolExecutor(max_workers=os.cpu_count() or 1) as ex:
        bad = [r for r in ex.map(fuzz_util.three_way, range(first, first + count), chunksize=16) if r is not None]
    print(json.dumps({"what": "tests/fuzz_util.three_way: C oracle vs Python transliteration vs groupby mesher, byte compare",
                      "first_seed": first, "cases": count, "disagreements": len(bad), "examples": bad[:5],
                      "seconds": round(time.time() - t0, 1), "workers": os.cpu_count()}))
