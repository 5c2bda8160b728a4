#!/usr/bin/env python
"""Dump per-SASS-instruction executed counts / samples of the first captured launch of an ncu report.
    python scripts/ncu_sass_dump.py rep.ncu-rep out.tsv"""
import csv, io, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], check=True,
                     capture_output=True, text=True).stdout
src = list(csv.reader(io.StringIO(txt)))
starts = [i for i, r in enumerate(src) if r and r[0] == "Address"]
h = src[starts[0]]
end = starts[1] - 1 if len(starts) > 1 else len(src)
b = [r for r in src[starts[0] + 1:end] if len(r) == len(h)]
si, ie = h.index("# Samples"), h.index("Instructions Executed")
with open(out, "w") as f:
    cum = 0
    for i, r in enumerate(b):
        e = int(r[ie]) if r[ie].isdigit() else 0
        cum += e
        f.write(f"{i}\t{e}\t{cum}\t{r[si]}\t{r[1].strip()}\n")
