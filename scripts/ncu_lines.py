#!/usr/bin/env python
"""Per-source-line instruction / sample shares of one captured launch of an ncu source page.
    ncu -i rep --page source --csv --print-source cuda,sass > src.csv
    python scripts/ncu_lines.py src.csv <launch index> [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = int(sys.argv[2]); top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
# launches are separated by "Function Name" of the first file repeating: count kernel blocks by 'Kernel Name'-like rows
blocks = []; cur = None; hdr = None; fname = None; fn = None
launch = -1; seen_files = set()
agg = {}
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split('/')[-1]
        continue
    if r and r[0] == "Function Name":
        key = (r[1], fname)
        if key in seen_files or launch < 0 or r[1] != fn:
            if r[1] != fn or key in seen_files:
                launch += 1; seen_files = set(); fn = r[1]
        seen_files.add(key)
        continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].strip().isdigit() and launch == want:
        k = (fname, int(r[0]), r[1].strip())
        a = agg.setdefault(k, [0, 0])
        a[0] += int(r[hdr.index("Instructions Executed")]); a[1] += int(r[hdr.index("# Samples")])
tot = sum(a[0] for a in agg.values()) or 1; ts = sum(a[1] for a in agg.values()) or 1
print("kernel", fn if launch >= want else None, "inst", tot, "samples", ts)
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]}:{k[1]:4d} inst {a[0]:8d} {100*a[0]/tot:5.1f}%  samp {a[1]:5d} {100*a[1]/ts:5.1f}%  {k[2][:80]}")
