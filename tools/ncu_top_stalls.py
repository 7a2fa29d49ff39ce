"""Summarise `ncu --page source --csv` output: instructions with the most stall samples per kernel."""
import csv
import sys

path = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 20
rows = list(csv.reader(open(path)))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
        continue
    if cur is not None:
        cur["rows"].append(r)
for b in blocks:
    hdr, data = b["rows"][0], b["rows"][1:]
    si, src = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Source")
    tot = sum(int(r[si] or 0) for r in data)
    print("==", b["name"][:60], "| total samples", tot, "| instructions", len(data))
    top = sorted(enumerate(data), key=lambda t: -int(t[1][si] or 0))[:topn]
    for i, r in sorted(top):
        print(f"{i:5d} {int(r[si]):6d} {100 * int(r[si]) / max(tot, 1):5.1f}%  {r[src][:100]}")
