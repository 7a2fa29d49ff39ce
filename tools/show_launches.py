import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv")) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot = {}
for r in rows[1:]:
    tot.setdefault(r[ki][:48], []).append(float(r[vi]))
s = 0
for k, v in tot.items():
    print(f"{k:50s} n={len(v):2d} avg={sum(v)/len(v)/1000:7.2f} us")
    s += sum(v) / len(v) * (len(v) / max(1, min(len(x) for x in tot.values())))
print(f"sum per step ~ {s/1000:.1f} us")
