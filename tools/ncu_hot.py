"""Print the hottest SASS instructions (warp-stall samples) per kernel of an ncu source-page CSV.
usage: ncu -i X.ncu-rep --page source --csv > src.csv; python tools/ncu_hot.py src.csv [name-substr] [min-pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
sub = sys.argv[2] if len(sys.argv) > 2 else ""
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.8
kern, names, cur = [], [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = []; kern.append(cur); names.append(r[1]); continue
    if cur is not None: cur.append(r)
for ki, k in enumerate(kern):
    if sub not in names[ki]: continue
    hdr, data = k[0], k[1:]
    si, so = hdr.index("# Samples"), hdr.index("Source")
    stall = [h for h in hdr if h.startswith("stall_") and "Not" not in h]
    tot = sum(int(r[si]) for r in data) or 1
    print("kernel", ki, names[ki][:70], "samples", tot, "instrs", len(data))
    agg = {h: sum(int(r[hdr.index(h)]) for r in data) for h in stall}
    print("  ", {k2: v for k2, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
    for idx, r in enumerate(data):
        if int(r[si]) > tot * minpct / 100:
            st = sorted(((int(r[hdr.index(h)]), h) for h in stall), reverse=True)[0]
            print(f"{idx:5d} {int(r[si]):6d} {100*int(r[si])/tot:5.1f}%  {r[so].strip()[:72]:72s} {st[1]}")
