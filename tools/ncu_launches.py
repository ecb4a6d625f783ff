"""Summarise an `ncu --csv --metrics ...` log: one line per launch (us, MB read/written, tensor %)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr, data = rows[hi], rows[hi + 1:]
ki, vi, mi, ii = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name"), hdr.index("ID")
out = {}
for r in data:
    if len(r) <= vi: continue
    d = out.setdefault(int(r[ii]), {"name": r[ki]})
    d[r[mi]] = float(r[vi].replace(",", ""))
for i in sorted(out):
    d = out[i]
    nm = d["name"].replace("bsb::<unnamed>::", "").replace("void ", "")[:34]
    print(f"{i:4d} {nm:34s} {d.get('gpu__time_duration.sum',0)/1e3:8.1f} us  rd {d.get('dram__bytes_read.sum',0)/1e6:7.1f} MB  wr {d.get('dram__bytes_write.sum',0)/1e6:7.1f} MB  tensor {d.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',0):5.1f}%")
