"""Tracked evidence files from one `bash tools/prof_r02.sh <tag> full` run (gpurun_out/r02_<tag>_*), written to profiles/:
    r02_bench_<tag>.json, r02_launches_<tag>.csv, r02_traffic.json (source-stamped), r02_<tag>_full.ncu-rep.xz,
    r02_<tag>_full_metrics.csv (per-launch extract of the raw page), r02_<tag>_stalls.txt (source page: stall shares).
Usage (here, no GPU needed):  python tools/extract_profiles.py final"""
import csv, os, re, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "final"
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
shutil.copy(f"{G}/r02_{tag}_bench.json", f"{P}/r02_bench_{tag}.json")
shutil.copy(f"{G}/r02_{tag}_launches.csv", f"{P}/r02_launches_{tag}.csv")
subprocess.run([sys.executable, f"{ROOT}/tools/traffic_from_launches.py", f"{G}/r02_{tag}_launches.csv", f"{P}/r02_traffic.json"],
               check=True, stdout=subprocess.DEVNULL)
rep = f"{G}/r02_{tag}_full.ncu-rep"
with open(f"{P}/r02_{tag}_full.ncu-rep.xz", "wb") as f:
    subprocess.run(["xz", "-9", "-T4", "-k", "-c", rep], check=True, stdout=f)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum"] + \
       [f"smsp__average_warps_issue_stalled_{r}_per_issue_active.ratio" for r in
        ("long_scoreboard", "short_scoreboard", "wait", "barrier", "math_pipe_throttle", "lg_throttle", "mio_throttle", "branch_resolving")]
idx = [h.index(w) for w in want if w in h]
names = ["stft", "pre.0", "pre.2", "fc.0", "fc.2", "fc.4", "band11 const", "rnn0.rec0", "rnn0.rec1", "rnn0.fc", "rnn1 (time axis)",
         "rnn2.rec0", "rnn2.rec1", "rnn2.fc", "rnn3 (time axis)", "back.0", "back.1", "back.2", "post.0", "post.2", "istft"]
with open(f"{P}/r02_{tag}_full_metrics.csv", "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["launch"] + [h[i] for i in idx])
    w.writerow([""] + [units[i] for i in idx])
    for n, r in zip(names, data):
        row = [r[i] for i in idx]
        row[0] = re.sub(r"\(.*", "", row[0]).replace("void ", "").replace("bsb::<unnamed>::", "")
        w.writerow([n] + row)
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], check=True, capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hd = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
out = ["Per-launch warp-stall sampling of ONE timed step (ncu --set full --import-source on, profiles/r02_%s_full.ncu-rep.xz):" % tag,
       "share of PC samples per stall reason, and the eight most-sampled instructions with their dominant reason.",
       "Tables in launch order: stft, pre.0, pre.2, fc.0, fc.2, fc.4, band-11 constant, rnn0.rec0, rnn0.rec1, rnn0.fc, rnn1, rnn2.rec0,",
       "rnn2.rec1, rnn2.fc, rnn3, back.0, back.1, back.2, post.0, post.2, istft.", ""]
last, k = None, 0
for t, hi in enumerate(hd):
    kn = [r for r in rows[max(0, hi - 4):hi] if r and r[0] == "Kernel Name"]
    kname = kn[-1][1] if kn else "?"
    short = re.sub(r"\(.*", "", kname).replace("void ", "").replace("bsb::<unnamed>::", "")
    end = hd[t + 1] - 1 if t + 1 < len(hd) else len(rows)
    hh = rows[hi]
    ci = {n: i for i, n in enumerate(hh)}
    d = [r for r in rows[hi + 1:end] if len(r) == len(hh)]
    def f(r, n):
        try:
            return float(r[ci[n]])
        except ValueError:
            return 0.0
    tot = sum(f(r, "# Samples") for r in d)
    if tot == 0 or (short, tot) == last:
        continue
    last = (short, tot)
    stalls = [n for n in hh if n.startswith("stall_") and "Not Issued" not in n]
    agg = sorted(((n, sum(f(r, n) for r in d)) for n in stalls), key=lambda kv: -kv[1])[:6]
    out.append(f"== {names[k] if k < len(names) else k}: {short}  ({tot:.0f} samples)")
    out.append("   stall reasons: " + ", ".join(f"{n[6:]} {100 * v / tot:.1f}%" for n, v in agg))
    for r in sorted(d, key=lambda r: -f(r, "# Samples"))[:8]:
        m = max(stalls, key=lambda n: f(r, n))
        out.append(f"   {100 * f(r, '# Samples') / tot:5.1f}%  {r[ci['Source']][:64]:64s} {m[6:]}")
    k += 1
open(f"{P}/r02_{tag}_stalls.txt", "w").write("\n".join(out) + "\n")
print("wrote", k, "stall tables;", len(data), "launch rows")
