"""Markdown block with the numbers of profiles/r02_bench_final.json + r02_launches_final.csv (pasted into r02_summary.md)."""
import csv, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
d = json.load(open(f"{ROOT}/profiles/r02_bench_final.json"))
e, r, c, g, s = d["e2e"], d["roofline"], d["cpu_baseline"]["detail"], d["gpu_comparator"], d["streaming"]
rows = list(csv.reader(open(f"{ROOT}/profiles/r02_launches_final.csv")))
hi = [i for i, x in enumerate(rows) if x and x[0] == "ID"][0]
h, data = rows[hi], rows[hi + 1:]
ki, vi, mi, ii = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name"), h.index("ID")
L = {}
for x in data:
    if len(x) > vi:
        L.setdefault(int(x[ii]), {"name": x[ki]})[x[mi]] = float(x[vi].replace(",", ""))
names = ["stft", "pre.0", "pre.2", "fc.0", "fc.2", "fc.4", "band-11 const", "rnn0.rec0", "rnn0.rec1", "rnn0.fc", "rnn1 (time axis)", "rnn2.rec0",
         "rnn2.rec1", "rnn2.fc", "rnn3 (time axis)", "back.0", "back.1", "back.2", "post.0", "post.2", "istft"]
tot = sum(v["gpu__time_duration.sum"] for v in L.values())
print(f"stamp {bench.source_stamp()} / traffic file {json.load(open(ROOT + '/profiles/r02_traffic.json'))['source_stamp']}")
print(f"* **{d['ms_per_step']:.3f} ms/step = {d['value']:,.0f} audio-s/s** device-resident; e2e fp32 {e['value']:,.0f} ({e['ms_per_step']:.2f} ms/step), "
      f"int16 PCM {e['pcm16_value']:,.0f}, serial {e['serial_value']:,.0f}; copy ceiling {e['host_copy_ceiling_gbs']:.1f} GB/s = {e['host_copy_ceiling_value']:,.0f} audio-s/s")
print(f"* parity max rel err {d['parity']['max_rel_err']:.3e}, dSI-SDR {d['parity']['d_si_sdr_db']:.1e} dB, SI-SDR(y, y_ref) {d['parity']['si_sdr_vs_oracle_db']:.1f} dB")
print(f"* gpu_comparator {g['ms_per_step']:.2f} ms/step = {g['value']:,.0f} audio-s/s -> {g['speedup_device_resident']:.1f}x")
print(f"* cpu ({d['cpu_baseline']['kind']}, {d['cpu_baseline']['cores']} cores): cfg2 all {d['cpu_baseline']['value']:.0f}, 1 thread {c['cfg2_1thread']['audio_s_per_s']:.0f}; "
      f"cfg1 {c['cfg1_all']['ms']:.1f} / {c['cfg1_1thread']['ms']:.1f} ms; hop {c['hop_all']['ms_p50']:.2f} / {c['hop_1thread']['ms_p50']:.2f} ms; "
      f"device hop {s['device_ms_p50']:.3f}, host {s['host_ms_p50']:.3f}, ladspa {s['ladspa_host']['run_ms_p50']:.3f} ms")
print(f"* roofline: {r['achieved']:.1f} TFLOP/s = {r['frac']:.3f} of {r['peak']}, pipe {r['tensor_pipe_frac']:.3f}, avg launch {r['avg_launch_ms']:.4f} ms, "
      f"share {r['share_of_step']:.3f}, traffic {r['traffic']}")
for n, v in d["kernels"].items():
    print(f"  {n}: {v['ms']:.4f} ms, share {v['share_of_step']:.3f}, hbm_frac {v['hbm_frac']:.3f}")
print("| launch | µs | tensor % | DRAM rd / wr MB |")
for i, n in zip(sorted(L), names):
    v = L[i]
    print(f"| {n} | {v['gpu__time_duration.sum'] / 1e3:.1f} | {v.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 0):.1f} | "
          f"{v['dram__bytes_read.sum'] / 1e6:.0f} / {v['dram__bytes_write.sum'] / 1e6:.0f} |")
agg = {}
for v in L.values():
    k = re.search(r"(\w+_kernel)", v["name"]).group(1)
    agg[k] = agg.get(k, 0) + v["gpu__time_duration.sum"]
print({k: (round(t / 1e3), round(100 * t / tot, 1)) for k, t in agg.items()})
