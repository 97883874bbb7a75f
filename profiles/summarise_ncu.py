"""Summarise an ncu launch list (csv) and a --set full report into markdown (run where ncu is installed).
usage: python profiles/summarise_ncu.py <launches.csv> <report.ncu-rep> "<title>" "<command>" > profiles/x.md"""
import collections
import csv
import subprocess
import sys

launch_csv, rep, title, cmd = sys.argv[1:5]
print(f"# {title}\n\nCommand: `{cmd}` (exited 0 without ncu first)\n")
print("## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`; cold-cache, serialised: compare shares)\n")
lines = [l for l in open(launch_csv) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(lines):
    n = row["Kernel Name"].split("(")[0]
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v[1] for v in agg.values())
print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"| `{n}` | {c} | {t:.1f} | {t / c:.1f} | {t / tot:.3f} |")
print(f"\n## `ncu --set full --clock-control none` ({rep}, not committed)\n")
out = subprocess.run(f"ncu -i {rep} --page raw --csv", shell=True, capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
hdr, units = r[0], r[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]
idx = [hdr.index(w) for w in want if w in hdr]
print("| " + " | ".join(hdr[i].split(".")[0] for i in idx) + " |\n|" + "---|" * len(idx))
for row in r[2:]:
    print("| " + " | ".join((row[i].split("(")[0][:28] + (" " + units[i] if units[i] else "")) for i in idx) + " |")
