"""Per-kernel mean durations from an `ncu --metrics gpu__time_duration.sum --csv` launch list (last N launches)."""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10 and r[0].isdigit()]
last = int(sys.argv[2]) if len(sys.argv) > 2 else 10
agg = collections.OrderedDict()
for r in rows[-last:]:
    k = r[4][:64]
    agg.setdefault(k, []).append(float(r[-1]) / 1e6)
tot = 0
for k, v in agg.items():
    print(f"{sum(v)/len(v):8.3f} ms x{len(v)}  {k}"); tot += sum(v)
print(f"{tot:8.3f} ms total of last {last}")
