"""ncu launch list (`--metrics gpu__time_duration.sum --csv --log-file X.csv`) -> per-kernel table for profiles/.
usage: python tools/ncu_launch_summary.py launches.csv "header line(s)" > profiles/rN_launches_<what>_summary.txt"""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
iname, ival, iunit = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
n = 0
for r in rows[1:]:
    if r[iname] == "Kernel Name":
        continue
    v = float(r[ival].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[iunit], 1e-3)
    a = agg.setdefault(r[iname], [0, 0.0])
    a[0] += 1
    a[1] += v
    n += 1
tot = sum(a[1] for a in agg.values())
for line in sys.argv[2:]:
    print("# " + line)
print("# %d launches captured, total %.2f ms device time; per-launch times are cold-cache/serialised -> compare SHARES" % (n, tot / 1e3))
print("# launches   total_ms  share   us/launch  kernel")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%6d   %8.3f  %5.1f%%  %8.1f  %s" % (c, t / 1e3, 100 * t / tot, t / c, k[:170]))
