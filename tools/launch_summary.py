"""Launch list (ncu --metrics gpu__time_duration.sum --csv) -> per-kernel share table:
   python tools/launch_summary.py launches.csv [out.md]"""
import collections
import csv
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
H = rows[0]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    n = r[ki].split("(")[0]
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else v * 1e3 if r[ui] == "ms" else v
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
out = ["%d launches, %.1f us of kernel time (cold-cache, serialised: shares are what to read)" % (len(rows) - 1, tot), "",
       "| kernel | launches | total us | us/launch | share |", "|---|---|---|---|---|"]
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    out.append("| %s | %d | %.1f | %.1f | %.1f%% |" % (n[:70], c, t, t / c, 100 * t / tot))
text = "\n".join(out) + "\n"
print(text)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(text)
