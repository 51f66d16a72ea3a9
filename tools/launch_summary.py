"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.  usage: python tools/launch_summary.py list.csv"""
import csv, re, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
col = {n: i for i, n in enumerate(rows[hi])}
agg = collections.OrderedDict(); tot = 0.0; n = 0
for r in rows[hi + 1:]:
    if len(r) < len(rows[hi]) or r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\((?!bool).*", "", r[col["Kernel Name"]])
    name = re.sub(r"^void ", "", name)[:70]
    ms = float(r[col["Metric Value"]].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[col["Metric Unit"]], 1e-6)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ms; tot += ms; n += 1
print("%d launches, %.2f ms of kernel time" % (n, tot))
for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%9.3f ms %5.1f%% %4d  %s" % (ms, 100 * ms / tot, c, k))
