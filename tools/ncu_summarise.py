"""Summarise `ncu -i X.ncu-rep --page raw --csv` (stdin or file) per kernel: launches, time, DRAM bytes, achieved GB/s,
pipe utilisation, top stall.  usage: python tools/ncu_summarise.py raw.csv [peak_GBs] > summary.txt"""
import csv, sys, re, collections, json
rows = list(csv.reader(open(sys.argv[1], newline="")))
peak = float(sys.argv[2]) if len(sys.argv) > 2 else 6448.1
hdr = rows[0]; data = rows[2:]
col = {n: i for i, n in enumerate(hdr)}
def num(r, name):
    i = col.get(name)
    if i is None: return None
    v = r[i].replace(",", "")
    try: return float(v)
    except ValueError: return None
def short(n):
    n = re.sub(r"\((?!bool).*", "", n)
    m = re.match(r"(?:void )?(k_\w+)(?:<(\w+))?", n)
    if m: return m.group(1) + ("<%s>" % m.group(2) if m.group(2) else "")
    return n[:60]
units = {n: rows[1][i] for n, i in col.items()}
def to_bytes(r, name):
    v = num(r, name)
    if v is None: return None
    u = units.get(name, "")
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
def to_ms(r, name):
    v = num(r, name)
    if v is None: return None
    u = units.get(name, "")
    return v * {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3, "usecond": 1e-3, "nsecond": 1e-6, "msecond": 1, "second": 1e3}.get(u, 1e-6)
stall_cols = [n for n in hdr if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")]
agg = collections.OrderedDict()
for r in data:
    k = short(r[col["Kernel Name"]])
    a = agg.setdefault(k, dict(n=0, ms=0.0, rd=0.0, wr=0.0, sm=0.0, fp64=0.0, occ=0.0, stalls=collections.Counter()))
    a["n"] += 1
    a["ms"] += to_ms(r, "gpu__time_duration.sum") or 0.0
    a["rd"] += to_bytes(r, "dram__bytes_read.sum") or 0.0
    a["wr"] += to_bytes(r, "dram__bytes_write.sum") or 0.0
    a["sm"] += num(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed") or 0.0
    a["fp64"] += num(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active") or num(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active") or 0.0
    a["occ"] += num(r, "sm__warps_active.avg.pct_of_peak_sustained_active") or 0.0
    for s_ in stall_cols:
        v = num(r, s_)
        if v: a["stalls"][s_[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] += v
print("%-34s %4s %9s %10s %10s %9s %6s %6s %6s %6s  %s" % ("kernel", "n", "ms/launch", "dram rd MB", "dram wr MB", "GB/s", "%peak", "sm%", "fp64%", "occ%", "top stalls (per issue)"))
out = {}
for k, a in agg.items():
    n = a["n"]; ms = a["ms"] / n; rd = a["rd"] / n; wr = a["wr"] / n
    gbs = (rd + wr) / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
    top = ", ".join("%s %.1f" % (s_, v / n) for s_, v in a["stalls"].most_common(3))
    print("%-34s %4d %9.3f %10.1f %10.1f %9.1f %6.1f %6.1f %6.1f %6.1f  %s" % (k, n, ms, rd / 1e6, wr / 1e6, gbs, 100 * gbs / peak, a["sm"] / n, a["fp64"] / n, a["occ"] / n, top))
    out[k] = {"launches": n, "ms_per_launch": ms, "dram_bytes_per_launch": rd + wr, "dram_gbs": gbs}
if len(sys.argv) > 3:
    json.dump(out, open(sys.argv[3], "w"), indent=1)
