"""median / min per-launch time of an `ncu --metrics gpu__time_duration.sum --csv` log.  usage: parse_launch_times.py file.csv [name filter]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
flt = sys.argv[2] if len(sys.argv) > 2 else ""
hdr = None; v = []
for r in rows:
    if "Kernel Name" in r: hdr = r; continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if flt not in d["Kernel Name"]: continue
        try: x = float(d["Metric Value"].replace(",", ""))
        except ValueError: continue
        u = d["Metric Unit"]; v.append(x / 1000 if u == "ns" else (x * 1000 if u == "ms" else x))
v.sort()
if not v: sys.exit("%s: no launch matches %r" % (sys.argv[1], flt))
print(sys.argv[1], flt, "launches", len(v), "median us %.1f min %.1f max %.1f" % (v[len(v) // 2], v[0], v[-1]))
