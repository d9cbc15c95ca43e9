"""python profiles/summarise_launches.py gpurun_out/<tag>_launches.csv "<command that was profiled>" > profiles/<tag>_launches_summary.csv
Groups the ncu launch list (--metrics gpu__time_duration.sum) by kernel (template arguments kept, parameters dropped)."""
import collections, csv, io, sys
rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
agg = collections.OrderedDict()
for x in csv.DictReader(io.StringIO("".join(rows))):
    k = x["Kernel Name"].split("(")[0]
    us = float(x["Metric Value"].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}[x["Metric Unit"]]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
print("# ncu --metrics gpu__time_duration.sum --clock-control none: %s" % (sys.argv[2] if len(sys.argv) > 2 else ""))
print("# (whole process: upload, t=0 exchange/traverse, warm-up, timed run, C-ABI round-trip arm; cold-cache serialised times: compare SHARES)")
print("kernel,launches,total_us,share")
for k, a in sorted(agg.items(), key=lambda t: -t[1][1]):
    print('"%s",%d,%.1f,%.3f' % (k, a[0], a[1], a[1] / tot))
