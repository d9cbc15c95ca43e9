"""ncu -i X.ncu-rep --page source --csv > src.csv ; python profiles/sass_counts.py src.csv  -> per-SASS-instruction counts of the first captured launch"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, ist = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
data = []
for r in rows[2:]:
    if len(r) <= ia or not r[ia].isdigit():
        if data:
            break
        continue
    data.append(r)
tot = sum(int(r[ia]) for r in data)
print('total warp instr', tot, 'n sass', len(data))
for k, r in enumerate(data):
    print(k, int(r[ia]), r[ist], r[isrc].strip()[:110])
