"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by (kernel, grid): python tools/launch_summary.py file.csv [skip_first_n_launches]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1], errors='ignore')))
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hdr, data = None, []
for r in rows:
    if 'Kernel Name' in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        data.append(dict(zip(hdr, r)))
data = data[skip:]
agg = collections.OrderedDict()
for d in data:
    key = (d['Kernel Name'][:64], d.get('Grid Size', ''))
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += float(d['Metric Value'].replace(',', ''))
tot = sum(a[1] for a in agg.values())
for (k, g), a in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
    print('%-66s %-16s n=%4d  avg %9.1f us  share %.3f' % (k, g, a[0], a[1] / a[0] / 1000, a[1] / tot))
