import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]
d = collections.OrderedDict()
for r in rows[hdr + 1:]:
    rec = dict(zip(H, r))
    e = d.setdefault(rec['ID'], {'k': rec['Kernel Name'].split('(')[0][-28:]})
    e[rec['Metric Name']] = float(rec['Metric Value'].replace(',', ''))
agg = collections.OrderedDict()
for e in d.values():
    key = (e['k'], round(e['smsp__inst_executed.sum'] / 1e6))
    agg.setdefault(key, []).append(e)
for (k, mi), es in agg.items():
    t = sorted(x['gpu__time_duration.sum'] for x in es)[len(es) // 2] / 1e3
    b = sorted(x['dram__bytes_read.sum'] for x in es)[len(es) // 2] / 1e6
    print(f"{k:30s} x{len(es):2d}  {t:9.1f} us  {b:9.1f} MB read  {mi:6d} M inst")
