"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and one step's sequence.
python tools/launch_summary.py gpurun_out/launches.csv [steps]"""
import csv, collections, sys
path = sys.argv[1]
lines = [l for l in open(path) if not l.startswith('==')]
agg = collections.OrderedDict(); seq = []
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    k = row['Kernel Name'][:64]; v = float(row['Metric Value'].replace(',', ''))
    u = row['Metric Unit']
    v = v / 1e3 if u == 'ns' else v * 1e3 if u == 'ms' else v
    seq.append((k, v)); a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print(f"{a[0]:4d} {a[1]:10.1f} us {100*a[1]/tot:5.1f}%  avg {a[1]/a[0]:8.1f}  {k}")
idx = [i for i, (k, v) in enumerate(seq) if 'csr_init' in k or 'csr_build_fused' in k]
if len(idx) > 2:
    i0, i1 = idx[-3], idx[-2]
    print("one step:", i1 - i0, "launches,", f"{sum(v for k, v in seq[i0:i1]):.1f} us")
