import csv, collections, sys
path = sys.argv[1]
with open(path) as f:
    lines = [l for l in f if not l.startswith('==')]
agg = collections.defaultdict(list); seq = []
for row in csv.DictReader(lines):
    v = float(row['Metric Value'].replace(',', '')); unit = row['Metric Unit']
    if unit == 'ns': v /= 1000
    elif unit == 'ms': v *= 1000
    name = row['Kernel Name'].split('(')[0].replace('void <unnamed>::', '')[:48]
    agg[name].append(v); seq.append((name, row.get('Grid Size', ''), row.get('Block Size', ''), v))
tot = sum(sum(v) for v in agg.values())
print("total us", round(tot, 1), "kernels", len(seq))
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{sum(v):9.1f} us  n={len(v):3d}  mean={sum(v)/len(v):7.1f}  max={max(v):7.1f}  {k}")
if len(sys.argv) > 2:
    for s in seq[:int(sys.argv[2])]: print(s)
