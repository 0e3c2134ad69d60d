"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name: count, total, mean, share."""
import collections, csv, re, sys
path = sys.argv[1]
lines = [l for l in open(path) if not l.startswith('==')]
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    v = float(row['Metric Value'].replace(',', ''))
    unit = row['Metric Unit']
    v = v / 1e3 if unit == 'ns' else (v * 1e3 if unit == 'ms' else v)
    name = re.sub(r'\(.*', '', row['Kernel Name'])
    name = re.sub(r'void |<unnamed>::', '', name)
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
print('total %.1f us over %d launches' % (tot, sum(n for n, _ in agg.values())))
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print('%-52s n=%5d total=%10.1f us avg=%8.1f us share=%5.1f%%' % (k[:52], n, t, t / n, 100 * t / tot))
