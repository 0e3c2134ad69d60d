"""Per-step kernel shares from an ncu launch list (gpu__time_duration.sum): the last complete critic step and generator step
of the run, found through their closing adam_kernel launches."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
rows = [r for r in csv.DictReader(lines) if r.get('Metric Name') == 'gpu__time_duration.sum']
def us(r):
    v = float(r['Metric Value'].replace(',', '')); u = r['Metric Unit']
    return v / 1e3 if u == 'ns' else (v * 1e3 if u == 'ms' else v)
names = [re.sub(r'void |<unnamed>::', '', re.sub(r'\(.*', '', r['Kernel Name'])) for r in rows]
idx = [i for i, n in enumerate(names) if n == 'adam_kernel']
segs = [(idx[k] + 1, idx[k + 1] + 1) for k in range(len(idx) - 1)]
def seg(a, b, title):
    agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0
    for i in range(a, b):
        agg[names[i]][0] += 1; agg[names[i]][1] += us(rows[i]); tot += us(rows[i])
    print('%s: %d launches, %.0f us (sum of serialised cold-cache launch times)' % (title, b - a, tot))
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 14]:
        print('   %-44s n=%3d tot=%7.1f avg=%7.1f %5.1f%%' % (k[:44], n, t, t / n, 100 * t / tot))
has_lstm = lambda a, b: any('lstm_bwd' in names[i] for i in range(a, b))
crit = [s for s in segs if not has_lstm(*s)]
gen = [s for s in segs if has_lstm(*s)]
if crit: seg(*crit[-1], 'critic step')
if gen: seg(*gen[-1], 'generator step')
