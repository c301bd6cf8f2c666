"""Aggregate an ncu --csv launch list (gpu__time_duration.sum): per kernel count / mean ms / share.
usage: ncu_launches.py launches.csv [steps]"""
import csv, sys, collections
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
agg = collections.OrderedDict()
for x in csv.DictReader(lines):
    if x.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    v = float(x['Metric Value'].replace(',', ''))
    v *= {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(x['Metric Unit'], 1.0)
    agg.setdefault(x['Kernel Name'].split('(')[0][:70], []).append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in agg.items():
    print('%-70s n=%4d mean %8.4f ms  per-step %8.4f ms  share %5.1f%%' % (k, len(v), sum(v) / len(v), sum(v) / steps, 100 * sum(v) / tot))
print('total per step %.4f ms' % (tot / steps))
