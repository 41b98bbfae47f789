"""Summarises an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
agg, grid = collections.defaultdict(list), {}
for row in csv.DictReader(lines):
    name = re.sub(r'\(.*', '', row['Kernel Name'])[:100]
    agg[name].append(float(row['Metric Value'])); grid[name] = row['Grid Size']
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print("%-100s n=%4d mean=%8.1f ns share=%.3f grid=%s" % (k, len(v), sum(v) / len(v), sum(v) / tot, grid[k]))
