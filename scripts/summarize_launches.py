import csv, re, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r'\(.*', '', r[4]).replace('void ', '').replace('vk::', '').replace('(anonymous namespace)::', '')[:70]
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += float(r[-1])
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print("%-72s n=%4d total %9.3f ms  avg %9.3f us  %5.1f%%" % (k, v[0], v[1] / 1e6, v[1] / v[0] / 1e3, 100 * v[1] / tot))
