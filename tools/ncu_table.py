"""Per-kernel table from an `ncu --csv --metrics ...` log (one row per launch)."""
import csv, collections, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr, rows = rows[0], rows[1:]
ix = {n: i for i, n in enumerate(hdr)}
K = collections.OrderedDict()
for r in rows:
    k = (r[ix['ID']], r[ix['Kernel Name']].replace('void ', '').replace('<unnamed>::', '')[:22])
    K.setdefault(k, {})[r[ix['Metric Name']]] = float(r[ix['Metric Value']].replace(',', ''))
cols = sys.argv[2:] or sorted(next(iter(K.values())))
short = lambda c: c.replace('l1tex__', '').replace('.sum', '').replace('smsp__', '').replace('.avg.pct_of_peak_sustained', '%')[:20]
print(f"{'kernel':24s}" + "".join(f"{short(c):>21s}" for c in cols))
tot = collections.Counter()
for k, m in K.items():
    print(f"{k[1]:24s}" + "".join(f"{m.get(c, 0) / (1 if '%' in short(c) or 'pct' in c else 1e6):21.3f}" for c in cols))
    for c in cols: tot[c] += m.get(c, 0)
print(f"{'total':24s}" + "".join(f"{tot[c] / 1e6:21.3f}" for c in cols))
