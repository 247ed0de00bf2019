"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
hdr = rows[hi]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0]); seq = []
for r in rows[hi + 1:]:
    if len(r) <= vi: continue
    v = float(r[vi].replace(',', '')); u = r[ui]
    v = v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
    short = re.sub(r'\(.*', '', r[ki]); short = re.sub(r'void |drn::<unnamed>::|at::native::', '', short)[:64]
    agg[short][0] += 1; agg[short][1] += v; seq.append((short, v))
tot = sum(v for _, v in agg.values())
print(f"total {tot:.1f} us over {len(seq)} launches")
for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 22]:
    print(f"{v:10.1f} us {100 * v / tot:5.1f}% n={n:4d} avg {v / n:8.1f}  {k}")
if len(sys.argv) > 3:
    for s, v in seq:
        if sys.argv[3] in s: print(f"{v:9.1f} {s}")
