#!/usr/bin/env python
"""Per-kernel share of one headline step from an ncu launch list (ncu --metrics gpu__time_duration.sum --csv): the launches between two
permute_kernel launches are one plan run; the last full-size run is reported. usage: python tools/launch_shares.py profiles/r2_launches.csv"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
data = rows[hdr + 1:]
names = [r[4].split("(")[0].replace("void ", "").replace("<unnamed>::", "") for r in data]
val = [float(r[-1].replace(",", "")) for r in data]
perm = [-1] + [i for i, n in enumerate(names) if "permute" in n]
runs = [(perm[k - 1] + 1, perm[k] + 1) for k in range(1, len(perm))]
full = [r for r in runs if sum(val[r[0]:r[1]]) > 5e6 and not any("dfma_probe" in n for n in names[r[0]:r[1]])]
a, b = full[-1]
agg = collections.OrderedDict()
for n, v in zip(names[a:b], val[a:b]):
    agg.setdefault(n, [0, 0.0]); agg[n][0] += 1; agg[n][1] += v
tot = sum(v for _, v in agg.values())
print(f"plan runs in the list: {len(runs)}, full-size: {len(full)}; last full-size run: {b - a} launches, {tot / 1e6:.3f} ms summed kernel time (serialised, cold cache)")
print("| kernel | launches | ms | share |\n|---|---|---|---|")
for n, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"| `{n}` | {c} | {v / 1e6:.3f} | {100 * v / tot:.1f} % |")
