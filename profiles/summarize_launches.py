import csv, sys, re, collections
for f in sys.argv[1:]:
    rows = list(csv.reader(open(f)))
    # find header
    hi = next(i for i,r in enumerate(rows) if 'Kernel Name' in r)
    hdr = rows[hi]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
    agg = collections.OrderedDict()
    for r in rows[hi+1:]:
        if len(r) <= vi: continue
        name = re.sub(r'\(.*', '', r[ki])
        name = re.sub(r'^void ', '', name)
        v = float(r[vi].replace(',', ''))
        if r[ui] == 'ns': v /= 1000.0
        elif r[ui] == 'ms': v *= 1000.0
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"== {f}: {sum(a[0] for a in agg.values())} launches, {tot:.1f} us total")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"  {t:9.1f} us {100*t/tot:5.1f}%  n={c:4d} avg={t/c:8.2f} us  {n[:90]}")
