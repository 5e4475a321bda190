"""Per-kernel summary of an ncu launch list (--metrics gpu__time_duration.sum --csv): count, average, max, share."""
import collections, csv, sys
for f in sys.argv[1:]:
    rows = [r for r in csv.reader(open(f)) if len(r) > 10]
    hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
    d = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] in ("ns", "nsecond") else (v * 1e3 if r[ui] in ("ms", "msecond") else v)
        d.setdefault(r[ki].split("(")[0][:52], []).append(v)
    tot = sum(sum(v) for v in d.values())
    print(f, "total %.1f us" % tot)
    for k, v in d.items():
        print(f"  {k:54s} n={len(v):4d} avg={sum(v)/len(v):9.2f} us  max={max(v):9.2f}  share={100*sum(v)/tot:5.1f}%")
