"""ncu `--metrics gpu__time_duration.sum --csv` launch list -> per-kernel launches, total time and share (development aid)."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict(); tot = 0.0
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", "")); unit = row["Metric Unit"]
    v = v / 1e6 if unit == "ns" else (v / 1e3 if unit == "us" else v)          # -> ms
    name = re.sub(r"\(.*", "", row["Kernel Name"])
    name = re.sub(r"\b(ntc|bld|lay)::", "", name)[:48]
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v; tot += v
print("%-48s %9s %12s %8s" % ("kernel", "launches", "total ms", "share"))
for k, (n, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-48s %9d %12.3f %7.1f%%" % (k, n, v, 100 * v / tot))
print("%-48s %9d %12.3f" % ("total", sum(n for n, _ in agg.values()), tot))
