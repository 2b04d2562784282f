"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: python tools/launch_summary.py launches.csv[.gz] [skip_first_n_launches] [--only REGEX]
       --only: aggregate only the kernels whose name matches (e.g. the kernels of the visibility step,
       so that a kernel's share of the STEP can be read off a launch list of the whole bench command)"""
import collections, csv, gzip, io, re, sys
only = None
if "--only" in sys.argv:
    i = sys.argv.index("--only")
    only = re.compile(sys.argv[i + 1])
    del sys.argv[i:i + 2]
raw = gzip.open(sys.argv[1], "rt").read() if sys.argv[1].endswith(".gz") else open(sys.argv[1]).read()
rows = list(csv.reader(io.StringIO(raw)))
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hi]; idx = {k: i for i, k in enumerate(h)}
agg = collections.OrderedDict(); total = 0.0; n = 0
for r in rows[hi + 1:]:
    if len(r) < len(h) or r[idx["Metric Name"]] != "gpu__time_duration.sum":
        continue
    if only is not None and not only.search(r[idx["Kernel Name"]]):
        continue
    n += 1
    if n <= skip:
        continue
    v = float(r[idx["Metric Value"]].replace(",", "")); u = r[idx["Metric Unit"]]
    v = v / 1e6 if u == "ns" else v / 1e3 if u in ("us", "usecond") else v
    k = r[idx["Kernel Name"]][:86]
    agg.setdefault(k, [0, 0.0]); agg[k][0] += 1; agg[k][1] += v; total += v
print("launches %d (skipped first %d), total kernel time %.3f ms (cold-cache, serialised under ncu)" % (n - skip, skip, total))
print("%6s %10s %7s  kernel" % ("count", "ms", "share"))
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%6d %10.3f %6.1f%%  %s" % (c, t, 100 * t / total, k))
