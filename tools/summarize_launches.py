"""Summarise an ncu --csv launch list (gpu__time_duration.sum) by kernel and grid."""
import csv, collections, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", "")) / 1e3
    key = (re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("pgn::", "").replace("(anonymous namespace)::", ""), row["Grid Size"])
    agg[key][0] += 1; agg[key][1] += v; tot += v
print("total %.1f ms over %d launches" % (tot / 1e3, sum(n for n, _ in agg.values())))
print("| kernel | grid | launches | total us | avg us | share |\n|---|---|---:|---:|---:|---:|")
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("| %s | %s | %d | %.1f | %.2f | %.1f%% |" % (k[0], k[1], n, t, t / n, 100 * t / tot))
