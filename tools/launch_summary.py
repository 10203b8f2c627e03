"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel: launches, total us, share."""
import collections
import csv
import io
import sys

rows = [l for l in open(sys.argv[1]) if not l.startswith("==")]
r = list(csv.DictReader(io.StringIO("".join(rows))))
agg = collections.OrderedDict()
for x in r:
    k = x["Kernel Name"].split("(")[0]
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += float(x["Metric Value"].replace(",", ""))
tot = sum(a[1] for a in agg.values())
print("| kernel | launches | total us | share |\n|---|---:|---:|---:|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{k}` | {a[0]} | {a[1] / 1e3:.1f} | {100 * a[1] / tot:.1f}% |")
