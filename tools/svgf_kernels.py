"""Per-kernel summary of an ncu launch list (csv with gpu__time_duration.sum and friends): one line per kernel name."""
import collections
import csv
import io
import sys

rows = [l for l in open(sys.argv[1]) if not l.startswith("==")]
names = collections.OrderedDict()
for x in csv.DictReader(io.StringIO("".join(rows))):
    k = x["Kernel Name"].split("(")[0]
    a = names.setdefault(k, collections.OrderedDict()).setdefault(x["Metric Name"], [0, 0.0])
    a[0] += 1
    a[1] += float(x["Metric Value"].replace(",", ""))
short = {"gpu__time_duration.sum": "avg_us", "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue%",
         "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram%", "sm__warps_active.avg.pct_of_peak_sustained_active": "occ%",
         "smsp__thread_inst_executed_per_inst_executed.ratio": "lanes"}
for k, v in names.items():
    n = list(v.values())[0][0]
    cols = []
    for m, a in v.items():
        val = a[1] / a[0] / (1e3 if m == "gpu__time_duration.sum" else 1)
        cols.append(f"{short.get(m, m)}={val:7.1f}")
    print(f"{k:44s} n={n:3d} " + " ".join(cols))
