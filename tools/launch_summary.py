"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
python tools/launch_summary.py launches.csv ["title line"]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ix = {k: i for i, k in enumerate(hdr)}
tot, cnt, mx = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows:
    if r is hdr or len(r) != len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(r[ix["Metric Value"]].replace(",", ""))
    u = r[ix["Metric Unit"]]
    ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
    name = r[ix["Kernel Name"]].split("(")[0]
    tot[name] += ms; cnt[name] += 1; mx[name] = max(mx[name], ms)
if len(sys.argv) > 2:
    print(sys.argv[2])
all_ms = sum(tot.values())
print("%d launches, %.3f ms of kernel time (cold-cache, serialised)" % (sum(cnt.values()), all_ms))
for name, ms in tot.most_common():
    print("%-44s n=%4d total %10.3f ms  max %9.3f ms  %5.1f%%" % (name[:44], cnt[name], ms, mx[name], 100 * ms / all_ms))
