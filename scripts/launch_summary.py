"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches and total time per kernel."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[h]
kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[h + 1:]:
    if len(r) <= mv:
        continue
    try:
        v = float(r[mv].replace(",", ""))
    except ValueError:
        continue
    name = r[kn].split("(")[0][:60]
    tot[name] += v
    cnt[name] += 1
total = sum(v for k, v in tot.items() if "int_peak" not in k)
print("| kernel | launches | total ms | share |")
print("|---|---|---|---|")
for k, v in tot.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 16):
    print("| `%s` | %d | %.1f | %.1f%% |" % (k, cnt[k], v / 1e6, 100 * v / total))
