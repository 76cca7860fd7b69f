"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name (second half of the launches = the measured repetition)."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
body = rows[1:]
if len(sys.argv) > 2 and sys.argv[2] == "half":
    body = body[len(body) // 2:]
agg, cnt = collections.OrderedDict(), collections.Counter()
for r in body:
    v = float(r[vi].replace(",", ""))
    v = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui] in ("us", "usecond") else v
    k = r[ki][:70]
    agg[k] = agg.get(k, 0) + v
    cnt[k] += 1
tot = sum(agg.values())
for k, v in sorted(agg.items(), key=lambda x: -x[1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 16]:
    print("%9.3f ms %5.1f %%  x%-4d %s" % (v, 100 * v / tot, cnt[k], k))
print("%9.3f ms total" % tot)
