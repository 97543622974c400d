"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.  Usage: ncu_summarize.py launches.csv"""
import csv, sys, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
h = rows[0]
kn, mn, mv, mu = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    if len(r) <= mv or r[mn] != "gpu__time_duration.sum":
        continue
    v = float(r[mv].replace(",", ""))
    us = v / 1e3 if r[mu] in ("ns", "nsecond") else (v if r[mu] in ("us", "usecond") else v * 1e3)
    name = r[kn].split("(")[0]
    tot[name] += us
    cnt[name] += 1
s = sum(tot.values())
for k, v in tot.most_common(24):
    print("%-62s n=%5d total_us=%11.1f share=%.4f mean_us=%9.1f" % (k[:62], cnt[k], v, v / s, v / cnt[k]))
