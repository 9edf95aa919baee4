"""Development aid: per-kernel totals from an `ncu --metrics gpu__time_duration.sum --csv` launch list.

    python tools/launch_summary.py gpurun_out/launches.csv "<command line that was profiled>" > profiles/rN_launches_summary.md
"""
import collections
import csv
import re
import sys

path, cmd = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
hdr = rows[0]
kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= mv:
        continue
    name = re.sub(r"\(.*", "", r[kn]).replace("<unnamed>::", "").replace("(anonymous namespace)::", "")
    v = float(r[mv].replace(",", ""))
    unit = r[mu]
    us = v / 1000.0 if unit in ("ns", "nsecond") else v * 1000.0 if unit in ("ms", "msecond") else v
    t = tot.setdefault(name, [0, 0.0])
    t[0] += 1
    t[1] += us
allus = sum(t[1] for t in tot.values())
print("# ncu launch list summary: `%s`" % cmd)
print("# per-launch times under ncu are cold-cache and serialised: compare shares, not absolutes\n")
print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
for name, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("| %s | %d | %.0f | %.1f | %.1f %% |" % (name, n, us, us / n, 100.0 * us / allus))
