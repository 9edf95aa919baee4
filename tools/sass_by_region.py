"""Development aid: executed warp instructions / stall samples of one kernel launch, grouped by
*region* of a top-level function (line ranges inside it), with inlined callees charged to the
line of the call site.

    python tools/sass_by_region.py <ncu --page source --csv export> <nvdisasm -gi dump> <kernel substr> \
        <launch index> <source file> <first>-<last> [name=first-last ...]

Instructions whose inline chain has a line inside [first, last] of <source file> are charged to
that line (the outermost one inside the range); the named sub-ranges are summed up, everything
else is listed by its own line.  The nvdisasm dump must come from the same build as the capture.
"""
import collections
import csv
import re
import sys

csv_path, sass_path, kname, which, src_path, span = sys.argv[1:7]
which = int(which)
lo_all, hi_all = [int(x) for x in span.split("-")]
regions = []
for spec in sys.argv[7:]:
    name, r = spec.split("=")
    a, b = r.split("-")
    regions.append((name, int(a), int(b)))
base = src_path.split("/")[-1]

lines = open(sass_path).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and kname in l)
ins = []          # per instruction: (line charged, innermost line)
chain = []
fresh = False
for l in lines[start + 1:]:
    if l.startswith("\t.section") or l.startswith(".text."):
        break
    m = re.search(r'//## File "(.*?)", line (\d+)', l)
    if m:
        if not fresh:          # an annotation block replaces the previous one; none = same lines as before
            chain = []
            fresh = True
        chain.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        charged = None
        for f, n in chain:           # innermost first; keep the outermost inside the span
            if f == base and lo_all <= n <= hi_all:
                charged = n
        inner = chain[0] if chain else ("?", 0)
        ins.append((charged, inner))
        fresh = False

kernels = []
rows = list(csv.reader(open(csv_path)))
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        hdr = rows[i + 1]
        j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if rows[j]:
                body.append(rows[j])
            j += 1
        kernels.append((rows[i][1], hdr, body))
        i = j
    else:
        i += 1
name, hdr, body = kernels[which]
print("kernel", name, "sass rows", len(body), "nvdisasm instructions", len(ins))
ci = hdr.index("Instructions Executed")
ct = hdr.index("Thread Instructions Executed")
cs = hdr.index("# Samples")
n = min(len(body), len(ins))
per_reg = collections.defaultdict(lambda: [0, 0, 0])
per_line = collections.defaultdict(lambda: [0, 0, 0])
for k in range(n):
    charged, inner = ins[k]
    v = (int(body[k][ci] or 0), int(body[k][ct] or 0), int(body[k][cs] or 0))
    reg = "(outside)"
    if charged is not None:
        reg = "(other lines)"
        for nm, a, b in regions:
            if a <= charged <= b:
                reg = nm
                break
    for x in range(3):
        per_reg[reg][x] += v[x]
        per_line[charged if charged is not None else inner][x] += v[x]
tot = [sum(p[x] for p in per_reg.values()) for x in range(3)]
print("total warp inst %d, thread inst %d (%.1f/32), samples %d" % (tot[0], tot[1], tot[1] / max(tot[0], 1), tot[2]))
print("%-28s %7s %7s %5s" % ("region", "inst%", "smpl%", "thr"))
order = [r[0] for r in regions] + ["(other lines)", "(outside)"]
for nm in order:
    p = per_reg.get(nm)
    if p:
        print("%-28s %6.2f%% %6.2f%% %5.1f" % (nm, 100.0 * p[0] / tot[0], 100.0 * p[2] / max(tot[2], 1), p[1] / max(p[0], 1)))
src = open(src_path).read().split("\n")
print("\ntop call-site lines")
for ln, p in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:45]:
    if isinstance(ln, tuple):
        text = "%s:%d" % ln
        label = "-"
    else:
        label, text = str(ln), (src[ln - 1].strip()[:100] if 0 < ln <= len(src) else "")
    print("%6s %6.2f%% %6.2f%% %5.1f  %s" % (label, 100.0 * p[0] / tot[0], 100.0 * p[2] / max(tot[2], 1), p[1] / max(p[0], 1), text))
