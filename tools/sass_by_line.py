"""Development aid: join an `ncu --page source --print-source sass --csv` export with the
line table from `nvdisasm -g` to get executed warp instructions / stall samples per source
line of one kernel.

    python tools/sass_by_line.py gpurun_out/kdec_sass.csv /tmp/cub/dec.sass k_decode [launch_index]
"""
import collections
import csv
import re
import sys

csv_path, sass_path, kname = sys.argv[1:4]
which = int(sys.argv[4]) if len(sys.argv) > 4 else 0

# --- nvdisasm: per-instruction (line, inlined-at chain head) in order -------------------
lines = open(sass_path).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and kname in l)
ins = []
cur = 0
for l in lines[start + 1:]:
    if l.startswith("\t.section") or l.startswith(".text."):
        break
    m = re.search(r'//## File "(.*?)", line (\d+)', l)
    if m:
        f = m.group(1)
        cur = int(m.group(2)) if f.endswith("zk_decode.cu") else (f.split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        ins.append(cur)

# --- ncu csv: kernels separated by "Kernel Name" rows -----------------------------------
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
per = collections.defaultdict(lambda: [0, 0, 0])
for k in range(n):
    p = per[ins[k]]
    p[0] += int(body[k][ci] or 0)
    p[1] += int(body[k][ct] or 0)
    p[2] += int(body[k][cs] or 0)
tot = [sum(p[x] for p in per.values()) for x in range(3)]
print("total warp inst %d, thread inst %d (%.1f/32), samples %d" % (tot[0], tot[1], tot[1] / max(tot[0], 1), tot[2]))
src = open("node-zkstream_b200/csrc/zk_decode.cu").read().split("\n")
print("%6s %7s %7s %5s  %s" % ("line", "inst%", "smpl%", "thr", "source"))
for ln, p in sorted(per.items(), key=lambda kv: -kv[1][0])[:70]:
    if isinstance(ln, tuple):
        label, text = "hdr", "%s:%d" % ln
    else:
        label, text = str(ln), (src[ln - 1].strip()[:110] if 0 < ln <= len(src) else "")
    print("%6s %6.2f%% %6.2f%% %5.1f  %s" % (label, 100.0 * p[0] / tot[0], 100.0 * p[2] / max(tot[2], 1), p[1] / max(p[0], 1), text))
