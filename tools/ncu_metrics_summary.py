"""Development aid: condense the CSV of an `ncu --metrics ... --csv --log-file X.csv` run (one row per
launch and metric) into the per-launch records kept under profiles/.

    python tools/ncu_metrics_summary.py gpurun_out/r2_all_metrics.csv "<how>" [kernel regex] > profiles/r2_xxx.json
"""
import csv
import json
import re
import sys

path, how = sys.argv[1], sys.argv[2]
pat = re.compile(sys.argv[3]) if len(sys.argv) > 3 else None
rows = [r for r in csv.reader(open(path)) if len(r) > 5]
hdr = rows[0]
iid, ik, im, iu, iv = (hdr.index(x) for x in ("ID", "Kernel Name", "Metric Name", "Metric Unit", "Metric Value"))
per = {}
for r in rows[1:]:
    name = re.sub(r"^void ", "", r[ik]).replace("<unnamed>::", "")
    name = re.sub(r"\(.*$", "", name).replace("(int)", "").replace("(bool)", "")
    d = per.setdefault(int(r[iid]), {"kernel": name})
    try:
        d[r[im]] = (float(r[iv].replace(",", "")), r[iu])
    except ValueError:
        pass


def val(d, m, unit=None):
    if m not in d:
        return None
    v, u = d[m]
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
    if unit in ("us", "MB"):
        return v * scale.get(u, 1.0)
    return v


out = []
for i in sorted(per):
    d = per[i]
    if pat and not pat.search(d["kernel"]):
        continue
    rec = {"id": i, "kernel": d["kernel"]}
    rec["duration_us"] = round(val(d, "gpu__time_duration.sum", "us") or 0, 1)
    rec["dram_read_MB"] = round(val(d, "dram__bytes_read.sum", "MB") or 0, 2)
    rec["dram_write_MB"] = round(val(d, "dram__bytes_write.sum", "MB") or 0, 2)
    for m, k in (("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct_of_peak"),
                 ("smsp__inst_executed.sum", "warp_inst"),
                 ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads_per_inst"),
                 ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
                 ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
                 ("launch__registers_per_thread", "registers"),
                 ("smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct", "ld_bytes_used_per_sector_pct"),
                 ("smsp__sass_average_data_bytes_per_sector_mem_global_op_st.pct", "st_bytes_used_per_sector_pct"),
                 ("lts__t_sector_hit_rate.pct", "l2_hit_pct")):
        v = val(d, m)
        if v is not None:
            rec[k] = round(v, 2)
    for a, b, k in (("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "ld_sectors_per_request"),
                    ("l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "st_sectors_per_request")):
        s, q = val(d, a), val(d, b)
        if s is not None and q:
            rec[k] = round(s / q, 2)
    out.append(rec)
print(json.dumps({"how": how, "launches": out}, indent=1))
