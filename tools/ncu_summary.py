"""Development aid: condense `ncu -i X.ncu-rep --page raw --csv` into one small JSON object per
kernel launch (the numbers profiles/*.json keep): duration, DRAM bytes, issue activity, active
threads per instruction, and the sector-efficiency counters of global loads/stores.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > /tmp/raw.csv
    python tools/ncu_summary.py /tmp/raw.csv [kernel regex] > profiles/r2_xxx.json
"""
import csv
import json
import re
import sys

WANT = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct_of_peak",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct_of_peak",
    "smsp__inst_executed.sum": "warp_inst",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_inst",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__shared_mem_per_block_static": "smem_static",
    "launch__occupancy_limit_registers": "occ_limit_regs",
    "launch__occupancy_limit_shared_mem": "occ_limit_smem",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum": "ld_sectors",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum": "ld_requests",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum": "st_sectors",
    "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum": "st_requests",
    "lts__t_sectors_op_read.sum": "l2_sectors_read",
    "lts__t_sectors_op_write.sum": "l2_sectors_write",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct": "ld_bytes_per_sector_pct",
    "smsp__sass_average_data_bytes_per_sector_mem_global_op_st.pct": "st_bytes_per_sector_pct",
}
rows = list(csv.reader(open(sys.argv[1])))
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
names, units = rows[hdr], rows[hdr + 1]
kcol = names.index("Kernel Name")
out = []
for r in rows[hdr + 2:]:
    if len(r) != len(names) or (pat and not pat.search(r[kcol])):
        continue
    o = {"kernel": re.sub(r"<unnamed>::|\(<unnamed>::\w+\)|void ", "", r[kcol])[:60], "id": int(r[0])}
    for i, n in enumerate(names):
        if n in WANT and r[i] != "":
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                continue
            o[WANT[n]] = v
            if n == "gpu__time_duration.sum":
                o["duration_unit"] = units[i]
            if n.startswith("dram__bytes"):
                o[WANT[n] + "_unit"] = units[i]
    if o.get("ld_requests"):
        o["ld_sectors_per_request"] = round(o["ld_sectors"] / o["ld_requests"], 2)
    if o.get("st_requests"):
        o["st_sectors_per_request"] = round(o["st_sectors"] / o["st_requests"], 2)
    out.append(o)
print(json.dumps(out, indent=1))
