#!/bin/bash
# Development aid: the command list of one gpurun call (edited per call; outputs under gpurun_out/,
# which must stay below 64 MiB or nothing is copied back).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum,l1tex__t_requests_pipe_lsu_mem_global_op_st.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sector_hit_rate.pct,smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct,smsp__sass_average_data_bytes_per_sector_mem_global_op_st.pct"
nproc > $O/box.txt; free -g | head -2 >> $O/box.txt
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/r2_tests.log
timeout 300 python tools/pcie_peak.py > $O/pcie_peak_n1.json 2> $O/pcie_peak_n1.err; cat $O/pcie_peak_n1.json
for c in cfg2 cfg1 cfg3 cfg4; do
  timeout 900 python bench.py --config $c --steps 5 --warmup 3 > $O/r2_bench_$c.json 2> $O/r2_bench_$c.err; echo "$c rc=$?"
  cut -c1-260 $O/r2_bench_$c.json; tail -2 $O/r2_bench_$c.err
done
timeout 900 python bench.py --config cfg2 --gb 2 --damage 1e-6 --steps 5 --no-cpu --no-encode > $O/r2_bench_cfg2_damage.json 2> $O/r2_bench_cfg2_damage.err; echo "damage rc=$?"
timeout 900 python bench.py --config cfg2 --gb 2 --steps 5 --no-cpu --no-encode > $O/r2_bench_cfg2_2gb.json 2> $O/r2_bench_cfg2_2gb.err; echo "2gb rc=$?"
timeout 1200 python bench.py --config cfg5 --steps 3 --warmup 3 > $O/r2_bench_cfg5.json 2> $O/r2_bench_cfg5.err; echo "cfg5 rc=$?"
cut -c1-260 $O/r2_bench_cfg5.json; tail -2 $O/r2_bench_cfg5.err
python tools/profile_all.py 1 1 > $O/r2_profall_plain.log 2>&1 && \
timeout 900 ncu --metrics $M --clock-control none -k "regex:k_" -c 80 --csv --log-file $O/r2_all_metrics.csv python tools/profile_all.py 1 1 > $O/r2_ncu_all.log 2>&1; echo "ncu rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_decode" -c 2 -o $O/r2_kdec -f python tools/profile_all.py 1 1 > $O/r2_ncu_kdec.log 2>&1; echo "ncu2 rc=$?"
du -sm $O
