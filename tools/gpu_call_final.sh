#!/bin/bash
# Development aid: the round's closing single-GPU call: tests, all five configs with CPU baselines,
# the damage run, launch list and ncu captures (outputs under gpurun_out/, below 64 MiB).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum,l1tex__t_requests_pipe_lsu_mem_global_op_st.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sector_hit_rate.pct,smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct,smsp__sass_average_data_bytes_per_sector_mem_global_op_st.pct"
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/r2_tests.log
python __graft_entry__.py smoke 2>&1 | tail -1
for c in cfg2 cfg1 cfg3 cfg4 cfg5; do
  timeout 1200 python bench.py --config $c --steps 10 --warmup 3 > $O/r2_bench_$c.json 2> $O/r2_bench_$c.err; echo "$c rc=$?"; python tools/bench_brief.py $O/r2_bench_$c.json; tail -2 $O/r2_bench_$c.err
done
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/r2_bench_reference_cfg2.json 2>/dev/null; cut -c1-200 $O/r2_bench_reference_cfg2.json
timeout 900 python bench.py --config cfg2 --gb 2 --damage 1e-6 --steps 5 --no-cpu --no-encode > $O/r2_bench_cfg2_2gb_damage.json 2> $O/r2_bench_cfg2_2gb_damage.err; echo "damage rc=$?"; python tools/bench_brief.py $O/r2_bench_cfg2_2gb_damage.json
timeout 900 python bench.py --config cfg2 --gb 2 --steps 5 --no-cpu --no-encode > $O/r2_bench_cfg2_2gb.json 2> /dev/null; python tools/bench_brief.py $O/r2_bench_cfg2_2gb.json
# every launch of two steps with its device time
BL="python bench.py --steps 2 --warmup 3 --no-cpu --no-encode --no-parity"
$BL > $O/launch_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 130 -c 104 --csv --log-file $O/r2_launches.csv $BL > $O/ncu_launch.log 2>&1; echo "launch list rc=$?"
# the decode kernels at full size, full set (DRAM traffic for roofline.traffic)
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_decode$" -s 6 -c 2 -o $O/r2_kdec_10GB -f $BL > $O/ncu_kdec10.log 2>&1; echo "kdec 10GB rc=$?"
# every kernel family at 1 GB: launch metrics with sector efficiency
python tools/profile_all.py 1 1 > $O/profall_plain.log 2>&1 && \
timeout 900 ncu --metrics $M --clock-control none -k "regex:k_" -c 90 --csv --log-file $O/r2_all_metrics.csv python tools/profile_all.py 1 1 > $O/ncu_all.log 2>&1; echo "all metrics rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_split|k_enc_write" -s 4 -c 5 -o $O/r2_ksplit_kenc -f python tools/profile_all.py 1 1 > $O/ncu_ks.log 2>&1; echo "ksplit/kenc rc=$?"
du -sm $O
