#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum,l1tex__t_requests_pipe_lsu_mem_global_op_st.sum,lts__t_sector_hit_rate.pct,smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct,smsp__sass_average_data_bytes_per_sector_mem_global_op_st.pct"
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?"; tail -2 $O/r2_tests.log
python __graft_entry__.py smoke 2>&1 | tail -1
for c in cfg2 cfg3; do
  timeout 900 python bench.py --config $c --steps 10 --warmup 3 > $O/r2_bench_$c.json 2> $O/r2_bench_$c.err; echo "$c rc=$?"; python tools/bench_brief.py $O/r2_bench_$c.json
done
BL="python bench.py --config cfg4 --steps 2 --warmup 1 --no-cpu --no-encode --no-parity"
timeout 600 ncu --metrics $M --clock-control none -k "regex:k_decode|k_long|k_split|k_chunk" -c 10 --csv --log-file $O/r2_cfg4_metrics.csv $BL > $O/ncu_l4.log 2>&1; echo "cfg4 metrics rc=$?"
du -sm $O
