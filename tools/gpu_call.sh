#!/bin/bash
# Development aid: the command list of one gpurun call (edited per call; outputs under gpurun_out/).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
nproc > $O/box.txt; free -g | head -2 >> $O/box.txt
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?" | tee -a $O/r2_tests.log
timeout 300 python tools/pcie_peak.py > $O/pcie_peak_n1.json 2> $O/pcie_peak_n1.err
for c in cfg2 cfg1 cfg3 cfg4; do
  timeout 900 python bench.py --config $c --steps 5 --warmup 3 > $O/r2_bench_$c.json 2> $O/r2_bench_$c.err; echo "$c rc=$?"
done
timeout 900 python bench.py --config cfg2 --gb 2 --damage 1e-6 --steps 5 --no-cpu --no-encode > $O/r2_bench_cfg2_damage.json 2> $O/r2_bench_cfg2_damage.err; echo "damage rc=$?"
timeout 1200 python bench.py --config cfg5 --steps 3 --warmup 3 > $O/r2_bench_cfg5.json 2> $O/r2_bench_cfg5.err; echo "cfg5 rc=$?"
python tools/profile_all.py 1 1 > $O/r2_profall_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_' -c 150 -o $O/r2_all_base -f python tools/profile_all.py 1 1 > $O/r2_ncu_all.log 2>&1; echo "ncu rc=$?"
tail -3 $O/r2_tests.log
