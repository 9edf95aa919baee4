#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/r2_tests.log
for c in cfg1 cfg3 cfg4; do
  timeout 900 python bench.py --config $c --steps 10 --warmup 3 > $O/r2_bench_$c.json 2> $O/r2_bench_$c.err; echo "$c rc=$?"; python tools/bench_brief.py $O/r2_bench_$c.json
done
