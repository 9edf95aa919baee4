#!/bin/bash
# Development aid: the command list of one gpurun call (edited per call; outputs under gpurun_out/,
# which must stay below 64 MiB or nothing is copied back).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/r2_tests.log
B="--steps 5 --warmup 3 --no-cpu --no-encode"
for c in cfg2 cfg4 cfg3 cfg1; do
  timeout 600 python bench.py --config $c $B > $O/b_$c.json 2> $O/b_$c.err; echo "$c rc=$?"; python tools/bench_brief.py $O/b_$c.json; tail -2 $O/b_$c.err
done
du -sm $O
