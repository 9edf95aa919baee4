#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/r2_tests.log
python tools/encode_bench.py 320 20000 > $O/encode_bench.json 2>&1; cat $O/encode_bench.json | tail -1 | cut -c1-300
du -sm $O
