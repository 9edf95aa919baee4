#!/bin/bash
# Development aid: the command list of one gpurun call (rewritten per experiment; started with
# tools/run_gpu.sh).  This one: parity tests, a short differential run and the default bench line.
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 600 python tools/fuzz_campaign.py 300000 1000 2>&1 | tail -2
python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 > gpurun_out/bench_line.json
python tools/bench_brief.py gpurun_out/bench_line.json
