#!/bin/bash
# Development aid: the command list of one gpurun call (rewritten per experiment; started with
# tools/run_gpu.sh).
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_golden_and_api.py tests/test_gpu_napi_addon.py -x -q -m gpu 2>&1 | tail -1
timeout 300 python tools/encode_bench.py 2>&1 | tail -2
