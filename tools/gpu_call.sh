#!/bin/bash
cd "$(dirname "$0")/.."
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["gpu_launches"], d["roofline"]["launch_ms"], d["roofline"]["split_ms_per_step"], round(d["roofline"]["frac"],3), d["parity"]["digest_ok"])'
timeout 900 python -m pytest tests/test_gpu_decode_parity.py -x -q -m gpu 2>&1 | tail -1
timeout 300 python tools/fuzz_campaign.py 1300000 600 2>&1 | tail -1
python bench.py --config cfg3 --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"
python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"
