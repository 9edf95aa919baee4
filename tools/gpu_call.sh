#!/bin/bash
cd "$(dirname "$0")/.."
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["gpu_launches"], d["roofline"]["launch_ms"], d["roofline"]["split_ms_per_step"], round(d["roofline"]["frac"],3), d["parity"]["digest_ok"])'
timeout 900 python -m pytest tests/test_gpu_decode_parity.py -x -q -m gpu -k "xid" 2>&1 | tail -2
python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"
python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"
python bench.py --config cfg5 --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"
