#!/bin/bash
cd "$(dirname "$0")/.."
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["gpu_launches"], d["roofline"]["launch_ms"], d["roofline"]["split_ms_per_step"], round(d["roofline"]["frac"],3), d["parity"]["digest_ok"])'
echo "== base"; python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"
GB=10 bash tools/variant_bench.sh sp8 sp9
