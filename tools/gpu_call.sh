#!/bin/bash
cd "$(dirname "$0")/.."
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["gpu_launches"], d["roofline"]["launch_ms"], d["roofline"]["split_ms_per_step"], round(d["roofline"]["frac"],3), d["parity"]["digest_ok"])'
for g in 32 128; do echo "== L2 fetch $g"; ZKBATCH_L2_FETCH=$g python bench.py --steps 5 --no-cpu --no-encode 2>/tmp/err.log | tail -1 | python -c "$P"; grep -m1 "L2 fetch" /tmp/err.log; done
