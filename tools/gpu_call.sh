#!/bin/bash
cd "$(dirname "$0")/.."
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["gpu_launches"], d["roofline"]["launch_ms"], d["roofline"]["split_ms_per_step"], round(d["roofline"]["frac"],3), d["parity"]["digest_ok"])'
for r in 1 2 4; do echo "== run $r"; ZKBATCH_LIB=tools/libzkbatch_prof.so ZKBATCH_SPLIT_RUN=$r python tools/split_iters.py 2 2>&1 | grep -v Warn | tail -12; done
for r in 1 2 4 8; do echo "== run $r"; ZKBATCH_SPLIT_RUN=$r python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"; done
for c in cfg3 cfg4; do for r in 2 4; do echo "== $c run $r"; ZKBATCH_SPLIT_RUN=$r python bench.py --config $c --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "$P"; done; done
