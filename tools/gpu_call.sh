#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
echo "== base"; python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['gpu_launches'], d['roofline']['launch_ms'], d['roofline']['split_ms_per_step'], round(d['roofline']['frac'],3))"
GB=10 bash tools/variant_bench.sh mb4 over512 hops2 fixext sp12 sp10
