#!/bin/bash
# Development aid: the multi-GPU call (gpurun --gpus 8).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
B="--gpus 8 --steps 5 --warmup 3 --no-cpu --no-encode"
timeout 900 python bench.py $B > $O/scale_cfg2_n8_compact.json 2> $O/scale_cfg2_n8_compact.err; echo "cfg2 n8 compact rc=$?"; tail -1 $O/scale_cfg2_n8_compact.json > $O/x.json; python tools/bench_brief.py $O/x.json
timeout 900 python bench.py $B --e2e-full-columns > $O/scale_cfg2_n8_full.json 2> $O/scale_cfg2_n8_full.err; echo "cfg2 n8 full rc=$?"; tail -1 $O/scale_cfg2_n8_full.json > $O/x.json; python tools/bench_brief.py $O/x.json
rm -f $O/x.json
