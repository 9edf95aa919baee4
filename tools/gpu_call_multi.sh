#!/bin/bash
# Development aid: the multi-GPU call (gpurun --gpus N): bench.py re-executes itself under torch.distributed.run.
#   tools/gpu_call_multi.sh <N> <config> [more bench.py switches]
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
N=$1; C=$2; shift 2
timeout 900 python bench.py --gpus $N --config $C --steps 5 --warmup 3 --no-cpu "$@" > $O/scale_${C}_n$N.json 2> $O/scale_${C}_n$N.err; echo "$C n$N rc=$?"
tail -1 $O/scale_${C}_n$N.json > $O/x.json; python tools/bench_brief.py $O/x.json; rm -f $O/x.json
