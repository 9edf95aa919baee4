#!/bin/bash
# Development aid: start one gpurun call detached from the calling shell (a killed shell then does not
# lose the results); the client's output goes to /tmp/w/gpu_call.log, results merge into gpurun_out/.
#   tools/run_gpu.sh <timeout seconds> '<command>' [--gpus N]
mkdir -p /tmp/w
T=$1; CMD=$2; shift 2
setsid nohup /usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD" > /tmp/w/gpu_call.log 2>&1 < /dev/null &
echo started
