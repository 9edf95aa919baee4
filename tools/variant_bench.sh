#!/bin/bash
# Development aid: step time for alternative builds of the library (tools/libzkbatch_*.so)
GB=${GB:-1}
for v in "$@"; do
  echo "== $v"
  ZKBATCH_LIB=tools/libzkbatch_$v.so python bench.py --gb $GB --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['gpu_launches'], d['roofline']['launch_ms'], d['roofline']['split_ms_per_step'], round(d['roofline']['frac'],3))"
done
