#!/bin/bash
# Development aid: time k_decode for alternative builds of the library (tools/libzkbatch_*.so)
for v in "$@"; do
  echo "== $v"
  ZKBATCH_LIB=tools/libzkbatch_$v.so python bench.py --gb 1 --steps 5 --no-cpu 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['launch_ms'], d['roofline']['split_emit_ms_per_step'], round(d['roofline']['frac'],3))"
done
