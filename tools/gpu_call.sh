#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
BL="python bench.py --steps 2 --warmup 3 --no-cpu --no-encode --no-parity"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_decode$" -s 6 -c 2 -o $O/r2_kdec_10GB -f $BL > $O/ncu_kdec10.log 2>&1; echo "kdec 10GB rc=$?"
ls -la $O/*.ncu-rep; du -sm $O
