#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
echo "== new"; python bench.py --steps 5 --no-cpu --no-encode 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['gpu_launches'], d['roofline']['launch_ms'], d['roofline']['split_ms_per_step'], round(d['roofline']['frac'],3), d['parity']['digest_ok'])"
python tools/profile_all.py 1 1 > $O/profall_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_decode" -c 2 -o $O/r2_kdec_1GB_b -f python tools/profile_all.py 1 1 > $O/ncu_kd.log 2>&1; echo "kdec rc=$?"
du -sm $O
