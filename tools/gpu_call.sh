#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/r2_tests.log
B="--steps 5 --warmup 3 --no-cpu --no-encode"
timeout 600 python bench.py $B > $O/b_cfg2.json 2> $O/b_cfg2.err; echo "cfg2 rc=$?"; python tools/bench_brief.py $O/b_cfg2.json; tail -2 $O/b_cfg2.err
timeout 600 python bench.py $B --e2e-full-columns > $O/b_cfg2_fullcols.json 2> $O/b_cfg2_fullcols.err; echo "cfg2 full rc=$?"; python tools/bench_brief.py $O/b_cfg2_fullcols.json
python -c "
import json
for f in ('b_cfg2','b_cfg2_fullcols'):
    l=json.load(open('$O/'+f+'.json')); e=l['e2e']; print(f, e['ms_per_step'], e['d2h_bytes_per_step'], e['h2d_bytes_per_step'], e['pcie_roofline'])"
du -sm $O
