#!/bin/bash
# Development aid: the multi-GPU call (gpurun --gpus 8): N-rank PCIe peaks, then the benches at N = 8.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
nproc > $O/box8.txt; free -g | head -2 >> $O/box8.txt; nvidia-smi topo -m >> $O/box8.txt 2>&1; numactl -H >> $O/box8.txt 2>&1
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29611 tools/pcie_peak.py > $O/pcie_peak_n$n.json 2> $O/pcie_peak_n$n.err; echo "pcie n=$n rc=$?"; tail -1 $O/pcie_peak_n$n.json | cut -c1-300
done
timeout 900 python bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu --no-encode > $O/scale_cfg2_n8.json 2> $O/scale_cfg2_n8.err; echo "cfg2 n8 rc=$?"; python tools/bench_brief.py $O/scale_cfg2_n8.json; tail -2 $O/scale_cfg2_n8.err
timeout 900 python bench.py --gpus 4 --steps 5 --warmup 3 --no-cpu --no-encode > $O/scale_cfg2_n4.json 2> $O/scale_cfg2_n4.err; echo "cfg2 n4 rc=$?"; python tools/bench_brief.py $O/scale_cfg2_n4.json
timeout 1500 python bench.py --gpus 8 --config cfg5 --steps 3 --warmup 3 --no-cpu > $O/scale_cfg5_n8.json 2> $O/scale_cfg5_n8.err; echo "cfg5 n8 rc=$?"; python tools/bench_brief.py $O/scale_cfg5_n8.json; tail -2 $O/scale_cfg5_n8.err
du -sm $O
