#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 900 python tools/fuzz_campaign.py 300000 6000 > $O/fuzz_campaign.log 2>&1; echo "fuzz rc=$?"; tail -5 $O/fuzz_campaign.log
