#!/bin/bash
TAG=${1:-x}; O=gpurun_out; mkdir -p $O
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'planarg_kernel' -c 2 \
    -o $O/${TAG}_planarg_shard8 python bench.py --emulate-shard 8 --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1; echo "ncu exit $?"
python tools/ncu_summary.py details $O/${TAG}_planarg_shard8.ncu-rep > $O/${TAG}_planarg_shard8_ncu_details.txt 2>&1; head -30 $O/${TAG}_planarg_shard8_ncu_details.txt | cut -c1-160
