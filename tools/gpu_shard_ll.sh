#!/bin/bash
# launch list of one rank's shard (1/8 of the angles) on one GPU
TAG=${1:-w}; O=gpurun_out; mkdir -p $O
python bench.py --emulate-shard 8 --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_c2_shard8.json 2>$O/${TAG}_shard.err
python -c "
import json; d=json.loads(open('$O/${TAG}_bench_c2_shard8.json').read().strip().splitlines()[-1]); r=d['roofline']; print(d['ms_per_step'], r['kernel_ms'], r['adjoint_ms'])"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
    --log-file $O/${TAG}_launches_raw.csv python bench.py --emulate-shard 8 --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_ll.log 2>&1
python tools/ncu_summary.py launches $O/${TAG}_launches_raw.csv > $O/${TAG}_launches_c2_shard8.csv 2>&1; head -30 $O/${TAG}_launches_c2_shard8.csv | cut -c1-160
