#!/bin/bash
O=gpurun_out; mkdir -p $O
export DRTVAM_PERSIST_BLOCKS=3 DRTVAM_PERSIST_SETUP_MIN=16 DRTVAM_PERSIST_MAX_WAIT=6
timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:general_persistent -c 2 \
    -o $O/r01h_persist_c5 python bench.py --workload C5 --n 160 --steps 1 --warmup 1 --no-cpu-baseline > $O/r01h_ncu.log 2>&1
tail -3 $O/r01h_ncu.log
