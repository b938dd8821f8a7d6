#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:planarg -c 2 \
    -o $O/r01t_planarg_tma python bench.py --workload C2 --steps 1 --warmup 3 --no-cpu-baseline > $O/r01l_ncu.log 2>&1
tail -3 $O/r01l_ncu.log
