#!/bin/bash
# ncu --set full of the general tracer kernels on a reduced C5 (scattering, 16 spp) and C2 --general
TAG=${1:-s}
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py --workload C5 --n 160 --steps 1 --warmup 1 --no-cpu-baseline > $O/${TAG}_bench_C5_n160.json 2> $O/${TAG}_bench_C5_n160.err
tail -c 600 $O/${TAG}_bench_C5_n160.json
timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:general_kernel -c 3 \
    -o $O/${TAG}_general_c5 python bench.py --workload C5 --n 160 --steps 1 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu_c5.log 2>&1
tail -3 $O/${TAG}_ncu_c5.log
ls -la $O | grep $TAG
