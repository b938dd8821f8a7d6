#!/bin/bash
# Bench lines of the other BASELINE.json configs (parity-test cases, reported for reference).
TAG=${1:-s}
O=gpurun_out
mkdir -p $O
for WL in C1 C3 C4; do
  timeout 300 python bench.py --workload $WL --steps 5 --warmup 3 --cpu-seconds 5 > $O/${TAG}_bench_$WL.json 2> $O/${TAG}_bench_$WL.err
  echo "$WL exit $?"; tail -c 1200 $O/${TAG}_bench_$WL.json
done
timeout 900 python bench.py --workload C5 --steps 2 --warmup 1 --cpu-seconds 10 > $O/${TAG}_bench_C5.json 2> $O/${TAG}_bench_C5.err
echo "C5 exit $?"; tail -c 1500 $O/${TAG}_bench_C5.json; tail -5 $O/${TAG}_bench_C5.err
timeout 300 python bench.py --workload C2 --jitter --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C2_jitter.json 2> $O/${TAG}_bench_C2_jitter.err
echo "C2 jitter exit $?"; tail -c 1200 $O/${TAG}_bench_C2_jitter.json
