#!/bin/bash
# parity tests + general-tracer benches (C5 scattering, C2 forced onto the general tracer)
TAG=${1:-s}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -15 $O/${TAG}_pytest.log
timeout 900 python bench.py --workload C5 --steps 2 --warmup 1 --no-cpu-baseline > $O/${TAG}_bench_C5.json 2> $O/${TAG}_bench_C5.err
echo "C5 exit $?"; tail -c 1500 $O/${TAG}_bench_C5.json; tail -5 $O/${TAG}_bench_C5.err
timeout 600 python bench.py --workload C2 --general --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C2_general.json 2> $O/${TAG}_bench_C2_general.err
echo "C2 general exit $?"; tail -c 1200 $O/${TAG}_bench_C2_general.json
