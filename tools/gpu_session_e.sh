#!/bin/bash
TAG=${1:-s}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -40 $O/${TAG}_pytest.log | cut -c1-300
timeout 900 python bench.py --workload C4 --surface-aware --steps 3 --warmup 3 --cpu-seconds 5 > $O/${TAG}_bench_C4_sa.json 2> $O/${TAG}_bench_C4_sa.err
echo "C4 SA exit $?"; tail -c 1500 $O/${TAG}_bench_C4_sa.json; tail -5 $O/${TAG}_bench_C4_sa.err
