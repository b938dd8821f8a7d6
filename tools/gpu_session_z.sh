#!/bin/bash
# re-entry check of HEAD: full GPU suite, smoke, C2 bench (both arms), launch list
TAG=${1:-r01z}; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log | cut -c1-250
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 400 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench_c2.json 2> $O/${TAG}_bench_c2.err; echo "bench exit $?"
cut -c1-400 $O/${TAG}_bench_c2.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_c2_reference_arm.json 2> $O/${TAG}_ref.err; echo "ref exit $?"
timeout 300 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/${TAG}_launches_raw.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_ll.log 2>&1; echo "ncu launch list exit $?"
python tools/ncu_summary.py launches $O/${TAG}_launches_raw.csv > $O/${TAG}_launches_c2.csv 2>&1; tail -14 $O/${TAG}_launches_c2.csv | cut -c1-200
