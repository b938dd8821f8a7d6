#!/bin/bash
# final state of the round: full GPU suite, smoke, C2 bench (both arms), launch list, one --set full capture
TAG=${1:-final}; O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log | cut -c1-250
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench_c2.json 2> $O/${TAG}_bench_c2.err; echo "bench exit $?"
cut -c1-400 $O/${TAG}_bench_c2.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_c2_reference_arm.json 2> $O/${TAG}_ref.err; echo "ref exit $?"
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/${TAG}_launches_raw.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_ll.log 2>&1; echo "ncu launch list exit $?"
python tools/ncu_summary.py launches $O/${TAG}_launches_raw.csv > $O/${TAG}_launches_c2.csv 2>&1; tail -14 $O/${TAG}_launches_c2.csv | cut -c1-200
timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'planarg|planarz' -c 4 \
    -o $O/${TAG}_planar python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
python tools/ncu_summary.py details $O/${TAG}_planar.ncu-rep > $O/${TAG}_planar_ncu_details.txt 2>&1
grep -A14 "planarg_kernel<4>" $O/${TAG}_planar_ncu_details.txt | grep "dram__bytes\|gpu__time\|issue_active" 
