#!/bin/bash
# HEAD after the bench fixes: C2 bench (own arm), reference arm under torchrun (rank 0 only), one --set full capture
TAG=${1:-r01z2}; O=gpurun_out; mkdir -p $O
timeout 400 python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench_c2.json 2> $O/${TAG}_bench_c2.err; echo "bench exit $?"
cut -c1-300 $O/${TAG}_bench_c2.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_c2_reference_arm.json 2> $O/${TAG}_ref.err; echo "ref exit $?"
cut -c1-200 $O/${TAG}_bench_c2_reference_arm.json
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'planarg|planarz' -c 4 \
    -o $O/${TAG}_planar python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
python tools/ncu_summary.py details $O/${TAG}_planar.ncu-rep > $O/${TAG}_planar_ncu_details.txt 2>&1
grep -A14 "planarg_kernel<4>" $O/${TAG}_planar_ncu_details.txt | grep "dram__bytes\|gpu__time\|issue_active"
