#!/bin/bash
# HEAD: one --set full capture of the planar kernels (patch-gather forward, gather adjoint) inside the C2 bench step
TAG=${1:-r01z4}; O=gpurun_out; mkdir -p $O
timeout 500 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'planarg|planarz' -c 4 \
    -o $O/${TAG}_planar python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
python tools/ncu_summary.py details $O/${TAG}_planar.ncu-rep > $O/${TAG}_planar_ncu_details.txt 2>&1
grep -A14 "planarg_kernel<4>" $O/${TAG}_planar_ncu_details.txt | grep "dram__bytes\|gpu__time\|issue_active"
rm -f $O/${TAG}_planar.ncu-rep.tmp
