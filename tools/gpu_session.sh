#!/bin/bash
# One gpurun session: GPU parity tests, the default bench, the reference arm, then
# (only after those exit 0) the ncu launch list and one --set full capture.
# Usage (from the repo root, under gpurun):  bash tools/gpu_session.sh TAG [WORKLOAD]
TAG=${1:-s}
WL=${2:-C2}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $O/${TAG}_gpu.txt 2>&1
nproc >> $O/${TAG}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
timeout 600 python bench.py --workload $WL > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; rc=$?
echo "bench exit $rc"; tail -c 3000 $O/${TAG}_bench.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 --workload $WL > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err
tail -c 1500 $O/${TAG}_bench_ref.json
if [ $rc -eq 0 ]; then
  timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/${TAG}_launches.csv python bench.py --workload $WL --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu1.log 2>&1
  timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:planar[gz]_kernel -c 6 \
    -o $O/${TAG}_planar python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu2.log 2>&1
  ls -la $O
fi
