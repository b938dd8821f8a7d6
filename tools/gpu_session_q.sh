#!/bin/bash
# the reference's double-cylindrical known answer + golden discretize, then the other configs' bench lines
TAG=${1:-q}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_optimization.py -m gpu -q -s > $O/${TAG}_pytest_optimization.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_optimization.log
grep -v "^Optimizing\|^Rendering\|^No \|^Pattern eff\|^Saving" $O/${TAG}_pytest_optimization.log | tail -30 | cut -c1-300
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
    print("%s ms/step %.2f e2e %.2f fwd %.2f adj %.2f plan %d launches %d loss %s"%(sys.argv[2], d["ms_per_step"], d["e2e"]["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["plan"], d["gpu_launches"], d["loss"]))
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
for WL in C1 C3 C4; do
timeout 600 python bench.py --workload $WL --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_$WL.json 2> $O/${TAG}_bench_$WL.err
show $O/${TAG}_bench_$WL.json $WL
done
timeout 900 python bench.py --workload C4 --surface-aware --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C4_surface_aware.json 2> $O/${TAG}_bench_C4sa.err
show $O/${TAG}_bench_C4_surface_aware.json C4sa
timeout 900 python bench.py --workload C5 --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C5.json 2> $O/${TAG}_bench_C5.err
show $O/${TAG}_bench_C5.json C5
