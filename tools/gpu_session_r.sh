#!/bin/bash
# full GPU suite + C2 / C3 bench (kernel experiments)
TAG=${1:-r}; O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -4 $O/${TAG}_pytest.log | cut -c1-250
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
    print("%s ms/step %.2f e2e %.2f fwd %.2f adj %.2f plan %d launches %d loss %s"%(sys.argv[2], d["ms_per_step"], d["e2e"]["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["plan"], d["gpu_launches"], d["loss"]))
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
for WL in C2 C3; do
timeout 600 python bench.py --workload $WL --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_$WL.json 2> $O/${TAG}_bench_$WL.err
show $O/${TAG}_bench_$WL.json $WL
done
