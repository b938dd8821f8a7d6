#!/bin/bash
# surface-aware planar path: parity tests, then C4 surface-aware bench (planar + inside records vs general tracer)
TAG=${1:-s}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_surface_aware.py -m gpu -q -x > $O/${TAG}_pytest_sa.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_sa.log
tail -25 $O/${TAG}_pytest_sa.log | cut -c1-300
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
    print("%s ms/step %.2f e2e %.2f fwd %.2f adj %.2f plan %d launches %d loss %s"%(sys.argv[2], d["ms_per_step"], d["e2e"]["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["plan"], d["gpu_launches"], d["loss"]))
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
DRTVAM_DEBUG=1 timeout 900 python bench.py --workload C4 --surface-aware --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C4sa.json 2> $O/${TAG}_bench_C4sa.err
show $O/${TAG}_bench_C4sa.json C4sa-planar; tail -3 $O/${TAG}_bench_C4sa.err
DRTVAM_SURFACE_AWARE_GENERAL=1 timeout 900 python bench.py --workload C4 --surface-aware --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C4sa_general.json 2> $O/${TAG}_bench_C4sa_general.err
show $O/${TAG}_bench_C4sa_general.json C4sa-general
timeout 600 python bench.py --workload C4 --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C4.json 2> $O/${TAG}_bench_C4.err
show $O/${TAG}_bench_C4.json C4-plain
