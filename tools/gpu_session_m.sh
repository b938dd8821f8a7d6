#!/bin/bash
# full GPU parity suite, then C2 / C3 / C4 / C4 surface-aware benches (short)
TAG=${1:-s}; O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -6 $O/${TAG}_pytest.log | cut -c1-250
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
    print("%s ms/step %.2f e2e %.2f fwd %.2f adj %.2f plan %d launches %d loss %s"%(sys.argv[2], d["ms_per_step"], d["e2e"]["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["plan"], d["gpu_launches"], d["loss"]))
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
for WL in C2 C3 C4; do
DRTVAM_DEBUG=1 timeout 600 python bench.py --workload $WL --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_$WL.json 2> $O/${TAG}_bench_$WL.err
show $O/${TAG}_bench_$WL.json $WL; grep drtvam $O/${TAG}_bench_$WL.err | tail -1
done
DRTVAM_DEBUG=1 timeout 900 python bench.py --workload C4 --surface-aware --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C4sa.json 2> $O/${TAG}_bench_C4sa.err
show $O/${TAG}_bench_C4sa.json C4sa; grep drtvam $O/${TAG}_bench_C4sa.err | tail -1
