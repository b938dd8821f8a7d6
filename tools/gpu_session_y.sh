#!/bin/bash
TAG=${1:-y}; O=gpurun_out; mkdir -p $O
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
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C2.json 2> $O/${TAG}_bench_C2.err; show $O/${TAG}_bench_C2.json C2
python bench.py --emulate-shard 8 --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C2_shard8.json 2> $O/${TAG}_bench_C2s.err; show $O/${TAG}_bench_C2_shard8.json C2-shard8
python bench.py --workload C4 --surface-aware --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C4sa.json 2> $O/${TAG}_bench_C4sa.err; show $O/${TAG}_bench_C4sa.json C4sa
