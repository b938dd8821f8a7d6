#!/bin/bash
# two-phase queue tracer for scattering: full GPU suite, then C5 with the queue and with the nested kernels
TAG=${1:-u}; O=gpurun_out; mkdir -p $O
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
timeout 900 python bench.py --workload C5 --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C5.json 2> $O/${TAG}_bench_C5.err
show $O/${TAG}_bench_C5.json C5-queue
DRTVAM_GENERAL_KERNELS=nested timeout 900 python bench.py --workload C5 --n 160 --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C5_n160_nested.json 2> $O/${TAG}_bench_C5b.err
show $O/${TAG}_bench_C5_n160_nested.json C5-n160-nested
timeout 900 python bench.py --workload C5 --n 160 --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_C5_n160_queue.json 2> $O/${TAG}_bench_C5c.err
show $O/${TAG}_bench_C5_n160_queue.json C5-n160-queue
