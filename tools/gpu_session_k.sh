#!/bin/bash
TAG=${1:-s}; O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -6 $O/${TAG}_pytest.log | cut -c1-250
for WL in C2 C3 C1; do
timeout 600 python bench.py --workload $WL --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_$WL.json 2> $O/${TAG}_bench_$WL.err
python - <<PY
import json
try:
    d=json.loads(open("$O/${TAG}_bench_$WL.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("$WL ms/step %.2f e2e %.2f fwd %.2f adj %.2f plan %d loss %s"%(d["ms_per_step"], d["e2e"]["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["plan"], d["loss"]))
except Exception as e: print("$WL ERR", e)
PY
tail -3 $O/${TAG}_bench_$WL.err
done
DRTVAM_PLANAR_FORWARD=scatter timeout 600 python bench.py --workload C2 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C2 scatter-forward ms/step %.2f fwd %.2f loss %s'%(d['ms_per_step'], d['roofline']['kernel_ms'], d['loss']))"
