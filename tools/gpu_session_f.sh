#!/bin/bash
# parity tests + general-tracer benches (C5 scattering, C2 forced general, C4 surface-aware)
TAG=${1:-s}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -30 $O/${TAG}_pytest.log | cut -c1-250
for spec in "C5:--steps 2 --warmup 1" "C2:--general --steps 3 --warmup 3" "C4:--surface-aware --steps 3 --warmup 3" "C2:--jitter --general --steps 3 --warmup 3"; do
  WL=${spec%%:*}; FL=${spec#*:}; NAME=$(echo "${WL}_${FL}" | tr -d ' -' | cut -c1-30)
  timeout 900 python bench.py --workload $WL $FL --no-cpu-baseline > $O/${TAG}_bench_$NAME.json 2> $O/${TAG}_bench_$NAME.err
  echo "$WL $FL exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/${TAG}_bench_$NAME.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("   ms/step %.1f fwd %.1f adj %.1f plan %d loss %s"%(d["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["plan"], d["loss"]))
except Exception as e: print("   ERR", e)
PY
  tail -3 $O/${TAG}_bench_$NAME.err
done
