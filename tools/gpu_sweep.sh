#!/bin/bash
# tuning sweep of the persistent general tracer on reduced workloads
O=gpurun_out; mkdir -p $O
run() {  # label, env..., -- bench args
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  out=$(env "${envs[@]}" timeout 300 python bench.py "$@" --no-cpu-baseline 2>/dev/null | tail -1)
  python - "$label" <<PY
import json,sys
try:
    d=json.loads('''$out'''); r=d["roofline"]
    print("%-44s ms/step %8.1f fwd %8.1f adj %8.1f"%(sys.argv[1], d["ms_per_step"], r["kernel_ms"], r["adjoint_ms"]))
except Exception as e: print(sys.argv[1], "ERR", e)
PY
}
for W in "C5 --n 160 --steps 1 --warmup 1" "C2 --n 256 --general --steps 2 --warmup 2"; do
  echo "## $W"
  run "simple" DRTVAM_GENERAL_KERNELS=simple -- --workload $W
  for B in 2 3 4; do
    run "persist blocks=$B burst=16 min=8" DRTVAM_PERSIST_BLOCKS=$B -- --workload $W
  done
  run "persist blocks=3 burst=32 min=8" DRTVAM_PERSIST_BLOCKS=3 DRTVAM_PERSIST_BURST=32 -- --workload $W
  run "persist blocks=3 burst=8 min=8" DRTVAM_PERSIST_BLOCKS=3 DRTVAM_PERSIST_BURST=8 -- --workload $W
  run "persist blocks=3 burst=16 min=16 wait=6" DRTVAM_PERSIST_BLOCKS=3 DRTVAM_PERSIST_SETUP_MIN=16 DRTVAM_PERSIST_MAX_WAIT=6 -- --workload $W
  run "persist blocks=3 burst=16 min=4 wait=2" DRTVAM_PERSIST_BLOCKS=3 DRTVAM_PERSIST_SETUP_MIN=4 DRTVAM_PERSIST_MAX_WAIT=2 -- --workload $W
  run "persist blocks=4 burst=32 min=12 wait=4" DRTVAM_PERSIST_BLOCKS=4 DRTVAM_PERSIST_BURST=32 DRTVAM_PERSIST_SETUP_MIN=12 DRTVAM_PERSIST_MAX_WAIT=4 -- --workload $W
done
