#!/bin/bash
# bench.py --workload WL at N GPUs of one node, launched as the driver does
N=${1:-8}; TAG=${2:-scale}; shift 2; O=gpurun_out; mkdir -p $O
for SPEC in "$@"; do
  WL=${SPEC%%:*}; EXTRA=""; NAME=$WL; STEPS=10
  if [ "$SPEC" = "C4:sa" ]; then EXTRA="--surface-aware"; NAME=C4_surface_aware; fi
  if [ "$WL" = "C5" ]; then STEPS=2; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus $N --workload $WL $EXTRA --steps $STEPS --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_${NAME}_${N}gpu.json 2> $O/${TAG}_bench_${NAME}_${N}gpu.err; echo "$NAME exit $?"
  python - "$O/${TAG}_bench_${NAME}_${N}gpu.json" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
    print("n_gpus %d ms/step %.2f e2e %.2f fwd %.2f adj %.2f allreduce %.2f value %.3g loss %s"%(d["n_gpus"], d["ms_per_step"], d["e2e"]["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["allreduce_ms"], d["value"], d["loss"]))
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1200:])
PY
done
