#!/bin/bash
# bench.py at N GPUs of one node, launched as the driver does
N=${1:-2}; TAG=${2:-scale}; O=gpurun_out; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 > $O/${TAG}_bench_c2_${N}gpu.json 2> $O/${TAG}_bench_c2_${N}gpu.err; echo "exit $?"
python - "$O/${TAG}_bench_c2_${N}gpu.json" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
    print("n_gpus %d ms/step %.2f e2e %.2f fwd %.2f adj %.2f allreduce %.2f value %.3g loss %s"%(d["n_gpus"], d["ms_per_step"], d["e2e"]["ms_per_step"], r["kernel_ms"], r["adjoint_ms"], r["allreduce_ms"], d["value"], d["loss"]))
except Exception as e:
    print("ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
