#!/bin/bash
for ZL in 4 2 1; do
DRTVAM_Z_LAYERS=$ZL timeout 600 python bench.py --workload C2 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('Z_LAYERS=$ZL ms/step %.2f fwd %.2f adj %.2f'%(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['adjoint_ms']))"
done
