#!/usr/bin/env python
"""One-line digest of a bench.py JSON line (tools/gpu_session.sh prints it after every bench step)."""
import json
import sys

try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline") or {}
    e = d.get("e2e") or {}
    ds = d.get("default_sampling") or {}
    print("n_gpus %s ms/step %.3f e2e %.3f fwd %s adj %s allreduce %s plan %s value %.4g loss %s | default_sampling %s | check %s" % (
        d.get("n_gpus"), d["ms_per_step"], e.get("ms_per_step", float("nan")), r.get("kernel_ms"), r.get("adjoint_ms"),
        r.get("allreduce_ms"), r.get("plan"), d["value"], d.get("loss"),
        {k: ds.get(k) for k in ("ms_per_step", "fwd_ms", "adj_ms", "plan")} if ds else None, d.get("multi_gpu_check")))
except Exception as exc:  # noqa: BLE001
    print("ERR", exc)
    try:
        print(open(sys.argv[1].rsplit(".", 1)[0] + ".err").read()[-2000:])
    except OSError:
        pass
