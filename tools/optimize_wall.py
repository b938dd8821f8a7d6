#!/usr/bin/env python
"""Wall time of a whole optimisation (the README's benchmark: 40 L-BFGS iterations, 400 angles x 400x400 DMD)
through drtvam_b200.optimize on the synthetic config of that shape.  Prints one JSON line."""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from drtvam_b200.optimize import optimize
from drtvam_b200.synthetic import make_config

name = sys.argv[1] if len(sys.argv) > 1 else "C2"
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
work = tempfile.mkdtemp(prefix="drtvam_opt_")
cfg = make_config(name, work, regular_sampling=True, n_steps=n_steps)
cfg["spp"] = cfg["spp_ref"] = 1
cfg["save_pattern_images"] = False
torch.cuda.init()
t0 = time.perf_counter()
vol = optimize(cfg, progress=False)
torch.cuda.synchronize()
t1 = time.perf_counter()
timing = np.load(os.path.join(cfg["output"], "timing.npy"))
loss = np.load(os.path.join(cfg["output"], "loss.npy"))
target = np.load(os.path.join(cfg["output"], "target.npy"))
thr = 0.5 * (0.9 + 0.95)
agree = float(np.mean((vol.cpu().numpy()[..., 0] > thr) == (target[..., 0] > 0.5)) * 100)
print(json.dumps({"workload": name, "n_steps": n_steps, "loop_s": float(timing[:, 1].sum()),
                  "s_per_iter_median": float(np.median(timing[:, 1])), "wall_s_incl_setup_and_io": t1 - t0,
                  "loss_first": float(loss[0]), "loss_last": float(loss[-1]), "thresholded_agreement_pct": agree}))
