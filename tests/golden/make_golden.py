#!/usr/bin/env python
"""Generates tests/golden/*.npz: small input/output vectors of the forward
projection and its adjoint from the f64 CPU oracle.

The reference (pure Python on Mitsuba 3 / Dr.Jit) cannot be imported in this
image, so these vectors are ORACLE-generated, not reference-generated (DESIGN.md
section 3, "parity unpinned").  They freeze the oracle: a later change of
either the oracle or the CUDA kernels is caught against a committed answer.

    python tests/golden/make_golden.py          # rewrites the fixtures

Each fixture stores the flat scene description (the dict both the oracle and
the C-ABI consume), the seeded inputs and the outputs, in float32 to keep the
files small (the 1e-4 parity band is far above f32 rounding of the outputs).
"""
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle as orc  # noqa: E402
from helpers import build, pixels_np  # noqa: E402

# (fixture name, synthetic config, edge, regular_sampling, spp, seed, integrator overrides)
CASES = [
    ("c1_index_matched_regular", "C1", 24, True, 1, 0, {}),
    ("c2_index_matched_regular", "C2", 24, True, 1, 0, {}),
    ("c3_cylindrical_regular", "C3", 24, True, 1, 0, {}),
    ("c4_square_regular", "C4", 24, True, 1, 0, {}),
    ("c3_cylindrical_jitter_spp2", "C3", 20, False, 2, 7, {}),
    ("c5_scattering_hg_spp4", "C5", 16, False, 4, 3, {}),
]


def jsonable(d):
    out = {}
    for k, v in d.items():
        if isinstance(v, np.ndarray):
            out[k] = v.tolist()
        elif isinstance(v, (list, tuple)):
            out[k] = [jsonable(x) if isinstance(x, dict) else (x.tolist() if isinstance(x, np.ndarray) else x) for x in v]
        elif isinstance(v, dict):
            out[k] = jsonable(v)
        elif isinstance(v, (np.floating, np.integer)):
            out[k] = v.item()
        else:
            out[k] = v
    return out


def main():
    orc.build()
    work = tempfile.mkdtemp(prefix="drtvam_golden_")
    for name, cfg_name, n, regular, spp, seed, over in CASES:
        scene, integ, sensor, desc, cfg = build(cfg_name, work, n, regular_sampling=regular, integrator=over)
        pix = pixels_np(scene)
        rng = np.random.default_rng(1000 + seed)
        p = (rng.random(pix.size) - 0.25).astype(np.float32)       # signed, as A d feeds (lbfgs.py:240-248)
        g = rng.standard_normal(desc.film_shape()).astype(np.float32)
        kind = "f32" if not regular else "f64"                      # stochastic cases: fp32 = the reference's type
        o = orc.Oracle(desc.as_dict(), kind)
        dose = o.forward(p, pix, spp, seed)
        grad = o.adjoint(g, pix, spp, seed)
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            desc=json.dumps(jsonable(desc.as_dict())), oracle_build=kind, spp=spp, seed=seed,
            active_pixels=pix, active_data=p, dose_grad=g,
            dose=np.asarray(dose, np.float32), pattern_grad=np.asarray(grad, np.float32),
            visits=np.int64(o.last_visits))
        print(f"{name}: n_active={pix.size} film={desc.film_shape()} |dose|={np.linalg.norm(dose):.6g} "
              f"|grad|={np.linalg.norm(grad):.6g} visits={o.last_visits}")


if __name__ == "__main__":
    main()
