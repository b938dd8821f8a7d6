"""Golden fixtures (tests/golden/*.npz, generator tests/golden/make_golden.py).

The fixtures are produced by the CPU oracle -- the reference itself (Mitsuba 3 /
Dr.Jit) cannot run in this image, see DESIGN.md section 3 -- and committed, so
both the oracle (CPU test below) and the CUDA path (GPU test below) are held
against a frozen answer.  Tolerance: 1e-4 relative L2 for the deterministic
cases (BASELINE.json north_star); the two stochastic cases use the same
PCG32/TEA streams on both sides and are compared against the fp32 oracle build
they were generated with: 1e-4 for the jittered absorptive case, 2e-2 for the
scattering case (an fp32 rounding difference can flip a free-flight decision
and re-route one of the 16 384 walks).
"""
import ctypes as C
import glob
import json
import os

import numpy as np
import pytest

import oracle as orc
from helpers import rel_l2

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURES = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))


def load(path):
    z = np.load(path, allow_pickle=False)
    desc = json.loads(str(z["desc"]))
    if desc.get("motion_table") is not None:
        desc["motion_table"] = np.asarray(desc["motion_table"], np.float64)
    return z, desc


def tolerance(name):
    return 2e-2 if "scattering" in name else 1e-4


def test_fixture_set_is_complete():
    names = {os.path.basename(p) for p in FIXTURES}
    for want in ("c1_index_matched_regular.npz", "c2_index_matched_regular.npz", "c3_cylindrical_regular.npz",
                 "c4_square_regular.npz", "c3_cylindrical_jitter_spp2.npz", "c5_scattering_hg_spp4.npz"):
        assert want in names


@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p)[:-4] for p in FIXTURES])
def test_oracle_reproduces_golden(path):
    z, desc = load(path)
    o = orc.Oracle(desc, str(z["oracle_build"]))
    spp, seed = int(z["spp"]), int(z["seed"])
    dose = o.forward(z["active_data"], z["active_pixels"], spp, seed)
    assert int(z["visits"]) == o.last_visits
    grad = o.adjoint(z["dose_grad"], z["active_pixels"], spp, seed)
    assert rel_l2(dose, z["dose"]) < 1e-6
    assert rel_l2(grad, z["pattern_grad"]) < 1e-6


def _desc_from_dict(d):
    from drtvam_b200._scene import SceneDesc, SurfaceDesc
    kw = {k: v for k, v in d.items() if k != "surfaces"}
    for k in ("bbox_min", "bbox_max", "res"):
        kw[k] = tuple(kw[k])
    sd = SceneDesc(**kw)
    sd.surfaces = [SurfaceDesc(**s) for s in d["surfaces"]]
    return sd


@pytest.mark.gpu
@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p)[:-4] for p in FIXTURES])
def test_cuda_reproduces_golden(path, general):
    import torch
    from drtvam_b200 import _lib
    z, desc = load(path)
    lib = _lib.load()
    sd = _desc_from_dict(desc)
    cdesc, keep = sd.to_c()
    h = C.c_void_p()
    _lib.check(lib.drtvam_scene_create(C.byref(cdesc), 0, C.byref(h)))
    try:
        _lib.check(lib.drtvam_scene_force_general(h, int(general)))
        spp, seed = int(z["spp"]), int(z["seed"])
        n = int(z["active_pixels"].size)
        # through the HOST-buffer entry points of the C-ABI
        p = np.ascontiguousarray(z["active_data"], np.float32)
        pix = np.ascontiguousarray(z["active_pixels"], np.uint32)
        dose = np.empty(sd.film_shape(), np.float32)
        _lib.check(lib.drtvam_forward_host(h, p.ctypes.data, pix.ctypes.data, n, spp, seed, 1.0, dose.ctypes.data))
        g = np.ascontiguousarray(z["dose_grad"], np.float32)
        grad = np.empty(n, np.float32)
        _lib.check(lib.drtvam_adjoint_host(h, g.ctypes.data, pix.ctypes.data, n, spp, seed, 1.0, grad.ctypes.data))
    finally:
        lib.drtvam_scene_destroy(h)
    tol = tolerance(os.path.basename(path))
    assert rel_l2(dose, z["dose"]) < tol
    assert rel_l2(grad, z["pattern_grad"]) < tol
    del torch
