"""The alternate dose sensors of the path: ratio tracking (sensor.py:193-295) and delta tracking
(sensor.py:112-191).  Both are unbiased estimators of the dose the DDA sensor integrates
analytically, so on the CPU the oracle's three sensors are checked against each other; on the GPU
the CUDA kernels are checked against the oracle with the same PCG32 streams (the reference's own
test, tests/test_integrators.py:69-110, runs its FD-vs-adjoint check for each of the three)."""
import ctypes as C

import numpy as np
import pytest

import oracle as orc
from helpers import build, pixels_np, random_inputs, rel_l2

SENSOR_DDA, SENSOR_RATIO, SENSOR_DELTA = 0, 1, 2


def _desc(workdir, n, kind, spp):
    scene, integ, sensor, d, _ = build("C5", workdir, n, spp=spp)
    d.sensor_kind = kind
    d.majorant = 4.0 * d.sigma_t if kind == SENSOR_RATIO else 0.0
    return scene, d


def test_oracle_sensors_estimate_the_same_dose(workdir):
    """Total deposited energy and a coarse (4x4x4 block) dose agree between the analytic DDA sensor
    and the two Monte Carlo sensors within their sampling noise."""
    films = {}
    for kind in (SENSOR_DDA, SENSOR_RATIO, SENSOR_DELTA):
        scene, d = _desc(workdir, 16, kind, 32)
        pix = pixels_np(scene)
        p = np.ones(pix.size, np.float32)
        films[kind] = orc.Oracle(d.as_dict(), "f64").forward(p, pix, 32, 3)
    ref = films[SENSOR_DDA]
    coarse = lambda f: f.reshape(4, 4, 4, 4, 4, 4, -1).sum(axis=(1, 3, 5))
    for kind, tol_sum, tol_l2 in ((SENSOR_RATIO, 0.01, 0.05), (SENSOR_DELTA, 0.03, 0.2)):
        assert abs(films[kind].sum() - ref.sum()) < tol_sum * ref.sum()
        assert rel_l2(coarse(films[kind]), coarse(ref)) < tol_l2


@pytest.mark.parametrize("kind", [SENSOR_RATIO, SENSOR_DELTA])
def test_oracle_sensor_adjoint_is_the_transpose(kind, workdir):
    """<A p, g> = <p, A^T g> with the same seed (f64): both passes make the same random walks."""
    scene, d = _desc(workdir, 12, kind, 4)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 1)
    o = orc.Oracle(d.as_dict(), "f64")
    lhs = float(np.vdot(o.forward(p, pix, 4, 7).astype(np.float64), g.astype(np.float64)))
    rhs = float(np.vdot(p.astype(np.float64), o.adjoint(g, pix, 4, 7).astype(np.float64)))
    assert abs(lhs - rhs) < 1e-9 * max(abs(lhs), 1e-30)


def test_delta_sensor_rejects_absorptive_medium(workdir):
    """volume.py:156-157."""
    scene, integ, sensor, d, _ = build("C2", workdir, 8)
    d.sensor_kind = SENSOR_DELTA
    with pytest.raises(ValueError, match="purely absorptive"):
        orc.Oracle(d.as_dict(), "f64")
    from drtvam_b200.sensor import sensors, DeltaVolumetricSensor
    assert sensors["delta"] is DeltaVolumetricSensor and DeltaVolumetricSensor.kind == "delta"
    delta = DeltaVolumetricSensor({"to_world": sensor.to_world, "film": sensor.film()})
    with pytest.raises(ValueError, match="purely absorptive"):
        scene.describe(integ, delta)


# ---------------------------------------------------------------------------- GPU

@pytest.mark.gpu
@pytest.mark.parametrize("kind", [SENSOR_RATIO, SENSOR_DELTA])
def test_sensor_matches_oracle_on_gpu(kind, workdir):
    """Same random streams on both sides; fp32-vs-fp64 branch flips move a few deposits, hence the
    statistical bound (the delta sensor has one deposit per collision, so a flip weighs more)."""
    from test_gpu_parity import Handle
    scene, d = _desc(workdir, 24, kind, 4)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 3)
    o = orc.Oracle(d.as_dict(), "f32")
    h = Handle(d)
    f_gpu, f_cpu = h.forward(p, pix, 4, 5), o.forward(p, pix, 4, 5)
    tol = 1e-2 if kind == SENSOR_RATIO else 2e-2  # measured: 5.5e-3 (ratio, adjoint), 1e-2 (delta)
    assert rel_l2(f_gpu, f_cpu) < tol
    assert abs(f_gpu.sum() - f_cpu.sum()) < 2e-3 * f_cpu.sum()
    assert rel_l2(h.adjoint(g, pix, 4, 5), o.adjoint(g, pix, 4, 5)) < tol
    # the adjoint is the transpose of the forward pass on the device too
    lhs = float(np.vdot(f_gpu.astype(np.float64), g.astype(np.float64)))
    rhs = float(np.vdot(p.astype(np.float64), h.adjoint(g, pix, 4, 5).astype(np.float64)))
    assert abs(lhs - rhs) < 1e-4 * max(abs(lhs), abs(rhs))


@pytest.mark.gpu
def test_delta_sensor_needs_scattering_on_gpu(workdir):
    from drtvam_b200 import _lib
    scene, integ, sensor, d, _ = build("C2", workdir, 8)
    d.sensor_kind = SENSOR_DELTA
    lib = _lib.load()
    cdesc, keep = d.to_c()
    h = C.c_void_p()
    assert lib.drtvam_scene_create(C.byref(cdesc), 0, C.byref(h)) != 0
    assert b"purely absorptive" in lib.drtvam_last_error()
