"""Surface-aware discretization (film.surface_aware): the target mesh is part
of the traced scene, its crossings select the film channel (volume.py:196,219;
sensor.py:405-409) and VolumetricSensor.compute_volume (sensor.py:47-110)
yields the (inside, outside) sub-voxel volumes.

CPU tests pin the oracle restatement against properties; GPU tests compare the
CUDA path (through the C-ABI) with the oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle as orc
from helpers import build, pixels_np, random_inputs, rel_l2

# target of the synthetic configs: box 80x80x40 with a r=20 through hole, scaled to max extent 7 mm
SCALE = 7.0 / 80.0
TARGET_VOLUME = (80 * 80 * 40 - np.pi * 20 ** 2 * 40) * SCALE ** 3


def test_oracle_channels_sum_to_plain_dose(workdir):
    """inside + outside deposits equal the single-channel dose up to the
    re-spawn gaps at the crossings ((1 + |p|) 1500 2^-24 per crossing)."""
    scene, integ, sensor, d, _ = build("C4", workdir, 24, surface_aware=True)
    scene2, _, _, d2, _ = build("C4", workdir, 24)
    assert d.film_shape()[-1] == 2 and d.target_triangles is not None and d2.target_triangles is None
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 0)
    o, o2 = orc.Oracle(d.as_dict(), "f64"), orc.Oracle(d2.as_dict(), "f64")
    f, f2 = o.forward(p, pix), o2.forward(p, pix)
    assert f[..., 0].sum() > 0 and f[..., 1].sum() > 0
    assert rel_l2(f.sum(-1), f2[..., 0]) < 5e-3
    # adjoint dot test with the two-channel film
    a = o.adjoint(g, pix)
    lhs, rhs = float((f * g).sum()), float((a * p.astype(np.float64)).sum())
    assert abs(lhs - rhs) < 1e-9 * max(abs(lhs), 1.0)
    # the inside channel only holds voxels that touch the target (up to slivers at mesh edges, where
    # the re-spawn offset lets a ray meet the neighbouring triangle again)
    vol = o.compute_volume(256)
    assert f[..., 0][vol[..., 0] == 0].sum() < 1e-4 * f[..., 0].sum()


def test_oracle_compute_volume(workdir):
    scene, integ, sensor, d, _ = build("C4", workdir, 20, surface_aware=True)
    o = orc.Oracle(d.as_dict(), "f64")
    vol = o.compute_volume(512)
    voxel = d.voxel_volume()
    assert vol.shape == d.film_shape()
    np.testing.assert_allclose(vol.sum(-1), voxel, rtol=1e-6)          # every sample is in or out
    assert abs(vol[..., 0].sum() - TARGET_VOLUME) < 0.01 * TARGET_VOLUME  # analytic volume of the mesh
    # deep inside / far outside voxels are exactly full / empty
    rz, ry, rx, _ = vol.shape
    assert vol[0, 0, 0, 0] == 0 and vol[0, 0, 0, 1] == pytest.approx(voxel)
    # a voxel well inside the solid part (x > 0 side, away from the hole at x = -0.875 r = 1.75)
    xs = (np.arange(rx) + 0.5) / rx * 10 - 5
    ix, iy, iz = int(np.argmin(abs(xs - 2.5))), int(np.argmin(abs(xs - 0.0))), int(np.argmin(abs(xs - 0.0)))
    assert vol[iz, iy, ix, 0] == pytest.approx(voxel)


def test_host_mirror_lowers_target_only_when_surface_aware(workdir):
    scene, integ, sensor, d, cfg = build("C4", workdir, 16, surface_aware=True)
    assert scene.target_active and "final_sensor" in cfg
    final = scene.sensor("final_sensor")
    assert not final.film().surface_aware
    # the mesh stays in the traced scene for the final sensor too (optimize.py:188-191 only moves it
    # away when the run is not surface-aware)
    assert scene.describe(integ, final).target_triangles is not None
    scene2, integ2, _, d2, _ = build("C4", workdir, 16)
    assert not scene2.target_active and d2.target_triangles is None


# ---------------------------------------------------------------------------- GPU

def _handle(d):
    from test_gpu_parity import Handle
    h = Handle(d)
    tri = d.target_triangles
    from drtvam_b200 import _lib
    _lib.check(h.lib.drtvam_scene_set_target(h.h, tri.ctypes.data, tri.shape[0]))
    return h


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["C2", "C3", "C4"])
def test_surface_aware_forward_adjoint_match_oracle(name, workdir):
    """Tolerance 1e-4 relative L2 on both channels together (deterministic 1-spp path)."""
    scene, integ, sensor, d, _ = build(name, workdir, 40, surface_aware=True)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 5)
    o = orc.Oracle(d.as_dict(), "f64")
    h = _handle(d)
    # planar kernels for the channel sum + the cached inside-target segments (inside_kernels.cu)
    assert h.lib.drtvam_scene_plan(h.h, pix.size, 1) == 2
    f_gpu, f_cpu = h.forward(p, pix), o.forward(p, pix)
    assert f_gpu.shape == f_cpu.shape and f_gpu.shape[-1] == 2
    assert rel_l2(f_gpu, f_cpu) < 1e-4
    assert rel_l2(f_gpu[..., 0], f_cpu[..., 0]) < 2e-4
    assert rel_l2(f_gpu[..., 1], f_cpu[..., 1]) < 2e-4
    a_gpu = h.adjoint(g, pix)
    assert rel_l2(a_gpu, o.adjoint(g, pix)) < 1e-4
    # the same scene through the general tracer (BVH walk in every pass)
    from drtvam_b200 import _lib
    _lib.check(h.lib.drtvam_scene_force_general(h.h, 1))
    assert h.lib.drtvam_scene_plan(h.h, pix.size, 1) == 0
    assert rel_l2(h.forward(p, pix), f_gpu) < 1e-4
    assert rel_l2(h.adjoint(g, pix), a_gpu) < 1e-4
    _lib.check(h.lib.drtvam_scene_force_general(h.h, 0))
    # an angle shard (whole patterns from angle 3 on) and a ragged list (general tracer through the gate)
    hw = d.res_x * d.res_y
    sl = slice(3 * hw, 11 * hw)
    assert rel_l2(h.forward(p[sl], pix[sl]), o.forward(p[sl], pix[sl])) < 1e-4
    assert rel_l2(h.adjoint(g, pix[sl]), a_gpu[sl]) < 1e-4
    rag = np.arange(5, pix.size, 3)
    assert rel_l2(h.forward(p[rag], pix[rag]), o.forward(p[rag], pix[rag])) < 1e-4
    assert rel_l2(h.adjoint(g, pix[rag]), a_gpu[rag]) < 1e-4


@pytest.mark.gpu
def test_surface_aware_jittered_and_scattering(workdir):
    """Jittered absorptive: same PCG32 streams; the knife-edge rays (here also: rays that graze a
    triangle edge of the target) are counted and bounded, everything else is held to 1e-4.
    Scattering: statistical band (fp32 branch flips change single walks)."""
    from helpers import assert_jitter_parity
    scene, integ, sensor, d, _ = build("C4", workdir, 32, surface_aware=True, regular_sampling=False)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 6)
    o = orc.Oracle(d.as_dict(), "f32")
    h = _handle(d)
    assert_jitter_parity(h.forward(p, pix, 2, 9), o.forward(p, pix, 2, 9), max(8, pix.size * 2 // 2000), 3 * 32, 1e-4,
                         what="C4 surface-aware jittered dose")
    scene, integ, sensor, d, _ = build("C5", workdir, 24, surface_aware=True, spp=4)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 7)
    o = orc.Oracle(d.as_dict(), "f32")
    h = _handle(d)
    f_gpu, f_cpu = h.forward(p, pix, 4, 3), o.forward(p, pix, 4, 3)
    assert rel_l2(f_gpu, f_cpu) < 5e-3
    assert abs(f_gpu.sum() - f_cpu.sum()) < 1e-3 * f_cpu.sum()


@pytest.mark.gpu
def test_compute_volume_matches_oracle(workdir):
    """Same per-voxel PCG32 streams on both sides: the Monte Carlo counts agree except where an
    fp32 intersection falls on the other side of a knife edge."""
    scene, integ, sensor, d, _ = build("C4", workdir, 24, surface_aware=True)
    o = orc.Oracle(d.as_dict(), "f32")
    n = 256
    ref = o.compute_volume(n)
    h = _handle(d)
    vol = torch.empty(d.film_shape(), device="cuda", dtype=torch.float32)
    from drtvam_b200 import _lib
    _lib.check(h.lib.drtvam_compute_volume(h.h, n, vol.data_ptr(), None))
    vol = vol.cpu().numpy()
    voxel = d.voxel_volume()
    np.testing.assert_allclose(vol.sum(-1), voxel, rtol=1e-5)
    assert abs(vol[..., 0].sum() - TARGET_VOLUME) < 0.01 * TARGET_VOLUME
    diff = np.abs(vol - ref) / voxel
    assert diff.max() <= 3.0 / n + 1e-6          # at most a few samples differ in any voxel
    assert (diff > 1e-6).mean() < 5e-3
    # without a target the call fails like sensor.py:62-63
    scene2, _, _, d2, _ = build("C4", workdir, 16)
    from test_gpu_parity import Handle
    h2 = Handle(d2)
    v2 = torch.empty((16, 16, 16, 2), device="cuda")
    assert h2.lib.drtvam_compute_volume(h2.h, 16, v2.data_ptr(), None) != 0
    assert b"No target shape" in h2.lib.drtvam_last_error()


@pytest.mark.gpu
def test_surface_aware_python_api_and_loss(workdir):
    """render() on a surface-aware sensor: both channels normalised by their own sub-voxel volume
    (volume.py:41-42,54), autograd through drtvam_adjoint, two-channel loss (loss.py:38-47)."""
    import drtvam_b200 as dv
    from drtvam_b200.loss import ThresholdedLoss
    scene, integ, sensor, d, _ = build("C4", workdir, 32, device="cuda", surface_aware=True)
    target = sensor.compute_volume(scene, sample_count=128)
    assert target.shape == d.film_shape() and sensor.compute_volume(scene) is target
    pix = pixels_np(scene)
    p_np, g_np = random_inputs(d, pix.size, 8)
    p = torch.as_tensor(p_np, device="cuda").requires_grad_(True)
    prm = dv.traverse(scene)
    prm["projector.active_data"] = p
    prm.update()
    vol = dv.render(scene, prm, integrator=integ, sensor=sensor, spp=1, seed=0)
    o = orc.Oracle(d.as_dict(), "f64")
    tv = target.cpu().numpy().astype(np.float64)
    inv = np.where(tv != 0, 1.0 / np.where(tv != 0, tv, 1.0), 0.0)
    ref = o.forward(p_np, pix) * inv
    assert rel_l2(vol.detach().cpu().numpy(), ref) < 1e-4
    loss = ThresholdedLoss({})(vol, target, p)
    loss.backward()
    assert torch.isfinite(loss) and torch.isfinite(p.grad).all() and float(p.grad.abs().sum()) > 0
    # gradient of <vol, g> against the oracle adjoint with dL = g * inv_vol
    p.grad = None
    vol = dv.render(scene, prm, integrator=integ, sensor=sensor, spp=1, seed=0)
    (vol * torch.as_tensor(g_np, device="cuda")).sum().backward()
    assert rel_l2(p.grad.cpu().numpy(), o.adjoint(g_np * inv, pix)) < 1e-4
    # the final sensor is single-channel and still sees the mesh
    final = scene.sensor("final_sensor")
    with torch.no_grad():
        v2 = dv.render(scene, prm, integrator=integ, sensor=final, spp=1, seed=0)
    assert v2.shape[-1] == 1
    scene.close()


@pytest.mark.gpu
def test_optimize_with_surface_aware_sensor_writes_the_reference_outputs(workdir):
    """optimize() with a surface-aware sensor (optimize.py:113-116,131-134,354-359): the fractional target and its
    two EXR channels, the binary target on the final sensor's grid, a loss that goes down through the fused
    two-channel loss, and a final dose from the plain final_sensor."""
    import os
    from drtvam_b200.optimize import optimize
    from drtvam_b200.synthetic import make_config
    cfg = make_config("C4", workdir, n=32, surface_aware=True, n_steps=6)
    cfg["output"] = os.path.join(workdir, "out_sa_optimize")
    cfg["save_pattern_images"] = False
    vol = optimize(cfg, progress=False)
    assert tuple(vol.shape) == (32, 32, 32, 1)
    for name in ("target.npy", "target_in.exr", "target_out.exr", "target_binary.npy", "target_binary.exr",
                 "final.npy", "patterns.npz", "loss.npy"):
        assert os.path.exists(os.path.join(cfg["output"], name)), name
    loss = np.load(os.path.join(cfg["output"], "loss.npy"))
    assert np.isfinite(loss).all() and loss[5] < 0.2 * loss[0]
    assert np.load(os.path.join(cfg["output"], "target.npy")).shape == (32, 32, 32, 2)
    assert np.load(os.path.join(cfg["output"], "target_binary.npy")).shape == (32, 32, 32, 1)
