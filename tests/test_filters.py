"""Pixel pre-filters of optimize.py:143-185: the Radon integrator (integrators/radon.py) and the
corner integrator (integrators/filter_corner.py).  CPU: properties of the oracle's restatement.
GPU: the CUDA kernels against the oracle through the C-ABI, and the host mirror through
drtvam_b200.render / optimize."""
import numpy as np
import pytest

import oracle as orc
from helpers import build, pixels_np, rel_l2


def _with_target(scene, d):
    """The filters run before the target is moved away (optimize.py:188-191): the mesh is part of
    the scene whether or not the film is surface-aware."""
    dd = d.as_dict()
    dd["target_triangles"] = scene.all_target_triangles()
    return dd


def test_oracle_radon_is_positive_exactly_where_the_ray_meets_the_target(workdir):
    """Index-matched vial, collimated beam: a pixel's Radon value is positive iff its ray crosses
    the box-with-hole target, i.e. iff the surface-aware film gets an inside-channel deposit."""
    scene, integ, sensor, d, _ = build("C2", workdir, 24, surface_aware=True)
    pix = pixels_np(scene)
    o = orc.Oracle(d.as_dict(), "f64")
    img = o.radon(pix, 1, 0)
    assert img.shape == (d.n_patterns, d.res_y, d.res_x) and (img >= 0).all()
    hit = np.zeros(pix.size, bool)
    g = np.zeros(d.film_shape(), np.float64)
    g[..., 0] = 1.0   # inside channel only
    hit = o.adjoint(g, pix) > 0
    assert ((img.reshape(-1)[pix] > 0) == hit).mean() > 0.999
    frac = (img > 0).mean()
    assert 0.2 < frac < 0.8
    # value: absorbed fraction inside the target, bounded by 1 - exp(-sigma_t * longest chord) times the weight
    w = d.pixel_x * d.pixel_y
    assert img.max() <= w * (1 - np.exp(-d.sigma_t * 7.0 * np.sqrt(3))) * 1.0001


def test_oracle_corner_filter_kills_the_rays_near_the_vial_edges(workdir):
    scene, integ, sensor, d, _ = build("C4", workdir, 32)
    pix = pixels_np(scene)
    o = orc.Oracle(d.as_dict(), "f64")
    dist = 0.5 * 12.408   # outer half width of the square vial of C4
    img = o.corner(pix, dist, 0.4)
    w = d.pixel_x * d.pixel_y
    vals = np.unique(np.round(img / w, 6))
    assert set(vals) <= {0.0, 1.0}
    killed = (img == 0).reshape(d.n_patterns, d.res_y, d.res_x)
    assert 0 < killed.mean() < 0.2
    assert (killed == killed[:, :1, :]).all()          # the edge is vertical: whole DMD columns go
    assert (o.corner(pix, dist, 1e-6) > 0).all()       # nothing within a vanishing radius


def test_host_mirror_registers_the_filter_integrators():
    from drtvam_b200.integrators import integrators, RadonIntegrator, CornerIntegrator
    assert integrators["radon"] is RadonIntegrator and integrators["corner"] is CornerIntegrator
    c = CornerIntegrator({"dist": 3.0})
    assert c.radius == 0.1 and c.to_string() == "CornerIntegrator"      # filter_corner.py:11-15
    with pytest.raises(KeyError):
        CornerIntegrator({})
    assert "RadonIntegrator" in RadonIntegrator({"max_depth": 5}).to_string()


# ---------------------------------------------------------------------------- GPU

def _gpu_image(d, tri, fn, pix, *args):
    import ctypes as C
    import torch
    from drtvam_b200 import _lib
    from test_gpu_parity import Handle
    h = Handle(d)
    if tri is not None:
        _lib.check(h.lib.drtvam_scene_set_target(h.h, tri.ctypes.data, tri.shape[0]))
    pixt = torch.as_tensor(pix.astype(np.int64), device="cuda").to(torch.int32)
    img = torch.full((d.n_patterns, d.res_y, d.res_x), 7.0, device="cuda")
    _lib.check(getattr(h.lib, fn)(h.h, pixt.data_ptr(), pix.size, *args, img.data_ptr(), None))
    torch.cuda.synchronize()
    return img.cpu().numpy(), h


@pytest.mark.gpu
@pytest.mark.parametrize("name,spp", [("C2", 1), ("C4", 1), ("C3", 4)])
def test_radon_matches_oracle_on_gpu(name, spp, workdir):
    scene, integ, sensor, d, _ = build(name, workdir, 32, regular_sampling=(spp == 1))
    d.max_depth = 5  # optimize.py:145-148
    pix = pixels_np(scene)
    tri = scene.all_target_triangles()
    dd = d.as_dict()
    dd["target_triangles"] = tri
    ref = orc.Oracle(dd, "f64").radon(pix, spp, 3)
    img, h = _gpu_image(d, tri, "drtvam_radon", pix, spp, 3)
    assert rel_l2(img, ref) < (1e-4 if spp == 1 else 2e-3)
    assert ((img > 0) != (ref > 0)).mean() < (1e-3 if spp == 1 else 5e-3)
    # a ragged pixel list only fills the listed pixels
    rag = pix[::3]
    img2, _ = _gpu_image(d, tri, "drtvam_radon", rag, spp, 3)
    mask = np.zeros(img.size, bool); mask[rag] = True
    assert (img2.reshape(-1)[~mask] == 0).all()
    if spp == 1:
        np.testing.assert_allclose(img2.reshape(-1)[mask], img.reshape(-1)[mask], rtol=1e-6, atol=0)
    # without the target the call fails like radon.py:57-58
    _, h0 = _gpu_image(d, None, "drtvam_corner_filter", pix, 1, 0, 1.0, 0.1)
    import torch
    buf = torch.empty(img.shape, device="cuda")
    pixt = torch.as_tensor(pix.astype(np.int64), device="cuda").to(torch.int32)
    assert h0.lib.drtvam_radon(h0.h, pixt.data_ptr(), pix.size, 1, 0, buf.data_ptr(), None) != 0
    assert b"No target shape" in h0.lib.drtvam_last_error()


@pytest.mark.gpu
def test_corner_filter_matches_oracle_on_gpu(workdir):
    scene, integ, sensor, d, _ = build("C4", workdir, 48)
    pix = pixels_np(scene)
    dist = 0.5 * 12.408
    ref = orc.Oracle(d.as_dict(), "f32").corner(pix, dist, 0.4)
    img, _ = _gpu_image(d, None, "drtvam_corner_filter", pix, 1, 0, dist, 0.4)
    assert ((img > 0) != (ref > 0)).mean() < 1e-3      # knife-edge pixels at the rim of the disc
    same = (img > 0) == (ref > 0)
    np.testing.assert_allclose(img[same], ref[same], rtol=1e-6)


@pytest.mark.gpu
def test_filters_through_the_python_api(workdir):
    """render() with the filter integrators, then the active-pixel update of optimize.py:151-159."""
    import torch
    import drtvam_b200 as dv
    from drtvam_b200.integrators import RadonIntegrator, CornerIntegrator
    scene, integ, sensor, d, cfg = build("C4", workdir, 32, device="cuda")
    n_all = scene.projector.active_pixels.numel()
    radon = dv.render(scene, integrator=RadonIntegrator({"max_depth": 5}), spp=4)
    assert radon.shape == scene.projector.size()
    keep = torch.nonzero(radon.reshape(-1) > 0).reshape(-1)
    assert 0 < keep.numel() < n_all
    prm = dv.traverse(scene)
    prm["projector.active_pixels"] = keep.to(torch.int32)
    prm["projector.active_data"] = torch.ones(keep.numel(), device="cuda")
    prm.update()
    corner = dv.render(scene, integrator=CornerIntegrator({"regular_sampling": True, "dist": 0.5 * 12.408,
                                                           "radius": 0.4}), spp=1)
    kept2 = torch.nonzero(corner.reshape(-1) > 0).reshape(-1)
    assert 0 < kept2.numel() < keep.numel()                       # only listed pixels can survive
    assert bool(torch.isin(kept2, keep).all())
    # the shrunken list renders through the general tracer (sparse pixel list) and matches the oracle
    prm["projector.active_pixels"] = kept2.to(torch.int32)
    prm["projector.active_data"] = torch.ones(kept2.numel(), device="cuda")
    prm.update()
    vol = dv.render(scene, prm, integrator=integ, sensor=sensor, spp=1)
    d2 = scene.describe(integ, sensor)
    pix = pixels_np(scene)
    ref = orc.Oracle(d2.as_dict(), "f64").forward(np.ones(pix.size, np.float32), pix) / d2.voxel_volume()
    assert rel_l2(vol.cpu().numpy(), ref) < 1e-4
    scene.close()
