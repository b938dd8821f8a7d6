"""Triangle-mesh interfaces of the traced scene: the 'ply' walls of a custom vial (CustomVial.to_dict,
geometry.py:113-134, dielectric with face normals) and the occluders of Container.add_occlusions
(geometry.py:55-72, black diffuse by default).  CPU tests pin the lowering and the oracle; GPU tests
compare the CUDA tracer (BVH walk, csrc/mesh.cuh) with the oracle (brute force over the triangles)."""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle as orc
from drtvam_b200 import _lib
from drtvam_b200._scene import BSDF_ABSORBER, BSDF_DIELECTRIC, SURF_BOX, SURF_MESH, create_handle
from drtvam_b200.integrators import VolumeIntegrator
from drtvam_b200.optimize import load_scene
from drtvam_b200.scene import load_dict
from helpers import pixels_np, random_inputs, rel_l2
from test_optimization import fixture_config


def small(name, workdir, n=24, device="cpu", **integrator):
    cfg = fixture_config(name, workdir)
    cfg["projector"].update({"n_patterns": n, "resx": n, "resy": 6, "pixel_size": 5.0 / n})
    cfg["sensor"]["film"].update({"resx": n, "resy": n, "resz": 8})
    scene = load_dict(load_scene(cfg), device=device)
    props = {"max_depth": 6, "rr_depth": 6, "regular_sampling": True, "print_time": 1.0}
    props.update(integrator)
    integ = VolumeIntegrator(props)
    sensor = scene.sensor("sensor")
    return scene, integ, sensor, scene.describe(integ, sensor)


def test_lowering_of_mesh_shapes(workdir):
    scene, integ, sensor, d = small("box_hole_occlusion", workdir)
    kinds = [s.kind for s in d.ordered_surfaces()]
    assert kinds == [SURF_BOX, SURF_BOX, SURF_MESH]
    occ = d.ordered_surfaces()[-1]
    assert occ.bsdf == BSDF_ABSORBER and occ.exterior_medium == 1 and occ.interior_medium == 0
    assert occ.triangles.shape == (12, 3, 3)
    cdesc, _ = d.to_c()
    assert cdesc.n_surf == 2                      # the mesh is added to the handle after creation
    scene, integ, sensor, d = small("box_hole_custom_cuvette", workdir)
    outer, inner = d.ordered_surfaces()
    assert outer.kind == inner.kind == SURF_MESH and outer.bsdf == inner.bsdf == BSDF_DIELECTRIC
    assert outer.eta == pytest.approx(1.4702 / 1.000277) and inner.eta == pytest.approx(1.33 / 1.4702)
    assert inner.interior_medium == 1 and outer.interior_medium == 0
    assert outer.triangles.shape == inner.triangles.shape == (8, 3, 3)
    assert d.to_c()[0].n_surf == 0


def test_oracle_occluder_casts_a_shadow(workdir):
    """Rays that meet the black box deposit up to it and nothing behind it; the adjoint is the
    transpose of the forward projection."""
    scene, integ, sensor, d = small("box_hole_occlusion", workdir)
    pix = pixels_np(scene)
    o = orc.Oracle(d.as_dict(), "f64")
    p, g = random_inputs(d, pix.size, 3)
    f = o.forward(p, pix)
    a = o.adjoint(g, pix)
    lhs, rhs = float((f * g).sum()), float((a * p.astype(np.float64)).sum())
    assert abs(lhs - rhs) < 1e-9 * max(abs(lhs), 1.0)
    # same scene without the occluder
    import copy
    d2 = copy.copy(d)
    d2.surfaces = [s for s in d.surfaces if s.kind != SURF_MESH]
    f2 = orc.Oracle(d2.as_dict(), "f64").forward(p, pix)
    assert f.sum() < 0.97 * f2.sum()
    # voxels strictly inside the occluder (|x| < 1, |y| < 0.5, |z| < 0.25) receive nothing
    rz, ry, rx, _ = f.shape
    xs = (np.arange(rx) + 0.5) / rx * 5 - 2.5
    zs = (np.arange(rz) + 0.5) / rz * 1.25 - 0.625
    inside = (np.abs(zs)[:, None, None] < 0.2) & (np.abs(xs)[None, :, None] < 0.25) & (np.abs(xs)[None, None, :] < 0.75)
    assert f[..., 0][inside].sum() == 0 and f2[..., 0][inside].sum() > 0
    assert np.all(f <= f2 + 1e-12)


class MeshHandle:
    def __init__(self, desc):
        self.lib = _lib.load()
        self.h, self.keep = create_handle(desc, 0)
        self.film_shape = desc.film_shape()

    def forward(self, p, pix, spp=1, seed=0):
        p = torch.as_tensor(p, device="cuda")
        pix = torch.as_tensor(pix.astype(np.int64), device="cuda").to(torch.int32)
        film = torch.empty(self.film_shape, device="cuda", dtype=torch.float32)
        _lib.check(self.lib.drtvam_forward(self.h, p.data_ptr(), pix.data_ptr(), p.numel(), spp, seed, 1.0,
                                           film.data_ptr(), 1, None))
        torch.cuda.synchronize()
        return film.cpu().numpy()

    def adjoint(self, g, pix, spp=1, seed=0):
        g = torch.as_tensor(g, device="cuda").contiguous()
        pix = torch.as_tensor(pix.astype(np.int64), device="cuda").to(torch.int32)
        out = torch.empty(pix.numel(), device="cuda", dtype=torch.float32)
        _lib.check(self.lib.drtvam_adjoint(self.h, g.data_ptr(), pix.data_ptr(), pix.numel(), spp, seed, 1.0,
                                           out.data_ptr(), None))
        torch.cuda.synchronize()
        return out.cpu().numpy()

    def __del__(self):
        self.lib.drtvam_scene_destroy(self.h)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["box_hole_occlusion", "box_hole_custom_cuvette"])
@pytest.mark.parametrize("regular", [True, False])
def test_mesh_interfaces_match_oracle(name, regular, workdir):
    scene, integ, sensor, d = small(name, workdir, n=32, regular_sampling=regular)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 5)
    spp = 1 if regular else 2
    o = orc.Oracle(d.as_dict(), "f32")
    h = MeshHandle(d)
    assert h.lib.drtvam_scene_plan(h.h, pix.size, spp) == 0      # general tracer
    f_gpu, f_cpu = h.forward(p, pix, spp=spp, seed=3), o.forward(p, pix, spp=spp, seed=3)
    a_gpu, a_cpu = h.adjoint(g, pix, spp=spp, seed=3), o.adjoint(g, pix, spp=spp, seed=3)
    assert f_cpu.sum() > 0
    # deterministic rays: 1e-4 (north_star); jittered telecentric rays can fall either side of a
    # wall edge in fp32, a handful of rays out of 10^4
    tol = 1e-4 if regular else 2e-3
    assert rel_l2(f_gpu, f_cpu) < tol
    assert rel_l2(a_gpu, a_cpu) < tol


@pytest.mark.gpu
def test_mesh_surface_abi_errors(workdir):
    scene, integ, sensor, d = small("box_hole_custom_cuvette", workdir, n=16)
    lib = _lib.load()
    cdesc, keep = d.to_c()
    h = C.c_void_p()
    _lib.check(lib.drtvam_scene_create(C.byref(cdesc), 0, C.byref(h)))
    # no surface yet: the projection calls refuse like common.py:36-37
    p = torch.ones(4, device="cuda")
    pix = torch.arange(4, device="cuda", dtype=torch.int32)
    film = torch.empty(d.film_shape(), device="cuda")
    assert lib.drtvam_forward(h.value, p.data_ptr(), pix.data_ptr(), 4, 1, 0, 1.0, film.data_ptr(), 1, None) != 0
    assert b"No printing medium" in lib.drtvam_last_error()
    tri = np.ascontiguousarray(d.ordered_surfaces()[0].triangles, np.float32)
    assert lib.drtvam_scene_add_mesh_surface(h.value, tri.ctypes.data, 0, 1, 1.5, 0, 0) != 0
    assert lib.drtvam_scene_add_mesh_surface(h.value, tri.ctypes.data, tri.shape[0], 1, 0.0, 0, 0) != 0
    for i in range(4):
        assert lib.drtvam_scene_add_mesh_surface(h.value, tri.ctypes.data, tri.shape[0], 1, 1.5, 1, 0) == 0
    assert lib.drtvam_scene_add_mesh_surface(h.value, tri.ctypes.data, tri.shape[0], 1, 1.5, 1, 0) != 0
    assert b"too many surfaces" in lib.drtvam_last_error()
    lib.drtvam_scene_destroy(h.value)
