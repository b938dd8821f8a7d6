"""The reference's own integrator test at the reference's own size and bound
(tests/test_integrators.py:7-110): 100 patterns of 128x128, a 128^3 grid, spp 128, a Rayleigh-scattering
resin (albedo 0.5) in a glass cylinder, sample_time, max_depth 32, rr_depth 3.  The finite-difference
derivative of mean(vol^2) with respect to a global pattern scale, taken with the `dda` sensor, must
agree to 2e-4 with the reverse-mode derivative through each of the `dda`, `ratio` and `delta`
sensors -- rendered, as in the reference, through the public render() entry point, so the adjoint
runs on its own de-correlated sample set (seed_grad = sample_tea_32(seed, 1)).

Also here: dose / gradient parity of the telecentric and lens projectors (projector.py:199-296) and
the forward-mode entry point render_forward (integrators/volume.py:58-95)."""
import numpy as np
import pytest
import torch

import oracle as orc
from helpers import build, pixels_np, random_inputs, rel_l2

pytestmark = pytest.mark.gpu

D_EXT, D_INT = 16.77, 15.33


def build_scene(method, n_pat=100, n=128):
    """tests/test_integrators.py:7-67, key for key (the emitter entry is called 'dmd' there)."""
    import drtvam_b200 as dv
    from drtvam_b200.transform import ScalarTransform4f
    patterns = torch.linspace(1, 10, n_pat * n * n, device="cuda").reshape(n_pat, n, n)
    scene_dict = {
        'type': 'scene',
        'dmd': {
            'type': 'collimated',
            'patterns': patterns,
            'pixel_size': D_EXT / n,
            'motion': 'circular',
            'distance': 1.5 * D_EXT,
        },
        'sensor': {
            'type': method,
            'to_world': ScalarTransform4f().scale(D_EXT),
            'film': {'type': 'vfilm', 'resx': n, 'resy': n, 'resz': n},
        },
        'integrator': {
            'type': 'volume',
            'print_time': 1,
            'max_depth': 32,
            'rr_depth': 3,
            'sample_time': True,
        },
        'vial_exterior': {
            'type': 'cylinder', 'p0': [0., 0., -10.], 'p1': [0., 0., 10.], 'radius': 0.5 * D_EXT,
            'bsdf': {'type': 'dielectric', 'int_ior': 1.514},
        },
        'vial_interior': {
            'type': 'cylinder', 'p0': [0., 0., -10.], 'p1': [0., 0., 10.], 'radius': 0.5 * D_INT,
            'bsdf': {'type': 'dielectric', 'ext_ior': 1.514, 'int_ior': 1.4849},
            'interior': {'type': 'homogeneous', 'sigma_t': 0.1, 'albedo': 0.5, 'phase': {'type': 'rayleigh'}},
        },
    }
    if method == 'ratio':
        scene_dict['sensor']['majorant'] = 10.
    return dv.load_dict(scene_dict, device="cuda")


def test_reverse_ad_at_reference_size():
    import drtvam_b200 as dv
    spp, eps = 128, 1e-3

    scene = build_scene('dda')
    params = dv.traverse(scene)
    key = 'projector.active_data'
    patterns = params[key].detach().clone()

    def loss_of(a):
        params[key] = patterns * a
        params.update()
        with torch.no_grad():
            vol = dv.render(scene, params, spp=spp, seed=0)
        return float(torch.mean(torch.square(vol.double())))   # f64 accumulation of the fp32 volume

    fd_grad = (loss_of(1.0 + eps) - loss_of(1.0 - eps)) / (2 * eps)
    scene.close()

    for method in ('dda', 'ratio', 'delta'):
        scene = build_scene(method)
        params = dv.traverse(scene)
        a = torch.ones((), device="cuda", dtype=torch.float32, requires_grad=True)
        params[key] = a * patterns
        params.update()
        vol = dv.render(scene, params, spp=spp)
        loss = torch.mean(torch.square(vol.double()))
        loss.backward()
        rel = abs((float(a.grad) - fd_grad) / fd_grad)
        print(f"{method}: d loss / d a = {float(a.grad):.8g}, finite differences {fd_grad:.8g}, rel {rel:.2e}")
        assert rel < 2e-4       # tests/test_integrators.py:110
        scene.close()


@pytest.mark.parametrize("kind", ["telecentric", "lens"])
def test_aperture_projectors_dose_and_gradient_match_oracle(kind, workdir):
    """Dose and pattern-gradient parity (not only ray generation) for the two projectors with an
    aperture: jittered pixel + aperture samples from the same PCG32 streams on both sides, general
    tracer, cylindrical vial with refraction."""
    from test_gpu_parity import Handle
    from helpers import assert_jitter_parity
    scene, integ, sensor, d, _ = build("C3", workdir, 40, regular_sampling=False)
    d.proj_kind = {"telecentric": 1, "lens": 2}[kind]
    d.aperture_radius, d.focus_distance = 0.05, 20.0
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 23)
    h = Handle(d)
    assert h.lib.drtvam_scene_plan(h.h, pix.size, 2) == 0
    o = orc.Oracle(d.as_dict(), "f32")
    spp, seed = 2, 4
    rays = pix.size * spp
    assert_jitter_parity(h.forward(p, pix, spp, seed), o.forward(p, pix, spp, seed), max(8, rays // 5000), 3 * 40,
                         1e-4, what=f"{kind} dose")
    assert_jitter_parity(h.adjoint(g, pix, spp, seed), o.adjoint(g, pix, spp, seed), max(8, rays // 5000), 1, 1e-4,
                         what=f"{kind} gradient")
    # deterministic pixel centres (regular_sampling): the aperture sample is still drawn, the 1e-4 bound
    # against the f64 oracle applies where no lane sits on a knife edge
    d.regular_sampling = 1
    h2, o64 = Handle(d), orc.Oracle(d.as_dict(), "f64")
    assert_jitter_parity(h2.forward(p, pix), o64.forward(p, pix), 8, 3 * 40, 1e-4, what=f"{kind} regular dose")
    assert_jitter_parity(h2.adjoint(g, pix), o64.adjoint(g, pix), 8, 1, 1e-4, what=f"{kind} regular gradient")


def test_render_forward_reference_signature(workdir):
    """render_forward(scene, params, sensor, seed, spp) (integrators/volume.py:58-66): the tangent is the
    gradient attached to params['projector.active_data'] (dr.set_grad in the reference); A is linear, so
    the result is A applied to the tangent."""
    import drtvam_b200 as dv
    scene, integ, sensor, d, _ = build("C3", workdir, 32, device="cuda")
    pix = pixels_np(scene)
    p_np, _ = random_inputs(d, pix.size, 8)
    t_np = (np.random.default_rng(2).random(pix.size) - 0.5).astype(np.float32)
    params = dv.traverse(scene)
    p = torch.as_tensor(p_np, device="cuda").requires_grad_(True)
    p.grad = torch.as_tensor(t_np, device="cuda")
    params["projector.active_data"] = p
    params.update()
    dvol = integ.render_forward(scene, params, sensor, seed=0, spp=1)
    assert tuple(dvol.shape) == d.film_shape()
    ref = orc.Oracle(d.as_dict(), "f64").forward(t_np, pix) / d.voxel_volume()
    assert rel_l2(dvol.cpu().numpy(), ref) < 1e-4
    # no tangent attached: the derivative is zero, as with a parameter that has no gradient enabled
    params["projector.active_data"] = torch.as_tensor(p_np, device="cuda")
    params.update()
    assert float(integ.render_forward(scene, params, sensor).abs().max()) == 0.0
