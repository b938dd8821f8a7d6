"""The reflection branch of the bounce loop: `transmission_only=False`
(integrators/volume.py:238-243 -- forced transmission at depth 0, Fresnel-sampled reflection or
refraction afterwards, weight 1 resp. eta_ti^2 when both lobes are sampled, (1 - F) eta_ti^2 when only
the transmission lobe is).

CPU: the oracle against an independent derivation -- the EXPECTED dose of an axial ray through a
glass tube, enumerated over the reflect / transmit tree with its probabilities (no random numbers),
compared with the oracle's mean over 4096 lanes.  GPU: the CUDA tracer against the oracle on the same
PCG32 streams, and finite differences against the adjoint."""
import numpy as np
import pytest

import oracle as orc
from drtvam_b200._scene import SceneDesc, SurfaceDesc
from helpers import rel_l2
from test_oracle import _motion_table

N_AIR, N_GLASS, N_MED = 1.000277, 1.9, 1.1
R_EXT, R_INT, HALF, SIGMA = 4.0, 3.0, 2.0, 0.15


def _tube(n_patterns, res, grid, transmission_only, max_depth, regular=1, **kw):
    surfs = [SurfaceDesc(radius=R_EXT, z0=-9, z1=9, bsdf=1, eta=N_GLASS / N_AIR),
             SurfaceDesc(radius=R_INT, z0=-9, z1=9, bsdf=1, eta=N_MED / N_GLASS, interior_medium=1)]
    pixel = 2.0 * HALF / res if res > 1 else 0.05
    return SceneDesc(proj_kind=0, n_patterns=n_patterns, res_x=res, res_y=res, pixel_x=pixel, pixel_y=pixel,
                     motion_table=_motion_table(20., n_patterns), regular_sampling=regular, sigma_t=SIGMA,
                     max_depth=max_depth, rr_depth=max_depth, transmission_only=transmission_only, surfaces=surfs,
                     bbox_min=(-HALF,) * 3, bbox_max=(HALF,) * 3, res=(grid,) * 3, **kw)


def _fresnel_normal(n1, n2):
    return ((n1 - n2) / (n1 + n2)) ** 2


def expected_axial_dose(max_depth, transmission_only, alpha=0.0):
    """E[sum of the film] per unit emitted radiance of a ray through the tube axis at projection angle
    `alpha`: every interface is met at normal incidence.  States: 'G_in' / 'G_out' = in the glass heading
    inwards / outwards, 'M' = about to cross the medium.  The cubic grid covers the middle
    2 HALF / max(|cos|, |sin|) of every medium chord."""
    f1, f2 = _fresnel_normal(N_AIR, N_GLASS), _fresnel_normal(N_GLASS, N_MED)
    chord = 2.0 * R_INT
    half = HALF / max(abs(np.cos(alpha)), abs(np.sin(alpha)))
    dep = np.exp(-SIGMA * (R_INT - half)) - np.exp(-SIGMA * (R_INT + half))   # sensor.py:400-404 summed over the chord

    def walk(state, depth, a):
        # `a` = attenuation (volume.py:263,265); returns the expected deposit from here on
        if depth >= max_depth or a == 0.0:
            return 0.0
        both = (not transmission_only) and depth != 0          # volume.py:238-243
        total = 0.0
        if state == "M":
            total += a * dep
            a = a * np.exp(-SIGMA * chord)                     # volume.py:263
            refl, eta_ti, nxt_t, nxt_r = f2, N_MED / N_GLASS, "G_out", "M"
        elif state == "G_in":
            refl, eta_ti, nxt_t, nxt_r = f2, N_GLASS / N_MED, "M", "G_out"
        elif state == "G_out":
            refl, eta_ti, nxt_t, nxt_r = f1, N_GLASS / N_AIR, "OUT", "G_in"
        else:                                                  # "AIR": the first interface
            refl, eta_ti, nxt_t, nxt_r = f1, N_AIR / N_GLASS, "G_in", "OUT"
        if both:
            if nxt_t != "OUT":
                total += (1 - refl) * walk(nxt_t, depth + 1, a * eta_ti ** 2)
            if nxt_r != "OUT":
                total += refl * walk(nxt_r, depth + 1, a)
        elif nxt_t != "OUT":
            total += walk(nxt_t, depth + 1, a * (1 - refl) * eta_ti ** 2)
        return total

    return walk("AIR", 0, 1.0)


@pytest.mark.parametrize("max_depth", [4, 9])
def test_oracle_reflection_expectation(max_depth):
    """Per-lane deposit of 4096 axial lanes (one central pixel, 4096 angles, each lane its own PCG32
    stream) against the enumerated expectation of its angle: the mean residual is within 4 standard
    errors; the transmission-only expectation differs by many standard errors, and
    transmission_only=True reproduces its own closed form lane by lane."""
    n = 4096
    pix = np.arange(n, dtype=np.uint32)
    ones = np.ones((32, 32, 32, 1))
    alphas = 2 * np.pi * np.arange(n) / n
    for t_only in (0, 1):
        d = _tube(n, 1, 32, t_only, max_depth)
        o = orc.Oracle(d.as_dict(), "f64")
        per_lane = o.adjoint(ones, pix) / (d.pixel_x * d.pixel_y)      # d sum(film) / d em of every lane
        expect = np.array([expected_axial_dose(max_depth, bool(t_only), a) for a in alphas])
        if t_only:
            # no lobe is sampled: every lane equals its closed form (spawn offsets ~1e-4 of a chord)
            np.testing.assert_allclose(per_lane, expect, rtol=3e-4)
        else:
            resid = per_lane - expect
            se = resid.std() / np.sqrt(n)
            assert se > 0
            assert abs(resid.mean()) < 4 * se + 2e-4 * expect.mean()
            only_t = np.array([expected_axial_dose(max_depth, True, a) for a in alphas])
            assert (expect - only_t).mean() > 5 * se   # the test can tell the two modes apart
    # the forward pass deposits exactly what the adjoint gathers
    d = _tube(n, 1, 32, 0, max_depth)
    o = orc.Oracle(d.as_dict(), "f64")
    p = np.linspace(0.5, 1.5, n).astype(np.float32)
    lhs = o.forward(p, pix).sum()
    rhs = float(np.sum(o.adjoint(ones, pix) * p))
    assert abs(lhs - rhs) < 1e-10 * abs(lhs)


def test_oracle_reflection_dot_test_off_axis():
    """<A p, g> = <p, A^T g> with reflections on a 24^2 DMD (oblique incidence, total internal
    reflection near the rim), jittered sampling, f64."""
    d = _tube(6, 24, 20, 0, 8, regular=0)
    pix = np.arange(6 * 24 * 24, dtype=np.uint32)
    rng = np.random.default_rng(3)
    p = rng.random(pix.size).astype(np.float32)
    g = rng.standard_normal(d.film_shape())
    o = orc.Oracle(d.as_dict(), "f64")
    lhs = float(np.sum(o.forward(p, pix, 2, 5) * g))
    rhs = float(np.sum(o.adjoint(g, pix, 2, 5) * p))
    assert abs(lhs - rhs) < 1e-10 * abs(lhs)
    # reflections add dose: the transmitted lobe has the same expectation in both modes
    d1 = _tube(6, 24, 20, 1, 8, regular=0)
    base = orc.Oracle(d1.as_dict(), "f64").forward(p, pix, 2, 5).sum()
    assert o.forward(p, pix, 2, 5).sum() > 1.01 * base


@pytest.mark.gpu
@pytest.mark.parametrize("regular", [1, 0])
def test_reflection_branch_matches_oracle_on_gpu(regular):
    """Same lanes, same PCG32 streams: a lane's lobe choice `u <= F` (Mitsuba dielectric) can only differ
    where u is within rounding of F, so the deterministic bound applies to the regular run; the jittered
    run is compared with the fp32 oracle like the other jittered tests."""
    from test_gpu_parity import Handle
    d = _tube(12, 32, 32, 0, 8, regular=regular)
    pix = np.arange(12 * 32 * 32, dtype=np.uint32)
    rng = np.random.default_rng(9)
    p = (rng.random(pix.size) - 0.2).astype(np.float32)
    g = rng.standard_normal(d.film_shape()).astype(np.float32)
    h = Handle(d)
    assert h.lib.drtvam_scene_plan(h.h, pix.size, 1) == 0          # reflections: general tracer
    spp, seed = (1, 0) if regular else (4, 13)
    o = orc.Oracle(d.as_dict(), "f64" if regular else "f32")
    f_gpu, f_cpu = h.forward(p, pix, spp, seed), o.forward(p, pix, spp, seed)
    a_gpu, a_cpu = h.adjoint(g, pix, spp, seed), o.adjoint(g, pix, spp, seed)
    # a reflected path exists in the result: more dose than transmission-only
    d1 = _tube(12, 32, 32, 1, 8, regular=regular)
    assert np.abs(f_gpu).sum() > 1.005 * np.abs(Handle(d1).forward(p, pix, spp, seed)).sum()
    if regular:
        assert rel_l2(f_gpu, f_cpu) < 1e-4 and rel_l2(a_gpu, a_cpu) < 1e-4
    else:
        from helpers import assert_jitter_parity
        assert_jitter_parity(f_gpu, f_cpu, 8, 3 * 32, 1e-4, what="reflection jittered dose")
        assert_jitter_parity(a_gpu, a_cpu, 8, 1, 1e-4, what="reflection jittered gradient")


@pytest.mark.gpu
def test_reflection_fd_vs_adjoint_on_gpu():
    """tests/test_integrators.py:69-110 with transmission_only=False and a scattering medium: central
    differences of mean(vol^2) against the reverse-mode gradient, same seed (exact transpose), 2e-4."""
    from test_gpu_parity import Handle
    n_pat, n = 16, 32
    d = _tube(n_pat, n, n, 0, 12, regular=0, albedo=0.6, phase_kind=2, hg_g=0.5)
    pix = np.arange(n_pat * n * n, dtype=np.uint32)
    patterns = np.linspace(1, 10, pix.size).astype(np.float32)
    h = Handle(d)
    spp, eps = 16, 1e-3
    inv_vol = float(1.0 / d.voxel_volume())

    def loss(a):
        vol = h.forward((patterns * a).astype(np.float32), pix, spp, 0, inv_vol).astype(np.float64)
        return np.mean(vol ** 2), vol

    fd = (loss(1 + eps)[0] - loss(1 - eps)[0]) / (2 * eps)
    _, vol = loss(1.0)
    g = h.adjoint((2 * vol / vol.size).astype(np.float32), pix, spp, 0, inv_vol)
    ad = float(np.sum(g.astype(np.float64) * patterns))
    assert abs((ad - fd) / fd) < 2e-4
