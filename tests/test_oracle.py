"""CPU tests of the oracle (oracle/drtvam_oracle.c): the pins available without
Mitsuba -- the reference's own ray-origin bound (tests/test_projector.py:7-38),
the FD-vs-adjoint consistency of tests/test_integrators.py:69-110, and analytic
identities of the forward operator.  PARITY UNPINNED for absolute dose values
(the reference holds no golden volumes)."""
import numpy as np
import pytest

import oracle as orc
from drtvam_b200._scene import SceneDesc, SurfaceDesc
from drtvam_b200.motion import CircularMotion
from helpers import build, pixels_np, random_inputs, rel_l2


def _motion_table(distance, n):
    m = CircularMotion({"distance": distance})
    return np.stack([m.eval(np.float32(k) / np.float32(n)).matrix for k in range(n)])


def test_crop_ray_origins_match_reference_bound():
    """tests/test_projector.py:9-38: 20x10 DMD, crop 4x4 at (8,3), pixel 1, distance 20
    => every origin has y, z in (-2, 2)."""
    from drtvam_b200.projector import CollimatedProjector
    proj = CollimatedProjector({"n_patterns": 1, "resx": 20, "resy": 10, "cropx": 4, "cropy": 4,
                                "crop_offset_x": 8, "crop_offset_y": 3, "pixel_size": 1.,
                                "motion": "circular", "distance": 20, "device": "cpu"})
    assert proj.active_size() == 16
    d = SceneDesc(proj_kind=0, n_patterns=1, res_x=20, res_y=10, pixel_x=1., pixel_y=1.,
                  motion_table=proj.motion_table(), regular_sampling=0, sigma_t=1.0,
                  surfaces=[SurfaceDesc(radius=1., z0=-1, z1=1, interior_medium=1)])
    pix = proj.active_pixels.numpy().astype(np.uint32)
    for prec in ("f32", "f64"):
        rays = orc.Oracle(d.as_dict(), prec).raygen(pix, spp=128, seed=0)
        assert rays.shape == (16 * 128, 6)
        o = rays[:, :3]
        assert np.all((o[:, 1] > -2) & (o[:, 1] < 2) & (o[:, 2] > -2) & (o[:, 2] < 2))
        np.testing.assert_allclose(o[:, 0], 20 - 0.01, rtol=1e-6)
        np.testing.assert_allclose(rays[:, 3:], np.tile([-1., 0., 0.], (len(rays), 1)), atol=1e-6)


def test_look_at_convention():
    """at alpha = 0 local (x, y, z) maps to world (D - z, -x, y)."""
    m = CircularMotion({"distance": 20.}).eval(0.0).matrix
    p = m @ np.array([1.5, 2.5, 0.25, 1.0])
    np.testing.assert_allclose(p[:3], [20 - 0.25, -1.5, 2.5], atol=1e-6)
    m = CircularMotion({"distance": 20., "clockwise": True}).eval(0.25).matrix
    np.testing.assert_allclose(m[:3, 3], [0, -20, 0], atol=1e-5)


def test_rng_stream_is_deterministic_and_uniform():
    a = orc.rng_stream(7, 123, 4096)
    b = orc.rng_stream(7, 123, 4096)
    c = orc.rng_stream(7, 124, 4096)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert a.min() >= 0 and a.max() < 1
    assert abs(a.mean() - 0.5) < 0.02 and abs(a.var() - 1 / 12) < 0.01


def test_pcg32_known_answer():
    """PCG32 reference stream (pcg32-demo, initstate 42, initseq 54)."""
    import ctypes as C
    lib, _ = orc._lib("f64")
    # the oracle seeds through TEA; check the raw generator by reproducing it here
    state, inc = 0, (54 << 1) | 1
    mask = (1 << 64) - 1

    def step():
        nonlocal state
        old = state
        state = (old * 0x5851f42d4c957f2d + inc) & mask
        xs = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        return ((xs >> rot) | (xs << ((-rot) & 31))) & 0xFFFFFFFF

    step(); state = (state + 42) & mask; step()
    got = [step() for _ in range(6)]
    assert got == [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]


@pytest.mark.parametrize("name", ["C1", "C2", "C3", "C4"])
def test_adjoint_dot_test_f64(name, workdir):
    """<A p, g> == <p, A^T g> to rounding in double precision."""
    scene, integ, sensor, d, _ = build(name, workdir, 24)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 1)
    o = orc.Oracle(d.as_dict(), "f64")
    lhs = float(np.sum(o.forward(p, pix) * g))
    rhs = float(np.sum(o.adjoint(g, pix) * p))
    assert abs(lhs - rhs) <= 1e-11 * abs(lhs)


@pytest.mark.parametrize("name", ["C1", "C3", "C4"])
def test_f32_build_tracks_f64(name, workdir):
    scene, integ, sensor, d, _ = build(name, workdir, 32)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 2)
    f64 = orc.Oracle(d.as_dict(), "f64").forward(p, pix)
    f32 = orc.Oracle(d.as_dict(), "f32").forward(p, pix)
    assert rel_l2(f32, f64) < 2e-5
    a64 = orc.Oracle(d.as_dict(), "f64").adjoint(g, pix)
    a32 = orc.Oracle(d.as_dict(), "f32").adjoint(g, pix)
    assert rel_l2(a32, a64) < 2e-5


def test_energy_identity_index_matched(workdir):
    """Every ray of an index-matched vial crosses the grid along one chord, so
    sum(film) = sum_rays em * (exp(-s t_in) - exp(-s t_out)), t measured from the
    vial entry point (sensor.py:400-404 summed over the voxels of a chord)."""
    scene, integ, sensor, d, cfg = build("C2", workdir, 16)
    pix = pixels_np(scene)
    p = np.ones(pix.size, np.float32)
    o = orc.Oracle(d.as_dict(), "f64")
    film = o.forward(p, pix)
    rays = o.raygen(pix)
    R, half, st = cfg["vial"]["r"], 5.0, cfg["vial"]["medium"]["extinction"]
    w = d.pixel_x * d.pixel_y
    total = 0.0
    for ox, oy, oz, dx, dy, dz in rays:
        # vial entry / exit
        b = ox * dx + oy * dy
        c = ox * ox + oy * oy - R * R
        disc = b * b - c
        if disc <= 0 or abs(oz) >= half:
            continue
        t0, t1 = -b - np.sqrt(disc), -b + np.sqrt(disc)
        # grid slab in xy
        lo, hi = t0, t1
        for o_, d_ in ((ox, dx), (oy, dy)):
            if abs(d_) < 1e-12:
                if abs(o_) > half:
                    lo, hi = 1, 0
                continue
            a, b2 = (-half - o_) / d_, (half - o_) / d_
            lo, hi = max(lo, min(a, b2)), min(hi, max(a, b2))
        if hi > lo:
            total += w * (np.exp(-st * (lo - t0)) - np.exp(-st * (hi - t0)))
    assert abs(film.sum() - total) < 2e-4 * total  # spawn_ray offset shifts t by ~1e-3 mm


def test_axis_aligned_single_ray_closed_form():
    """One ray along -x through the centre row of a 8^3 grid: voxel k receives
    em * e^{-s t_k} (1 - e^{-s h}) (sensor.py:404)."""
    n, ext, st, R = 8, 8.0, 0.2, 6.0
    d = SceneDesc(proj_kind=0, n_patterns=1, res_x=1, res_y=1, pixel_x=0.5, pixel_y=0.5,
                  motion_table=_motion_table(20., 1), regular_sampling=1, sigma_t=st, max_depth=6, rr_depth=6,
                  surfaces=[SurfaceDesc(radius=R, z0=-20, z1=20, interior_medium=1)],
                  bbox_min=(-ext / 2,) * 3, bbox_max=(ext / 2,) * 3, res=(n, n, n))
    o = orc.Oracle(d.as_dict(), "f64")
    film = o.forward(np.array([3.0], np.float32), np.array([0], np.uint32))
    # pixel centre is at local (0,0): world y = z = 0 -> truncation puts it in voxel n/2 on both axes
    row = film[n // 2, n // 2, :, 0]
    assert np.count_nonzero(film) == n
    em = 3.0 * 0.25
    h = ext / n
    eps_off = (1 + R) * 1500 * 2.0 ** -24  # spawn_ray offset shortens the lead-in
    for k in range(n):
        x_hi = ext / 2 - k * h           # ray travels towards -x, enters voxel n-1-k at x_hi
        t_k = (R - eps_off) - x_hi
        expect = em * np.exp(-st * t_k) * (1 - np.exp(-st * h))
        assert abs(row[n - 1 - k] - expect) < 1e-6 * expect


def test_fresnel_normal_incidence_weight():
    """A ray through the axis of a cylindrical vial hits every wall at normal
    incidence: weight = prod (1 - ((n1-n2)/(n1+n2))^2) * eta_ti^2 (App. A.9)."""
    n_air, n_glass, n_med, st = 1.000277, 1.54, 1.49, 0.1
    r_int, r_ext, n, ext = 3.0, 4.0, 4, 4.0
    surfs = [SurfaceDesc(radius=r_ext, z0=-9, z1=9, bsdf=1, eta=n_glass / n_air),
             SurfaceDesc(radius=r_int, z0=-9, z1=9, bsdf=1, eta=n_med / n_glass, interior_medium=1)]
    d = SceneDesc(proj_kind=0, n_patterns=1, res_x=1, res_y=1, pixel_x=1., pixel_y=1.,
                  motion_table=_motion_table(20., 1), regular_sampling=1, sigma_t=st, surfaces=surfs,
                  bbox_min=(-ext / 2,) * 3, bbox_max=(ext / 2,) * 3, res=(n, n, n))
    film = orc.Oracle(d.as_dict(), "f64").forward(np.array([1.0], np.float32), np.array([0], np.uint32))

    def T(a, b):
        return 1 - ((a - b) / (a + b)) ** 2

    weight = T(n_air, n_glass) * (n_air / n_glass) ** 2 * T(n_glass, n_med) * (n_glass / n_med) ** 2
    eps_off = (1 + r_int) * 1500 * 2.0 ** -24
    t_in = (r_int - eps_off) - ext / 2
    expect = weight * (np.exp(-st * t_in) - np.exp(-st * (t_in + ext)))
    assert abs(film.sum() - expect) < 1e-6 * expect


def test_total_internal_reflection_kills_the_ray():
    """Ray at height y in a glass ring (n_g) around a medium (n_m): the sine law
    gives sin(theta) = y / (n_g r_int) at the inner wall, so r_int n_m < y < r_ext is
    totally reflected: weight 0, no dose."""
    surfs = [SurfaceDesc(radius=4.0, z0=-9, z1=9, bsdf=1, eta=1.9 / 1.000277),
             SurfaceDesc(radius=3.0, z0=-9, z1=9, bsdf=1, eta=1.0 / 1.9, interior_medium=1)]
    d = SceneDesc(proj_kind=0, n_patterns=1, res_x=64, res_y=1, pixel_x=8. / 64, pixel_y=1.,
                  motion_table=_motion_table(20., 1), regular_sampling=1, sigma_t=0.5, surfaces=surfs,
                  bbox_min=(-4., -4., -4.), bbox_max=(4., 4., 4.), res=(16, 16, 1))
    o = orc.Oracle(d.as_dict(), "f64")
    edge = o.forward(np.array([1.0], np.float32), np.array([1], np.uint32))   # |y| = 3.81 in (3, 4): TIR at the inner wall
    mid = o.forward(np.array([1.0], np.float32), np.array([32], np.uint32))
    assert edge.sum() == 0.0 and mid.sum() > 0.0


def test_fd_vs_adjoint_scattering(workdir):
    """tests/test_integrators.py:69-110 restated: d/da mean(vol(a p)^2) by central
    differences equals the reverse-mode gradient, for a scattering medium in a
    glass cylinder with sample_time and Russian roulette."""
    n_pat, n = 12, 16
    d_ext, d_int = 16.77, 15.33
    surfs = [SurfaceDesc(radius=0.5 * d_ext, z0=-10, z1=10, bsdf=1, eta=1.514 / 1.000277),
             SurfaceDesc(radius=0.5 * d_int, z0=-10, z1=10, bsdf=1, eta=1.4849 / 1.514, interior_medium=1)]
    d = SceneDesc(proj_kind=0, n_patterns=n_pat, res_x=n, res_y=n, pixel_x=d_ext / n, pixel_y=d_ext / n,
                  motion_analytic=1, distance=1.5 * d_ext, motion_table=_motion_table(1.5 * d_ext, n_pat),
                  max_depth=32, rr_depth=3, sample_time=1, regular_sampling=0,
                  sigma_t=0.1, albedo=0.5, phase_kind=1, surfaces=surfs,
                  bbox_min=(-d_ext / 2,) * 3, bbox_max=(d_ext / 2,) * 3, res=(n, n, n))
    pix = np.arange(n_pat * n * n, dtype=np.uint32)
    patterns = np.linspace(1, 10, pix.size).astype(np.float32)
    o = orc.Oracle(d.as_dict(), "f64")
    spp, eps = 8, 1e-3
    inv_vol = 1.0 / (d_ext / n) ** 3

    def loss(a):
        vol = o.forward((patterns * a).astype(np.float32), pix, spp=spp, seed=0) * inv_vol
        return np.mean(vol ** 2), vol

    l1, _ = loss(1 + eps)
    l2, _ = loss(1 - eps)
    fd = (l1 - l2) / (2 * eps)
    _, vol = loss(1.0)
    dL = 2 * vol / vol.size * inv_vol           # d loss / d film
    g = o.adjoint(dL, pix, spp=spp, seed=0)     # same seed -> exact transpose
    ad = float(np.sum(g * patterns))
    assert abs((ad - fd) / fd) < 2e-4           # tolerance of test_integrators.py:110


def test_scattering_conserves_less_than_absorptive(workdir):
    """With albedo > 0 part of the light is scattered instead of absorbed; the
    expected-value estimator deposits (1 - albedo) of the extinction."""
    scene, integ, sensor, d, _ = build("C3", workdir, 16)
    pix = pixels_np(scene)
    p = np.ones(pix.size, np.float32)
    base = orc.Oracle(d.as_dict(), "f64").forward(p, pix).sum()
    d2 = build("C5", workdir, 16, spp=4)[3]
    sc = orc.Oracle(d2.as_dict(), "f64").forward(p, pix, spp=4, seed=3).sum()
    assert 0.05 * base < sc < 0.6 * base
