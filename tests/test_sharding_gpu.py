"""Multi-GPU building blocks of the C-ABI, exercised on one GPU by running the ranks one after the other:
the z-slab window (drtvam_scene_set_z_window), the sampler index that makes a shard draw the samples of
the unsharded run (drtvam_scene_set_sampler_index), the pinned pixel list (drtvam_scene_pin_pixels) and
the NCCL entry points (world size 1).  The N-rank == 1-rank check on N real GPUs is bench.py's
`multi_gpu_check`; the host-side sharding logic runs on CPU over gloo (tests/test_parallel.py).
Reference: SURVEY.md section 8(e); every ray of a planar scene keeps its voxel layer (sensor.py:345)."""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle as orc
from drtvam_b200 import _lib
from helpers import build, flip_report, pixels_np, random_inputs, rel_l2
from test_gpu_parity import Handle

pytestmark = pytest.mark.gpu


def _window(h, z0, z1):
    r0, r1 = C.c_int32(0), C.c_int32(0)
    _lib.check(h.lib.drtvam_scene_set_z_window(h.h, z0, z1, C.byref(r0), C.byref(r1)))
    return r0.value, r1.value


def _rows_index(d, r0, r1):
    k = np.arange(d.n_patterns, dtype=np.int64)[:, None, None]
    r = np.arange(r0, r1, dtype=np.int64)[None, :, None]
    c = np.arange(d.res_x, dtype=np.int64)[None, None, :]
    return (k * (d.res_x * d.res_y) + r * d.res_x + c).reshape(-1)


class SlabHandle(Handle):
    def set_film(self, nz, d):
        self.film_shape = (nz,) + tuple(d.film_shape()[1:])


@pytest.mark.parametrize("name,n,grid", [("C2", 64, None), ("C3", 72, None), ("C4", 96, 48)])
def test_z_window_equals_slab_of_full_projection(name, n, grid, workdir):
    """Three z windows, one after the other: each forward fills exactly its layers of the unsharded dose,
    each adjoint (fed its slab of the dose gradient) returns exactly the gradient of its DMD rows.  The
    grid of C4 holds two DMD rows per layer."""
    scene, integ, sensor, d, _ = build(name, workdir, n, n_patterns=10, grid=grid)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 51)
    p = (p - 0.3).astype(np.float32)
    full = Handle(d)
    assert full.lib.drtvam_scene_plan(full.h, pix.size, 1) == 2
    f_full, a_full = full.forward(p, pix), full.adjoint(g, pix)
    rz = d.res[2]
    h = SlabHandle(d)
    covered = np.zeros(pix.size, bool)
    for rank in range(3):
        z0, z1 = rz * rank // 3, rz * (rank + 1) // 3
        r0, r1 = _window(h, z0, z1)
        idx = _rows_index(d, r0, r1)
        covered[idx] = True
        h.set_film(z1 - z0, d)
        assert h.lib.drtvam_scene_plan(h.h, idx.size, 1) == 2
        f = h.forward(p[idx], pix[idx])
        assert f.shape[0] == z1 - z0
        assert rel_l2(f, f_full[z0:z1]) < 1e-6
        a = h.adjoint(np.ascontiguousarray(g[z0:z1]), pix[idx])
        # the adjoint adds ~10 tile partials per ray with fp32 atomics in arbitrary order (and the window
        # regroups its layers): equal to rounding, not bit for bit
        assert rel_l2(a, a_full[idx]) < 1e-5
    # rows that belong to no window see no layer of the film: their gradient is zero
    assert np.abs(a_full[~covered]).max(initial=0.0) == 0.0
    # back to the whole film
    assert _window(h, 0, rz) == (0, d.res_y)
    h.film_shape = d.film_shape()
    assert rel_l2(h.forward(p, pix), f_full) < 1e-7
    # a window needs a planar scene; sparse lists inside a window go to the general tracer through the gate
    r0, r1 = _window(h, 0, rz // 2)
    idx = _rows_index(d, r0, r1)[::3]
    h.set_film(rz // 2, d)
    o = orc.Oracle(d.as_dict(), "f64")
    assert rel_l2(h.forward(p[idx], pix[idx]), o.forward(p[idx], pix[idx])[:rz // 2]) < 1e-4


def test_z_window_jittered_with_sampler_index(workdir):
    """regular_sampling=false: with the sampler index every lane of a window draws the samples the
    unsharded run draws for the same pixel, so the windows reproduce the unsharded result - except lanes
    whose jitter sample is exactly 0, which sit on a layer boundary and may belong to the next window
    (2^-23 of the lanes); they are counted."""
    scene, integ, sensor, d, _ = build("C3", workdir, 64, n_patterns=10, regular_sampling=False)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 52)
    full = Handle(d)
    spp, seed = 2, 6
    assert full.lib.drtvam_scene_plan(full.h, pix.size, spp) == 3
    f_full, a_full = full.forward(p, pix, spp, seed), full.adjoint(g, pix, spp, seed)
    h = SlabHandle(d)
    rz = d.res[2]
    for rank in range(2):
        z0, z1 = rz * rank // 2, rz * (rank + 1) // 2
        r0, r1 = _window(h, z0, z1)
        idx = _rows_index(d, r0, r1)
        gidx = torch.as_tensor(idx, device="cuda").to(torch.int32)
        _lib.check(h.lib.drtvam_scene_set_sampler_index(h.h, gidx.data_ptr()))
        h.set_film(z1 - z0, d)
        assert h.lib.drtvam_scene_plan(h.h, idx.size, spp) == 3
        flagged, rest, _ = flip_report(h.forward(p[idx], pix[idx], spp, seed), f_full[z0:z1])
        assert flagged <= 3 * 64 and rest < 1e-5
        flagged, rest, _ = flip_report(h.adjoint(np.ascontiguousarray(g[z0:z1]), pix[idx], spp, seed), a_full[idx])
        assert flagged <= 2 and rest < 1e-5
    # a grid whose rows straddle layers cannot be z-sharded with jitter
    scene, integ, sensor, d2, _ = build("C3", workdir, 48, n_patterns=4, grid=20, regular_sampling=False)
    h2 = Handle(d2)
    assert h2.lib.drtvam_scene_set_z_window(h2.h, 0, 10, None, None) == -1
    assert b"straddles" in h2.lib.drtvam_last_error()


def test_sampler_index_makes_angle_shards_reproduce_the_full_run(workdir):
    """Scattering resin (general tracer), two angle shards: with the sampler index the shards' doses sum to
    the unsharded dose and their gradients are slices of the unsharded gradient."""
    scene, integ, sensor, d, _ = build("C5", workdir, 32, n_patterns=8, spp=4)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 53)
    h = Handle(d)
    f_full, a_full = h.forward(p, pix, 4, 3), h.adjoint(g, pix, 4, 3)
    total = np.zeros_like(f_full)
    hw = d.res_x * d.res_y
    for lo, hi in ((0, 3 * hw), (3 * hw, 8 * hw)):
        gidx = torch.arange(lo, hi, device="cuda", dtype=torch.int32)
        _lib.check(h.lib.drtvam_scene_set_sampler_index(h.h, gidx.data_ptr()))
        total += h.forward(p[lo:hi], pix[lo:hi], 4, 3)
        assert rel_l2(h.adjoint(g, pix[lo:hi], 4, 3), a_full[lo:hi]) < 1e-6
    assert rel_l2(total, f_full) < 1e-5      # fp32 atomics in another order
    _lib.check(h.lib.drtvam_scene_set_sampler_index(h.h, None))
    # without the index the second shard draws other samples
    assert rel_l2(h.adjoint(g, pix[3 * hw:], 4, 3), a_full[3 * hw:]) > 1e-3
    # a scattering medium is planar only in its first segments (plan 4): scattered rays leave their layers, so it
    # must refuse a z window and be sharded by angle
    assert h.lib.drtvam_scene_plan(h.h, pix.size, 4) == 4
    assert h.lib.drtvam_scene_set_z_window(h.h, 0, 16, None, None) == -1
    assert b"shard by projection angle" in h.lib.drtvam_last_error()
    from drtvam_b200.parallel import zslab_shard
    scene_g, integ_g, sensor_g, _, _ = build("C5", workdir, 32, n_patterns=8, spp=4, device="cuda")
    assert zslab_shard(scene_g, 0, 2, integ_g, sensor_g) is None


def test_pinned_pixel_list(workdir):
    """A pinned list is checked once; planar and non-planar lists both keep giving the unpinned results,
    and re-pinning another list re-decides."""
    scene, integ, sensor, d, _ = build("C2", workdir, 64, n_patterns=12)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 54)
    h = Handle(d)
    f_ref, a_ref = h.forward(p, pix), h.adjoint(g, pix)
    pt = torch.as_tensor(p, device="cuda")
    pixt = torch.as_tensor(pix.astype(np.int64), device="cuda").to(torch.int32)
    gt = torch.as_tensor(g, device="cuda")
    film = torch.empty(d.film_shape(), device="cuda")
    out = torch.empty(pix.size, device="cuda")

    def run(pixels, data):
        _lib.check(h.lib.drtvam_forward(h.h, data.data_ptr(), pixels.data_ptr(), pixels.numel(), 1, 0, 1.0,
                                        film.data_ptr(), 1, None))
        _lib.check(h.lib.drtvam_adjoint(h.h, gt.data_ptr(), pixels.data_ptr(), pixels.numel(), 1, 0, 1.0,
                                        out.data_ptr(), None))
        torch.cuda.synchronize()
        return film.cpu().numpy(), out[:pixels.numel()].cpu().numpy()

    _lib.check(h.lib.drtvam_scene_pin_pixels(h.h, pixt.data_ptr(), pixt.numel()))
    n0 = h.lib.drtvam_launch_count()
    f1, a1 = run(pixt, pt)
    n1 = h.lib.drtvam_launch_count()
    f2, a2 = run(pixt, pt)
    n2 = h.lib.drtvam_launch_count()
    assert np.array_equal(f1, f_ref) and np.array_equal(f2, f_ref)     # the patch gather is deterministic
    assert rel_l2(a1, a_ref) < 1e-5 and rel_l2(a2, a_ref) < 1e-5       # fp32 atomics in arbitrary order
    assert n2 - n1 < n1 - n0                      # the layout check ran once
    # a permuted list of the same length, pinned: the general tracer alone serves it
    perm = torch.as_tensor(np.random.default_rng(0).permutation(pix.size), device="cuda")
    pix2, p2 = pixt[perm].contiguous(), pt[perm].contiguous()
    _lib.check(h.lib.drtvam_scene_pin_pixels(h.h, pix2.data_ptr(), pix2.numel()))
    f3, a3 = run(pix2, p2)
    assert rel_l2(f3, f_ref) < 1e-4
    assert rel_l2(a3, a_ref[perm.cpu().numpy()]) < 1e-4
    _lib.check(h.lib.drtvam_scene_pin_pixels(h.h, None, 0))
    f4, _ = run(pixt, pt)
    assert np.array_equal(f4, f_ref)


def test_jittered_segment_table_is_reused_for_the_direction_pass(workdir):
    """One L-BFGS iteration renders A p and A d with the same seed (lbfgs.py:248), with A^T g (another seed)
    in between: for a pinned list the per-lane segment table of A p serves A d.  Same result as a handle
    that rebuilds it, fewer launches' worth of work, and a new pin voids the table."""
    scene, integ, sensor, d, _ = build("C4", workdir, 48, n_patterns=8, regular_sampling=False)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 55)
    dvec = (np.random.default_rng(1).random(pix.size) - 0.5).astype(np.float32)
    ref = Handle(d)
    f_p, f_d = ref.forward(p, pix, 2, 5), ref.forward(dvec, pix, 2, 5)
    h = Handle(d)
    pixt = torch.as_tensor(pix.astype(np.int64), device="cuda").to(torch.int32)
    film = torch.empty(d.film_shape(), device="cuda")
    out = torch.empty(pix.size, device="cuda")
    gt = torch.as_tensor(g, device="cuda")

    def fwd(vec, seed):
        v = torch.as_tensor(vec, device="cuda")
        _lib.check(h.lib.drtvam_forward(h.h, v.data_ptr(), pixt.data_ptr(), pixt.numel(), 2, seed, 1.0, film.data_ptr(), 1,
                                        None))
        torch.cuda.synchronize()
        return film.cpu().numpy()

    _lib.check(h.lib.drtvam_scene_pin_pixels(h.h, pixt.data_ptr(), pixt.numel()))
    assert rel_l2(fwd(p, 5), f_p) < 1e-5
    _lib.check(h.lib.drtvam_adjoint(h.h, gt.data_ptr(), pixt.data_ptr(), pixt.numel(), 2, 77, 1.0, out.data_ptr(), None))
    assert rel_l2(out.cpu().numpy(), ref.adjoint(g, pix, 2, 77)) < 1e-5
    assert rel_l2(fwd(dvec, 5), f_d) < 1e-5             # table of the first pass, new pattern values
    assert rel_l2(fwd(dvec, 6), ref.forward(dvec, pix, 2, 6)) < 1e-5   # another seed: rebuilt
    # the same address pinned again (its contents may have changed): rebuilt as well
    _lib.check(h.lib.drtvam_scene_pin_pixels(h.h, pixt.data_ptr(), pixt.numel()))
    assert rel_l2(fwd(p, 6), ref.forward(p, pix, 2, 6)) < 1e-5


def test_nccl_entry_points_single_rank(workdir):
    """drtvam_comm_* with a one-rank communicator: libnccl is found at run time, the all-reduce is the
    identity, the asynchronous-error query is clean, and calls without a communicator fail loudly."""
    scene, integ, sensor, d, _ = build("C1", workdir, 16)
    h = Handle(d)
    x = torch.arange(1000, device="cuda", dtype=torch.float32)
    assert h.lib.drtvam_allreduce_dose(h.h, x.data_ptr(), x.numel(), None) == -1
    assert b"drtvam_comm_init" in h.lib.drtvam_last_error()
    uid = (C.c_char * 128)()
    _lib.check(h.lib.drtvam_comm_unique_id(uid))
    _lib.check(h.lib.drtvam_comm_init(h.h, uid, 0, 1))
    assert h.lib.drtvam_comm_init(h.h, uid, 0, 1) == -1          # one communicator per handle
    _lib.check(h.lib.drtvam_allreduce_dose(h.h, x.data_ptr(), x.numel(), None))
    s = torch.tensor([1.5, -2.0], device="cuda", dtype=torch.float64)
    _lib.check(h.lib.drtvam_allreduce_scalars(h.h, s.data_ptr(), 2, None))
    torch.cuda.synchronize()
    assert torch.equal(x, torch.arange(1000, device="cuda", dtype=torch.float32))
    assert s.tolist() == [1.5, -2.0]
    _lib.check(h.lib.drtvam_comm_destroy(h.h))


def test_streamed_render_of_a_scattering_scene(workdir):
    """The same for the scattering hybrid (plan 4): every block runs the planar per-lane kernels for its first
    segments and the tracer for the rest, both on the lanes' positions in the whole wavefront."""
    scene, integ, sensor, d, _ = build("C5", workdir, 48, n_patterns=10, device="cuda", spp=2)
    n = scene.projector.active_size()
    host = (torch.rand(n, generator=torch.Generator().manual_seed(5)) - 0.2).pin_memory()
    scene.projector.active_data = host.cuda()
    plain = integ.render(scene, sensor, seed=3, spp=2).clone()
    streamed = integ.render(scene, sensor, seed=3, spp=2, host_patterns=host, blocks=3)
    assert rel_l2(streamed.cpu().numpy(), plain.cpu().numpy()) < 1e-5


@pytest.mark.parametrize("regular", [True, False])
def test_streamed_render_equals_plain_render(regular, workdir):
    """render(host_patterns=...) uploads the patterns in blocks of projection angles under the projection of
    the previous block and accumulates the blocks (A = sum_k A_k).  Same dose as one upload followed by one
    projection - for jittered rays too, whose PCG32 streams are seeded with their position in the whole
    wavefront, not in the block."""
    import drtvam_b200 as dv
    scene, integ, sensor, d, _ = build("C3", workdir, 48, n_patterns=10, device="cuda", regular_sampling=regular)
    n = scene.projector.active_size()
    host = (torch.rand(n, generator=torch.Generator().manual_seed(5)) - 0.2).pin_memory()
    scene.projector.active_data = host.cuda()
    plain = integ.render(scene, sensor, seed=3, spp=2).clone()
    for blocks in (3, 10, 64):
        streamed = integ.render(scene, sensor, seed=3, spp=2, host_patterns=host, blocks=blocks)
        assert torch.equal(scene.projector.active_data.cpu(), host)
        assert rel_l2(streamed.cpu().numpy(), plain.cpu().numpy()) < (1e-6 if regular else 1e-5)
    # pageable memory: one upload, same result
    streamed = integ.render(scene, sensor, seed=3, spp=2, host_patterns=host.clone())
    assert rel_l2(streamed.cpu().numpy(), plain.cpu().numpy()) < 1e-6


@pytest.mark.parametrize("regular", [True, False])
def test_streamed_gradient_download_equals_plain_adjoint(regular, workdir):
    """render_backward(host_out=...) runs the adjoint in blocks of projection angles and copies the gradient
    of a block to the host under the adjoint of the next one (A^T g is block-diagonal in the angles).  Same
    gradient as one call; jittered configurations are not cut (their lanes are numbered through the whole
    call) but still deliver the host copy."""
    scene, integ, sensor, d, _ = build("C3", workdir, 48, n_patterns=10, device="cuda", regular_sampling=regular)
    n = scene.projector.active_size()
    g = torch.randn(scene.film_shape(sensor), generator=torch.Generator().manual_seed(7)).cuda()
    plain = integ.render_backward(scene, None, g, sensor, seed=3, spp=2).clone()
    for blocks in (1, 3, 10, 64):
        host = torch.full((n,), float("nan")).pin_memory()
        out = integ.render_backward(scene, None, g, sensor, seed=3, spp=2, host_out=host, blocks=blocks)
        torch.cuda.current_stream().synchronize()
        assert rel_l2(out.cpu().numpy(), plain.cpu().numpy()) < 1e-5
        assert torch.equal(host, out.cpu())
    # wait=False: the caller waits for the scene's copy stream itself
    host = torch.full((n,), float("nan")).pin_memory()
    out = integ.render_backward(scene, None, g, sensor, seed=3, spp=2, host_out=host, wait=False)
    torch.cuda.current_stream().wait_stream(scene.copy_stream())
    torch.cuda.current_stream().synchronize()
    assert torch.equal(host, out.cpu())
