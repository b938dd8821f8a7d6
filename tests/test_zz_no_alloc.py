"""drtvam_forward_noalloc: the deterministic planar forward on the static scratch of the library image
(include/drtvam_b200.h; no counterpart in the reference - an entry point for callers that cannot call
cudaMalloc).  Checked through the C-ABI against drtvam_forward, which keeps the same scratch in a
cudaMalloc'ed buffer, on two calls in a row that share the static array; the size limit and the
"tables first" rule are checked too.  This file is collected last on purpose."""
import numpy as np
import pytest
import torch

from drtvam_b200 import _lib
from drtvam_b200._scene import create_handle
from helpers import build, rel_l2

pytestmark = pytest.mark.gpu

TOL = 1e-5   # the two entry points run the same kernels; only the atomics of the frozen-axis rays reorder
CAPACITY_FLOATS = 4 << 20   # include/drtvam_b200.h: 4 Mi floats = 16 MiB
E_INVALID = -1              # DRTVAM_E_INVALID


class Handle:
    def __init__(self, desc):
        self.lib = _lib.load()
        self.h, self.keep = create_handle(desc, 0, desc.target_triangles)
        self.film_shape = desc.film_shape()

    def call(self, fn, p, pix, film=None, zero_first=1):
        if film is None:
            film = torch.empty(self.film_shape, device="cuda", dtype=torch.float32)
        rc = fn(self.h, p.data_ptr(), pix.data_ptr(), p.numel(), 1, 0, 1.0, film.data_ptr(), zero_first, None)
        return rc, film

    def close(self):
        if self.h:
            self.lib.drtvam_scene_destroy(self.h)
            self.h = None


def _pixels(desc, k0, nk):
    hw = desc.res_x * desc.res_y
    return torch.arange(k0 * hw, (k0 + nk) * hw, device="cuda", dtype=torch.int64).to(torch.int32)


def _free_bytes():
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info()[0]


def test_noalloc_forward_matches_forward_on_two_calls_in_a_row(workdir):
    n = 64
    scene, integ, sensor, d, _ = build("C3", workdir, n)
    h = Handle(d)
    lib = h.lib
    pix = _pixels(d, 0, n)
    rng = np.random.default_rng(5)
    p1 = torch.as_tensor((rng.random(pix.numel()) - 0.3).astype(np.float32), device="cuda")
    p2 = torch.as_tensor((rng.random(pix.numel()) + 0.1).astype(np.float32), device="cuda")

    # before the tables exist the call refuses (building them allocates)
    rc, _ = h.call(lib.drtvam_forward_noalloc, p1, pix)
    assert rc == E_INVALID and b"drtvam_scene_plan" in lib.drtvam_last_error()

    assert lib.drtvam_scene_plan(h.h, pix.numel(), 1) == 2
    assert (n * d.res_x * ((n + 31) // 32 * 32)) <= CAPACITY_FLOATS
    rc, ref1 = h.call(lib.drtvam_forward, p1, pix)
    _lib.check(rc)
    rc, ref2 = h.call(lib.drtvam_forward, p2, pix)
    _lib.check(rc)

    out1 = torch.empty(h.film_shape, device="cuda", dtype=torch.float32)
    out2 = torch.empty(h.film_shape, device="cuda", dtype=torch.float32)
    rc, _ = h.call(lib.drtvam_forward_noalloc, p1, pix, out1)   # first use on the device resolves the address
    _lib.check(rc)
    free0 = _free_bytes()
    rc, _ = h.call(lib.drtvam_forward_noalloc, p2, pix, out2)   # second call in a row: same static array
    _lib.check(rc)
    free1 = _free_bytes()
    assert free1 == free0, "the call changed the free device memory"
    e1, e2 = rel_l2(out1.cpu().numpy(), ref1.cpu().numpy()), rel_l2(out2.cpu().numpy(), ref2.cpu().numpy())
    print("noalloc vs drtvam_forward:", e1, e2)
    assert e1 < TOL and e2 < TOL
    assert float(ref1.abs().sum()) > 0 and rel_l2(out2.cpu().numpy(), ref1.cpu().numpy()) > 1e-2   # not a stale result

    # a third call with the first input again: the array holds the second call's repack until it is overwritten
    rc, out3 = h.call(lib.drtvam_forward_noalloc, p1, pix)
    _lib.check(rc)
    assert rel_l2(out3.cpu().numpy(), ref1.cpu().numpy()) < TOL

    # an angle shard accumulated onto another one (zero_first = 0), as a caller splits an oversized call
    half = n // 2
    hw = d.res_x * d.res_y
    acc = torch.empty(h.film_shape, device="cuda", dtype=torch.float32)
    rc, _ = h.call(lib.drtvam_forward_noalloc, p1[:half * hw], pix[:half * hw], acc, 1)
    _lib.check(rc)
    rc, _ = h.call(lib.drtvam_forward_noalloc, p1[half * hw:], pix[half * hw:], acc, 0)
    _lib.check(rc)
    # (the ray weight, pixel_area * n_active / (n_active * spp), does not depend on the split)
    assert rel_l2(acc.cpu().numpy(), ref1.cpu().numpy()) < TOL

    # empty call: clears the film, nothing else
    rc = lib.drtvam_forward_noalloc(h.h, None, None, 0, 1, 0, 1.0, acc.data_ptr(), 1, None)
    _lib.check(rc)
    assert float(acc.abs().max()) == 0.0
    h.close()


def test_noalloc_forward_rejects_sizes_beyond_the_static_scratch(workdir):
    n = 176   # 176 angles x 176 columns x 192 padded layers = 5.9 M floats > 4 Mi
    scene, integ, sensor, d, _ = build("C2", workdir, n)
    h = Handle(d)
    lib = h.lib
    pix = _pixels(d, 0, n)
    assert lib.drtvam_scene_plan(h.h, pix.numel(), 1) == 2
    p = torch.rand(pix.numel(), device="cuda")
    assert n * d.res_x * 192 > CAPACITY_FLOATS
    rc, _ = h.call(lib.drtvam_forward_noalloc, p, pix)
    assert rc == E_INVALID
    msg = lib.drtvam_last_error()
    assert b"exceed the static maximum" in msg and b"split the call by angles" in msg
    # the split the message asks for gives the result of the allocating entry point
    rc, ref = h.call(lib.drtvam_forward, p, pix)
    _lib.check(rc)
    hw = d.res_x * d.res_y
    acc = torch.empty(h.film_shape, device="cuda", dtype=torch.float32)
    for i, k0 in enumerate(range(0, n, 88)):
        sl = slice(k0 * hw, (k0 + 88) * hw)
        rc, _ = h.call(lib.drtvam_forward_noalloc, p[sl], pix[sl], acc, 1 if i == 0 else 0)
        _lib.check(rc)
    assert rel_l2(acc.cpu().numpy(), ref.cpu().numpy()) < TOL
    h.close()


def test_noalloc_forward_refuses_other_plans(workdir):
    scene, integ, sensor, d, _ = build("C3", workdir, 32, regular_sampling=False)   # jittered: plan 3
    h = Handle(d)
    pix = _pixels(d, 0, 32)
    assert h.lib.drtvam_scene_plan(h.h, pix.numel(), 1) == 3
    rc, _ = h.call(h.lib.drtvam_forward_noalloc, torch.rand(pix.numel(), device="cuda"), pix)
    assert rc == E_INVALID and b"deterministic planar plan" in h.lib.drtvam_last_error()
    h.close()
