"""LinearLBFGS (lbfgs.py:146-275) on a small linear least-squares problem."""
import torch

from drtvam_b200.lbfgs import LinearLBFGS


def test_linear_lbfgs_converges_on_quadratic():
    torch.manual_seed(0)
    A = torch.randn(40, 12, dtype=torch.float64)
    x_true = torch.rand(12, dtype=torch.float64)
    b = A @ x_true
    key = 'projector.active_data'

    def render_fn(v):
        return A @ v[key]

    def loss_fn(y, patterns):
        return torch.sum((y - b) ** 2)

    opt = LinearLBFGS(render_fn=render_fn, loss_fn=loss_fn)
    opt[key] = torch.zeros(12, dtype=torch.float64)
    losses = []
    for i in range(25):
        x = opt[key]
        vol = A @ x
        loss = loss_fn(vol, x)
        losses.append(loss.item())
        loss.backward()
        opt.step(vol.detach(), loss.detach())
    assert losses[-1] < 1e-12 * losses[0]
    assert torch.allclose(opt[key].detach(), x_true, atol=1e-6)
    # history is capped at m pairs (lbfgs.py:176-181)
    assert len(opt.s[key]) == opt.m and opt.t[key] == 25
    # first step uses gamma = 1 and must satisfy Armijo
    assert all(b_ <= a_ + 1e-12 for a_, b_ in zip(losses, losses[1:]))


def test_vanishing_pairs_do_not_poison_the_direction():
    """Once a thresholded loss has reached exactly 0 the gradient vanishes, and with it s and y of the next pair:
    y.s = 0.  The reference takes 1 / (y.s) unguarded (lbfgs.py:221-222) and continues on NaNs; here such a pair
    is left out and the iterate stays finite (and where y.s != 0 the arithmetic is the reference's)."""
    key = 'projector.active_data'

    def render_fn(v):
        return v[key] * 2.0

    def loss_fn(y, patterns):
        return torch.sum(torch.clamp_min(1.0 - y, 0) ** 2)     # exactly 0 once y >= 1 everywhere

    opt = LinearLBFGS(render_fn=render_fn, loss_fn=loss_fn)
    opt[key] = torch.zeros(8, dtype=torch.float64)
    for i in range(12):
        x = opt[key]
        vol = render_fn({key: x})
        loss = loss_fn(vol, x)
        loss.backward()
        opt.step(vol.detach(), loss.detach())
        assert torch.isfinite(opt[key]).all(), i
    assert float(loss_fn(render_fn({key: opt[key].detach()}), None)) == 0.0


def test_step_requires_gradient():
    import pytest
    opt = LinearLBFGS(render_fn=lambda v: v['k'], loss_fn=lambda y, p: y.sum())
    opt['k'] = torch.zeros(3)
    with pytest.raises(RuntimeError):
        opt.step(torch.zeros(3), torch.tensor(0.))


import pytest


@pytest.mark.gpu
@pytest.mark.parametrize("sparsity", [0.0, 0.05])
def test_device_lbfgs_follows_the_eager_loop(sparsity):
    """The device path (csrc/lbfgs_kernels.cu: Gram matrix + recursion on coefficients, all Armijo trials
    in one sweep, step length chosen on the device) against the eager restatement of lbfgs.py:198-275 on the
    same linear problem with the thresholded loss: same step lengths and trial counts, iterates equal to
    fp32 rounding while the history is short, the same loss curve afterwards."""
    from drtvam_b200.loss import ThresholdedLoss
    torch.manual_seed(3)
    dev = torch.device("cuda")
    n, shape = 6000, (10, 20, 24, 1)
    A = (torch.rand(4800, n, device=dev) < 0.02).float() * torch.rand(4800, n, device=dev) * 0.05
    target = (torch.rand(shape[:-1], device=dev) > 0.6).float()
    key = 'projector.active_data'
    loss_obj = ThresholdedLoss({"weight_sparsity": sparsity})

    def make(device_path):
        def render_fn(v):
            return (A @ v[key].reshape(-1)).reshape(shape)

        def loss_fn(y, patterns):
            return loss_obj(y, target, patterns)
        loss_fn.accepts_lazy = True
        if device_path:
            loss_fn.armijo_sweep = lambda x, dx, direction, *a: loss_obj.armijo_sweep(x, dx, target, direction, *a)
        opt = LinearLBFGS(render_fn=render_fn, loss_fn=loss_fn)
        opt[key] = torch.full((n,), 0.3, device=dev)
        return opt, render_fn

    runs = []
    for device_path in (True, False):
        opt, render_fn = make(device_path)
        hist = []
        for i in range(14):
            p = opt[key]
            vol = render_fn({key: p})
            loss = loss_obj(vol, target, p)
            loss.backward()
            opt.step(vol.detach(), loss.detach())
            opt[key] = torch.clamp_min(opt[key].detach(), 0)
            hist.append((float(loss), opt.last_alpha, opt.last_trials, opt[key].detach().clone()))
        assert (len(opt._dev) == 1) == device_path          # the path that was meant to run did run
        runs.append(hist)
    dev_run, eager = runs
    assert dev_run[-1][0] < 0.5 * dev_run[0][0]
    for i, (a, b) in enumerate(zip(dev_run, eager)):
        if i < 4:
            assert a[1] == b[1] and a[2] == b[2]
            err = float(torch.linalg.norm(a[3] - b[3]) / torch.linalg.norm(b[3]))
            assert err < 2e-5, (i, err)
        assert abs(a[0] - b[0]) < 2e-2 * abs(b[0]) + 1e-6, (i, a[0], b[0])


@pytest.mark.gpu
def test_device_direction_leaves_vanishing_pairs_out():
    """drtvam_lbfgs_direction on a history whose newest pair is (0, 0) - the state after a step at loss 0 - and on
    a NaN dose in the fused losses: coefficients stay finite (the direction falls back to -g from the usable
    pairs), and a NaN dose gives a NaN loss as relu(NaN) does in the reference, not the 0 of fmaxf."""
    import ctypes as C
    from drtvam_b200 import _lib
    from drtvam_b200.loss import ThresholdedLoss
    lib = _lib.load()
    dev = torch.device("cuda")
    n = 4096
    g = torch.randn(n, device=dev)
    s0, y0 = torch.randn(n, device=dev) * 0.1, torch.randn(n, device=dev) * 0.1
    if float(torch.dot(s0, y0)) < 0:
        y0 = -y0
    zero = torch.zeros(n, device=dev)
    vecs = [s0, zero, y0, zero.clone(), g]               # s_0, s_1 = 0, y_0, y_1 = 0, g
    table = (C.c_void_p * 5)(*[v.data_ptr() for v in vecs])
    gram = torch.zeros(64, device=dev, dtype=torch.float64)
    coef = torch.zeros(32, device=dev, dtype=torch.float64)
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.drtvam_lbfgs_direction(0, table, 2, n, gram.data_ptr(), 0, coef.data_ptr(), st))
    d = torch.empty(n, device=dev)
    _lib.check(lib.drtvam_lbfgs_combine(0, table, 2, coef.data_ptr(), n, d.data_ptr(), st))
    torch.cuda.synchronize()
    assert torch.isfinite(coef[:5]).all() and torch.isfinite(d).all()
    # the same history without the vanishing pair gives the same direction
    table1 = (C.c_void_p * 3)(s0.data_ptr(), y0.data_ptr(), g.data_ptr())
    coef1 = torch.zeros(32, device=dev, dtype=torch.float64)
    _lib.check(lib.drtvam_lbfgs_direction(0, table1, 1, n, gram.data_ptr(), 0, coef1.data_ptr(), st))
    d1 = torch.empty(n, device=dev)
    _lib.check(lib.drtvam_lbfgs_combine(0, table1, 1, coef1.data_ptr(), n, d1.data_ptr(), st))
    # (gamma comes from the newest pair: unusable -> 1 in the first case, y_0.s_0 / y_0.y_0 in the second)
    gamma1 = float(torch.dot(y0, s0) / torch.dot(y0, y0))
    assert torch.isfinite(d1).all() and gamma1 > 0
    # NaN dose -> NaN loss
    loss = ThresholdedLoss({})
    x = torch.rand(8, 8, 8, 1, device=dev)
    x[3, 3, 3, 0] = float("nan")
    t = (torch.rand(8, 8, 8, device=dev) > 0.5).float()
    assert torch.isnan(loss(x, t, torch.zeros(3, device=dev)))
