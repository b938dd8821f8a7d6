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
