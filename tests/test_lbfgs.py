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
