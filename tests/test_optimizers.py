"""Adam / SGD of the pattern vector: Mitsuba's update rules (mi.ad.Adam / mi.ad.SGD, which the reference builds
from config['optimizer'] at optimize.py:205-209), checked against the formulas written out by hand, and one
GPU optimisation per optimizer through optimize()."""
import copy
import math
import os

import numpy as np
import pytest
import torch

from drtvam_b200.optimizers import SGD, Adam


def _grad_step(opt, key, g):
    p = opt[key]
    (p * torch.as_tensor(g, dtype=p.dtype)).sum().backward()
    opt.step()


def test_sgd_and_momentum():
    opt = SGD(lr=0.1)
    opt['x'] = torch.tensor([1.0, 2.0], dtype=torch.float64)
    _grad_step(opt, 'x', [1.0, -2.0])
    assert torch.allclose(opt['x'].detach(), torch.tensor([0.9, 2.2], dtype=torch.float64))
    opt = SGD(lr=0.1, momentum=0.5)
    opt['x'] = torch.tensor([1.0], dtype=torch.float64)
    _grad_step(opt, 'x', [1.0])          # v = 1      -> 0.9
    _grad_step(opt, 'x', [1.0])          # v = 1.5    -> 0.75
    assert abs(float(opt['x']) - 0.75) < 1e-12
    with pytest.raises(ValueError):
        SGD(lr=0.1, momentum=1.0)


def test_adam_follows_mitsubas_rule():
    lr, b1, b2, eps = 0.05, 0.9, 0.999, 1e-8
    opt = Adam(lr=lr)
    x0 = torch.tensor([0.3, -0.7, 2.0], dtype=torch.float64)
    opt['x'] = x0
    m = np.zeros(3)
    v = np.zeros(3)
    x = x0.numpy().copy()
    rng = np.random.default_rng(0)
    for t in range(1, 6):
        g = rng.standard_normal(3)
        _grad_step(opt, 'x', g)
        m = b1 * m + (1 - b1) * g
        v = b2 * v + (1 - b2) * g * g
        x = x - lr * math.sqrt(1 - b2 ** t) / (1 - b1 ** t) * m / (np.sqrt(v) + eps)
        assert np.allclose(opt['x'].detach().numpy(), x, rtol=0, atol=1e-14)
    # assigning a clamped value of the same shape keeps the moments (the loop clamps after every step)
    opt['x'] = torch.clamp_min(opt['x'].detach(), 0)
    assert opt.state['x'][0] == 5
    # masked updates leave entries without a gradient alone
    opt = Adam(lr=0.1, mask_updates=True)
    opt['x'] = torch.tensor([1.0, 1.0], dtype=torch.float64)
    _grad_step(opt, 'x', [1.0, 0.0])
    assert float(opt['x'][1]) == 1.0 and float(opt['x'][0]) < 1.0


@pytest.mark.gpu
@pytest.mark.parametrize("optimizer", [{"type": "adam", "lr": 0.2}, {"type": "sgd", "lr": 0.2, "momentum": 0.5}])
def test_first_order_optimizers_through_optimize(optimizer, workdir):
    from drtvam_b200.optimize import optimize
    from drtvam_b200.synthetic import make_config
    cfg = make_config("C2", workdir, n=32, n_steps=25)
    cfg["optimizer"] = dict(optimizer)
    cfg["output"] = os.path.join(workdir, "out_" + optimizer["type"])
    cfg["save_pattern_images"] = False
    vol = optimize(cfg, progress=False)
    loss = np.load(os.path.join(cfg["output"], "loss.npy"))
    assert torch.isfinite(vol).all() and np.isfinite(loss).all()
    assert loss[-1] < (0.5 if optimizer["type"] == "adam" else 0.8) * loss[0], loss
    with pytest.raises(ValueError):
        cfg["optimizer"] = {"type": "newton"}
        optimize(cfg, progress=False)


@pytest.mark.gpu
def test_psf_analysis_projects_the_listed_pixels(workdir):
    """optimize.py:245-264: with a `psf_analysis` list in the config nothing is optimised; the listed single
    pixels are projected at their intensities (a sparse pixel list, so the general tracer) and the result is
    linear in them."""
    from drtvam_b200.optimize import optimize
    from drtvam_b200.synthetic import make_config
    cfg = make_config("C3", workdir, n=32)
    cfg["save_pattern_images"] = False
    entries = [{"x": 16, "y": 16, "index_pattern": 0, "intensity": 1.0},
               {"x": 10, "y": 20, "index_pattern": 8, "intensity": 2.0}]
    vols = []
    for k, sub in enumerate((entries, entries[:1], entries[1:])):
        c = copy.deepcopy(cfg)          # load_scene pops the sensor's scale keys, as the reference does
        c.update(psf_analysis=sub, output=os.path.join(workdir, f"out_psf{k}"))
        vols.append(optimize(c, progress=False).cpu().numpy())
        pat = np.load(os.path.join(c["output"], "patterns.npz"))["patterns"]
        assert np.count_nonzero(pat) == len(sub)
    assert pat[8, 20, 10] == 2.0
    assert vols[0].sum() > 0 and np.count_nonzero(vols[1]) < 0.05 * vols[1].size      # one thin beam
    assert np.allclose(vols[0], vols[1] + vols[2], rtol=0, atol=1e-6 * vols[0].max())
