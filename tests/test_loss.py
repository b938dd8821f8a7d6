"""Known answers of the reference's tests/test_loss.py:6-108, in torch."""
import torch

from drtvam_b200.loss import L2Loss, ThresholdedLoss


def T(vals, shape):
    return torch.tensor(vals, dtype=torch.float32).reshape(shape)


def run(loss_fn, pred_vals, pred_shape, target):
    pred = T(pred_vals, pred_shape).requires_grad_(True)
    loss = loss_fn(pred, target, 0 * target)
    loss.backward()
    return loss.item(), pred.grad.reshape(-1)


def close(a, b):
    return torch.allclose(torch.as_tensor(a, dtype=torch.float32), torch.as_tensor(b, dtype=torch.float32),
                          rtol=1e-5, atol=1e-7)


def test_l2():
    target = T([1, 1, 0, 0], (2, 2, 1))
    l, g = run(L2Loss({'reduction': 'sum'}), [1, 2, 3, 4], (2, 2, 1), target)
    assert l == 26 and close(g, [0, 2, 6, 8])
    l, g = run(L2Loss({'reduction': 'mean'}), [1, 2, 3, 4], (2, 2, 1), target)
    assert l == 6.5 and close(g, [0, 0.5, 1.5, 2])
    target = T([0.2, 0.8, 0.5, 0.], (2, 2, 1))
    l, g = run(L2Loss({'reduction': 'sum'}), [1., 1., 1., 1.], (2, 2, 1), target)
    assert close(l, 0.8 ** 2 + 0.2 ** 2 + 0.5 ** 2 + 1.) and close(g, [1.6, 0.4, 1., 2.])
    # surface-aware target
    target = T([0.2, 0.8], (1, 1, 2))
    l, g = run(L2Loss({'reduction': 'sum'}), [0.4, 0.3], (1, 1, 2), target)
    assert close(l, 0.2 * 0.6 ** 2 + 0.8 * 0.3 ** 2) and close(g, [-2 * 0.2 * 0.6, 2 * 0.8 * 0.3])


def test_thresholded():
    target = T([1, 1, 0, 0], (2, 2))
    P = [0.5, 0.97, 0.92, 0.5]
    l, g = run(ThresholdedLoss({'K': 2, 'tl': 0.9, 'tu': 0.95, 'reduction': 'sum'}), P, (2, 2, 1), target)
    assert close(l, 0.45 ** 2 + 0.02 ** 2) and close(g, [-0.9, 0., 0.04, 0.])
    l, g = run(ThresholdedLoss({'K': 2, 'tl': 0.9, 'tu': 0.95, 'reduction': 'mean'}), P, (2, 2, 1), target)
    assert close(l, (0.45 ** 2 + 0.02 ** 2) / 4) and close(g, [-0.225, 0., 0.01, 0.])
    l, g = run(ThresholdedLoss({'K': 1, 'tl': 0.9, 'tu': 0.95, 'reduction': 'sum'}), [0.5, 1.1, 0.92, 0.5], (2, 2, 1), target)
    assert close(l, 0.57) and close(g, [-1, 1., 1., 0.])
    l, g = run(ThresholdedLoss({'K': 2, 'tl': 0.4, 'tu': 0.95, 'reduction': 'sum'}), P, (2, 2, 1), target)
    assert close(l, 0.45 ** 2 + 0.52 ** 2 + 0.1 ** 2) and close(g, [-0.9, 0., 1.04, 0.2])
    l, g = run(ThresholdedLoss({'K': 2, 'tl': 0.9, 'tu': 0.99, 'reduction': 'sum'}), P, (2, 2, 1), target)
    assert close(l, 0.49 ** 2 + 0.02 ** 2 + 0.02 ** 2) and close(g, [-0.98, -0.04, 0.04, 0.])
    # surface-aware target
    target = T([0.2, 0.8, 2, 2], (2, 1, 2))
    l, g = run(ThresholdedLoss({'K': 2, 'tl': 0.9, 'tu': 0.95, 'reduction': 'sum'}), [0.2, 0.1, 0.96, 0.92], (2, 1, 2), target)
    assert close(l, 0.2 * 0.75 ** 2 + 0.5 * 0.02 ** 2) and close(g, [-2 * 0.2 * 0.75, 0., 0., 2 * 0.5 * 0.02])


def test_invalid():
    import pytest
    with pytest.raises(ValueError):
        ThresholdedLoss({'tl': 0.95, 'tu': 0.9})
    with pytest.raises(ValueError):
        L2Loss({'reduction': 'median'})
    with pytest.raises(ValueError):
        L2Loss({})(torch.zeros(2, 2, 1), torch.zeros(3, 2, 1), torch.zeros(1))
