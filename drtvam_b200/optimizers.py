"""First-order optimizers of the pattern vector.  The reference hands `config['optimizer']` of type 'adam' / 'sgd'
to Mitsuba's `mi.ad.Adam` / `mi.ad.SGD` (optimize.py:205-209); these are their update rules (Mitsuba 3.6
`python/ad/optimizers.py`) on torch tensors, with the dict-like access the optimisation loop uses
(`opt[key]`, `opt[key] = value`, `opt.step()`).  The updates are elementwise, so an angle- or slab-sharded pattern
vector needs no collective."""
from __future__ import annotations

import math

import torch


class _Optimizer:
    def __init__(self, lr, params=None):
        if lr is None or lr < 0:
            raise ValueError("'lr' must be >= 0")
        self.lr = float(lr)
        self.variables = {}
        self.state = {}
        if params:
            for k, v in params.items():
                self[k] = v

    def __contains__(self, key):
        return key in self.variables

    def __getitem__(self, key):
        return self.variables[key]

    def __setitem__(self, key, value):
        if not isinstance(value, torch.Tensor) or not value.is_floating_point():
            raise TypeError("Optimizer variables must be floating-point tensors")
        new = value.detach().clone() if value.requires_grad or value.grad_fn is not None else value.detach()
        new.requires_grad_(True)
        old = self.variables.get(key)
        self.variables[key] = new
        if old is None or old.shape != new.shape:
            self.reset(key)

    def keys(self):
        return self.variables.keys()

    def items(self):
        return self.variables.items()

    def set_learning_rate(self, lr):
        self.lr = float(lr)

    def reset(self, key):
        raise NotImplementedError

    def step(self):
        raise NotImplementedError


class SGD(_Optimizer):
    """value <- value - lr g, or with momentum: v <- momentum v + g, value <- value - lr v."""

    def __init__(self, lr, momentum=0.0, mask_updates=False, params=None):
        if momentum < 0 or momentum >= 1:
            raise ValueError("'momentum' must be in [0, 1)")
        self.momentum = float(momentum)
        self.mask_updates = bool(mask_updates)
        super().__init__(lr, params)

    def reset(self, key):
        self.state[key] = torch.zeros_like(self.variables[key]) if self.momentum != 0 else None

    @torch.no_grad()
    def step(self):
        for k, p in list(self.variables.items()):
            g = p.grad
            if g is None:
                continue
            if self.momentum == 0:
                value = p - self.lr * g
            else:
                v = self.momentum * self.state[k] + g
                if self.mask_updates:      # entries without a gradient keep their velocity and value
                    v = torch.where(g != 0, v, self.state[k])
                self.state[k] = v
                value = p - self.lr * torch.where(g != 0, v, torch.zeros_like(v)) if self.mask_updates else p - self.lr * v
            value = value.detach()
            value.requires_grad_(True)
            self.variables[k] = value


class Adam(_Optimizer):
    """m <- b1 m + (1 - b1) g, v <- b2 v + (1 - b2) g^2, value <- value - lr sqrt(1 - b2^t) / (1 - b1^t) m / (sqrt(v) + eps)."""

    def __init__(self, lr, beta_1=0.9, beta_2=0.999, epsilon=1e-8, mask_updates=False, uniform=False, params=None):
        if not (0 <= beta_1 < 1 and 0 <= beta_2 < 1):
            raise ValueError("'beta_1' and 'beta_2' must be in [0, 1)")
        if epsilon <= 0:
            raise ValueError("'epsilon' must be > 0")
        self.beta_1, self.beta_2, self.epsilon = float(beta_1), float(beta_2), float(epsilon)
        self.mask_updates, self.uniform = bool(mask_updates), bool(uniform)
        super().__init__(lr, params)

    def reset(self, key):
        p = self.variables[key]
        self.state[key] = (0, torch.zeros_like(p), torch.zeros_like(p))

    @torch.no_grad()
    def step(self):
        for k, p in list(self.variables.items()):
            g = p.grad
            if g is None:
                continue
            t, m, v = self.state[k]
            t += 1
            lr_t = self.lr * math.sqrt(1 - self.beta_2 ** t) / (1 - self.beta_1 ** t)
            m_new = self.beta_1 * m + (1 - self.beta_1) * g
            v_new = self.beta_2 * v + (1 - self.beta_2) * g * g
            if self.mask_updates:
                nz = g != 0
                m_new = torch.where(nz, m_new, m)
                v_new = torch.where(nz, v_new, v)
            self.state[k] = (t, m_new, v_new)
            denom = torch.sqrt(v_new.max() if self.uniform else v_new) + self.epsilon
            step = lr_t * m_new / denom
            if self.mask_updates:
                step = torch.where(g != 0, step, torch.zeros_like(step))
            value = (p - step).detach()
            value.requires_grad_(True)
            self.variables[k] = value
