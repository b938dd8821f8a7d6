"""L-BFGS for a *linear* forward model.  Mirrors ``LinearLBFGS`` of
``src/drtvam/lbfgs.py:146-275``: history update (:176-196), two-loop recursion
(:213-234), one extra forward projection of the search direction (:239-249)
and an Armijo back-tracking search that re-evaluates only the loss thanks to
``vol(p + a d) = vol + a dvol`` (:256-266).  torch tensors; ``dot`` can be
swapped for an all-reducing dot when the patterns are angle-sharded.
"""
from __future__ import annotations

import torch

from .loss import LazyAxpy


class Optimizer:
    """Dict-like holder of optimised variables (the role of ``mi.ad.Optimizer``)."""

    def __init__(self, lr=1.0, params=None):
        self.lr = lr
        self.variables = {}
        if params is not None:
            for k, v in params.items():
                self[k] = v

    def __setitem__(self, key, value):
        v = value.detach().clone() if value.requires_grad or value.grad_fn is not None else value.detach()
        v.requires_grad_(True)
        self.variables[key] = v

    def __getitem__(self, key):
        return self.variables[key]

    def __contains__(self, key):
        return key in self.variables

    def keys(self):
        return self.variables.keys()

    def items(self):
        return self.variables.items()


class LinearLBFGS(Optimizer):
    def __init__(self, lr=1.0, m=5, params=None, render_fn=None, loss_fn=None, search_it=20, dot=None):
        super().__init__(lr, params)
        self.m = m                # history size
        self.p_old, self.g_old = {}, {}
        self.s, self.y, self.ys, self.t = {}, {}, {}, {}
        self.render_fn = render_fn
        self.loss_fn = loss_fn
        self.search_it = search_it  # back-tracking iterations
        self.dot = dot if dot is not None else (lambda a, b: torch.dot(a.reshape(-1), b.reshape(-1)))
        self.last_alpha = None
        self.last_trials = 0

    def reset(self, k):
        self.s[k], self.y[k], self.ys[k] = [], [], []
        self.t[k] = 0

    def update_history(self, k):
        if self.t[k] > self.m:  # discard the oldest pair
            self.s[k].pop(0)
            self.y[k].pop(0)
            self.ys[k].pop(0)
        p = self.variables[k].detach().reshape(-1)
        g = self.variables[k].grad
        if g is None:
            raise RuntimeError(f"variable '{k}' has no gradient; call loss.backward() before step()")
        g = g.detach().reshape(-1).clone()
        if self.t[k] > 0:
            s = p - self.p_old[k]
            y = g - self.g_old[k]
            self.s[k].append(s)
            self.y[k].append(y)
            self.ys[k].append(self.dot(y, s))
        self.p_old[k] = p.clone()
        self.g_old[k] = g
        self.t[k] += 1

    @torch.no_grad()
    def _direction(self, k):
        """Two-loop recursion; returns -H g."""
        s, y, ys = self.s[k], self.y[k], self.ys[k]
        q = self.g_old[k].clone()
        n = len(s)
        alphas = [None] * n
        for i in range(n - 1, -1, -1):
            a = self.dot(s[i], q) / ys[i]
            q = q - a * y[i]
            alphas[i] = a
        gamma = 1.0 if self.t[k] == 1 else ys[-1] / self.dot(y[-1], y[-1])
        z = gamma * q
        for i in range(n):
            b = self.dot(y[i], z) / ys[i]
            z = z + (alphas[i] - b) * s[i]
        return -z

    def step(self, vol, loss):
        search_dirs = {}
        with torch.no_grad():
            for k in self.variables:
                if k not in self.s:
                    self.reset(k)
                self.update_history(k)
                search_dirs[k] = self._direction(k)

            # one forward projection of the (signed) search direction
            params = {k: search_dirs[k].reshape(self.variables[k].shape) for k in self.variables}
            dvol = self.render_fn(params)

            g_dot_z = 0.0
            for k in params:
                g_dot_z = g_dot_z + self.dot(self.g_old[k], search_dirs[k])

            alpha = 1.0
            c1 = 1e-4
            vol = vol.detach()
            loss = loss.detach()
            trials = 0
            for _ in range(self.search_it):
                trials += 1
                # note: the pattern term sees the *direction*, as in lbfgs.py:260
                y_trial = LazyAxpy(vol, dvol, alpha) if getattr(self.loss_fn, 'accepts_lazy', False) \
                    else vol + alpha * dvol
                f_new = self.loss_fn(y_trial, params['projector.active_data']
                                     if 'projector.active_data' in params else next(iter(params.values())))
                if bool((f_new <= loss + c1 * alpha * g_dot_z).all()):
                    break
                alpha *= 0.5
            self.last_alpha, self.last_trials = alpha, trials

            for k, p in list(self.variables.items()):
                new = (p.detach() + alpha * params[k]).detach()
                new.requires_grad_(True)
                self.variables[k] = new
