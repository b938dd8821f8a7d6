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
    def __init__(self, lr=1.0, m=5, params=None, render_fn=None, loss_fn=None, search_it=20, dot=None, allsum=None,
                 consensus=None):
        super().__init__(lr, params)
        self.m = m                # history size
        self.p_old, self.g_old = {}, {}
        self.s, self.y, self.ys, self.t = {}, {}, {}, {}
        self.render_fn = render_fn
        self.loss_fn = loss_fn
        self.search_it = search_it  # back-tracking iterations
        self._default_dot = dot is None
        self.dot = dot if dot is not None else (lambda a, b: torch.dot(a.reshape(-1), b.reshape(-1)))
        # sum of a small tensor over the ranks that share the (sharded) pattern vector: with it the device
        # path needs ONE collective per iteration (the Gram matrix) instead of one per dot product
        self.allsum = allsum
        # rank-mean of redundantly computed scalars, so that every rank takes the same Armijo decision
        self.consensus = consensus
        self._alpha = None        # python float, or a device tensor {alpha, trials, loss} of the last step
        self._dev = {}            # device-path state per variable

    @property
    def last_alpha(self):
        a = self._alpha
        return float(a[0]) if isinstance(a, torch.Tensor) else a

    @property
    def last_trials(self):
        a = self._alpha
        return int(a[1]) if isinstance(a, torch.Tensor) else getattr(self, '_trials', 0)

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
        # a pair with y.s = 0 (both vanish when the loss has all but reached 0) would give an infinite 1 / (y.s) and
        # a NaN direction, as it does in the reference (lbfgs.py:221-222); it carries no curvature and is left out
        use = [bool(torch.isfinite(torch.as_tensor(v)).all()) and float(v) != 0.0 for v in ys]
        for i in range(n - 1, -1, -1):
            if not use[i]:
                continue
            a = self.dot(s[i], q) / ys[i]
            q = q - a * y[i]
            alphas[i] = a
        gamma = 1.0
        if self.t[k] != 1 and use[-1]:
            yy = self.dot(y[-1], y[-1])
            if float(yy) != 0.0:
                gamma = ys[-1] / yy
        z = gamma * q
        for i in range(n):
            if not use[i]:
                continue
            b = self.dot(y[i], z) / ys[i]
            z = z + (alphas[i] - b) * s[i]
        return -z

    # -- device path ----------------------------------------------------------------------------------
    def _device_ok(self, vol):
        """One CUDA fp32 variable, a loss that can sweep all Armijo trials at once, and - when the pattern
        vector is sharded - a way to sum the Gram matrix over ranks."""
        if len(self.variables) != 1 or not 1 <= self.m <= 6 or not 1 <= self.search_it <= 32:
            return False
        v = next(iter(self.variables.values()))
        if not (v.is_cuda and v.dtype == torch.float32 and v.grad is not None and v.grad.is_cuda):
            return False
        if not (self._default_dot or self.allsum is not None):
            return False
        return getattr(self.loss_fn, 'armijo_sweep', None) is not None and isinstance(vol, torch.Tensor) and vol.is_cuda

    @torch.no_grad()
    def _step_device(self, vol, loss):
        """lbfgs.py:198-275 as a queue of kernels (csrc/lbfgs_kernels.cu): no host round trip, every history
        vector read twice per iteration (Gram matrix, combination) instead of ~13 times."""
        import ctypes as C
        from . import _lib
        lib = _lib.load()
        k, var = next(iter(self.variables.items()))
        dev = var.device
        st = torch.cuda.current_stream(dev).cuda_stream
        di = dev.index or 0
        p = var.detach().reshape(-1).contiguous()
        g = var.grad.detach().reshape(-1).to(torch.float32).contiguous()
        n = p.numel()
        S = self._dev.get(k)
        if S is None or S['n'] != n:
            S = {'n': n, 't': 0, 's': [], 'y': [], 'p_old': torch.empty_like(p), 'g_old': torch.empty_like(p),
                 'gram': torch.empty(13 * 13, device=dev, dtype=torch.float64),
                 'coef': torch.empty(16, device=dev, dtype=torch.float64),
                 'out': torch.zeros(4, device=dev, dtype=torch.float32), 'd': torch.empty_like(p),
                 'sums': torch.zeros(32, device=dev, dtype=torch.float64)}
            self._dev[k] = S
        # ---- history (lbfgs.py:176-196)
        first = S['t'] == 0
        s_new = y_new = None
        if not first:
            if S['t'] > self.m:      # discard the oldest pair, reuse its storage
                s_new, y_new = S['s'].pop(0), S['y'].pop(0)
            else:
                s_new, y_new = torch.empty_like(p), torch.empty_like(p)
            S['s'].append(s_new)
            S['y'].append(y_new)
        _lib.check(lib.drtvam_lbfgs_history(di, p.data_ptr(), g.data_ptr(), S['p_old'].data_ptr(), S['g_old'].data_ptr(),
                                            s_new.data_ptr() if s_new is not None else None,
                                            y_new.data_ptr() if y_new is not None else None, n, int(first), st))
        S['t'] += 1
        # ---- direction (lbfgs.py:213-234): Gram matrix of [s.., y.., g], recursion on coefficients, combination
        h = len(S['s'])
        vecs = S['s'] + S['y'] + [S['g_old']]
        table = (C.c_void_p * len(vecs))(*[v.data_ptr() for v in vecs])
        if self.allsum is None:
            _lib.check(lib.drtvam_lbfgs_direction(di, table, h, n, S['gram'].data_ptr(), 0, S['coef'].data_ptr(), st))
        else:
            _lib.check(lib.drtvam_lbfgs_direction(di, table, h, n, S['gram'].data_ptr(), 2, S['coef'].data_ptr(), st))
            nn = (2 * h + 1) ** 2
            S['gram'][:nn] = self.allsum(S['gram'][:nn])
            _lib.check(lib.drtvam_lbfgs_direction(di, table, h, n, S['gram'].data_ptr(), 1, S['coef'].data_ptr(), st))
        d = S['d']
        _lib.check(lib.drtvam_lbfgs_combine(di, table, h, S['coef'].data_ptr(), n, d.data_ptr(), st))
        # ---- one forward projection of the direction, then every Armijo trial in one sweep
        params = {k: d.reshape(var.shape)}
        dvol = self.render_fn(params)
        # stages of 4 and then up to 16 trials: alpha = 1 is accepted in almost every iteration, and a stage
        # that finds the search over (S['out'][3]) returns in its first instruction
        f0 = loss.detach().to(torch.float64).reshape(1)
        if self.consensus is not None:
            f0 = self.consensus(f0)
        S['sums'].zero_()
        first = 0
        for count in (4, 16, 12):
            count = min(count, self.search_it - first)
            if count <= 0:
                break
            stage = self.loss_fn.armijo_sweep(vol.detach(), dvol, d, first, count, S['sums'],
                                              S['out'][3:4] if first else None)
            if stage is None:
                if first:
                    raise RuntimeError("the fused Armijo sweep became unavailable between two stages")
                return None
            scale, term = stage
            if self.consensus is not None:
                S['sums'][first:first + count] = self.consensus(S['sums'][first:first + count])
            _lib.check(lib.drtvam_armijo_select(di, S['sums'].data_ptr(), first, first + count, float(scale), f0.data_ptr(),
                                                term.data_ptr() if term is not None else None,
                                                S['coef'][13:14].data_ptr(), 1e-4, S['out'].data_ptr(), st))
            first += count
        new = torch.empty_like(p)
        _lib.check(lib.drtvam_lbfgs_update(di, p.data_ptr(), d.data_ptr(), S['out'].data_ptr(), n, new.data_ptr(), st))
        self._alpha = S['out'].clone()
        new = new.reshape(var.shape)
        new.requires_grad_(True)
        self.variables[k] = new
        return True

    def step(self, vol, loss):
        if self._device_ok(vol):
            with torch.no_grad():
                if self._step_device(vol, loss):
                    return
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
            if self.consensus is not None:
                loss = self.consensus(loss)
            trials = 0
            for _ in range(self.search_it):
                trials += 1
                # note: the pattern term sees the *direction*, as in lbfgs.py:260
                y_trial = LazyAxpy(vol, dvol, alpha) if getattr(self.loss_fn, 'accepts_lazy', False) \
                    else vol + alpha * dvol
                f_new = self.loss_fn(y_trial, params['projector.active_data']
                                     if 'projector.active_data' in params else next(iter(params.values())))
                if self.consensus is not None:
                    f_new = self.consensus(f_new.detach())
                if bool((f_new <= loss + c1 * alpha * g_dot_z).all()):
                    break
                alpha *= 0.5
            self._alpha, self._trials = alpha, trials

            for k, p in list(self.variables.items()):
                new = (p.detach() + alpha * params[k]).detach()
                new.requires_grad_(True)
                self.variables[k] = new
