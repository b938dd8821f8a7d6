"""Objective functions on the dose volume.  Mirrors ``src/drtvam/loss.py``:
``Loss.__call__`` shape handling and surface-aware weighting (:28-59),
``L2Loss`` (:63-80), ``ThresholdedLoss`` (:82-132), registry (:135-138).
torch elementwise + reductions; autograd supplies d loss / d dose.
"""
from __future__ import annotations

import torch


def relu(x):
    return torch.clamp_min(x, 0)


class LazyAxpy:
    """`x + alpha * dx` that has not been formed yet.  LinearLBFGS hands it to a
    loss that declares `accepts_lazy` for the Armijo trials (lbfgs.py:258), so
    the fused CUDA loss can read both volumes once instead of materialising
    their sum; any other loss gets `materialize()`."""

    def __init__(self, x, dx, alpha):
        self.x, self.dx, self.alpha = x, dx, float(alpha)

    @property
    def shape(self):
        return self.x.shape

    def dim(self):
        return self.x.dim()

    def materialize(self):
        return self.x + self.alpha * self.dx


class _FusedThresholded(torch.autograd.Function):
    """ThresholdedLoss on a CUDA dose volume through drtvam_thresholded_loss
    (csrc/loss_kernels.cu): loss and d loss / d dose in one pass."""

    @staticmethod
    def forward(ctx, x, dx, alpha, target, loss):
        from . import _lib
        lib = _lib.load()
        # surface-aware: two channels per voxel, the loss is per VOXEL (Loss.__call__, loss.py:44-53)
        two = target.dim() == 4 and target.shape[-1] == 2 and x.shape == target.shape
        n = x.numel() // 2 if two else x.numel()
        n_mean = loss.dose_count if (loss.dose_allsum is not None and loss.dose_count) else n
        need_grad = x.requires_grad and torch.is_grad_enabled()
        ctx.has_grad = bool(x.requires_grad)
        grad = torch.empty_like(x) if x.requires_grad else None
        out = torch.empty(1, device=x.device, dtype=torch.float64)
        mean = loss.reduction is torch.mean
        fn = lib.drtvam_thresholded_loss_sa if two else lib.drtvam_thresholded_loss
        _lib.check(fn(
            x.device.index or 0, x.data_ptr(), dx.data_ptr() if dx is not None else None, float(alpha),
            target.data_ptr(), n, float(loss.tl), float(loss.tu), float(loss.weight_object), float(loss.weight_void),
            float(loss.weight_limit), int(loss.K), (1.0 / n_mean) if mean else 1.0,
            grad.data_ptr() if grad is not None else None, out.data_ptr(),
            torch.cuda.current_stream(x.device).cuda_stream))
        ctx.grad = grad
        del need_grad
        total = out[0]
        if loss.dose_allsum is not None:      # z-slab shard: the dose term is a sum over ranks
            total = loss.dose_allsum(total)
        res = total / n_mean if mean else total
        return res.to(torch.float32)

    @staticmethod
    def backward(ctx, grad_out):
        g = ctx.grad * grad_out if ctx.grad is not None else None
        return g, None, None, None, None


class Loss:
    def __init__(self, props):
        reduction = props.get('reduction', 'sum')
        if reduction == 'sum':
            self.reduction = torch.sum
        elif reduction == 'mean':
            self.reduction = torch.mean
        else:
            raise ValueError(f"Invalid reduction method: '{reduction}'.")
        # sum of the (angle-sharded) pattern term over ranks; identity on one GPU
        self.pattern_allsum = None
        # z-slab sharding: every rank holds a slab of the dose, so the dose term is summed over ranks too
        self.dose_allsum = None
        self.dose_count = None   # total number of dose elements over all ranks (reduction 'mean')

    def eval_in(self, x):
        raise NotImplementedError

    def eval_out(self, x):
        raise NotImplementedError

    def eval_sparsity(self, patterns):
        raise NotImplementedError

    def eval(self, x, target, patterns):
        raise NotImplementedError

    def _reduce_patterns(self, lp):
        if not isinstance(lp, torch.Tensor):
            return lp
        if self.pattern_allsum is None:
            return self.reduction(lp)
        if self.reduction is torch.mean:
            n = self.pattern_allsum(torch.tensor(float(lp.numel()), device=lp.device))
            return self.pattern_allsum(lp.sum()) / n
        return self.pattern_allsum(lp.sum())

    def _reduce_dose(self, l):
        """Reduction of the per-voxel dose term; with a z-slab shard the value is the sum over all ranks
        while the gradient still flows to the local voxels (d total / d local = d local sum / d local)."""
        if self.dose_allsum is None:
            return self.reduction(l)
        local = l.sum()
        total = local + (self.dose_allsum(local.detach()) - local.detach())
        if self.reduction is torch.mean:
            n = self.dose_count if self.dose_count else float(
                self.dose_allsum(torch.tensor(float(l.numel()), device=l.device)))
            return total / n
        return total

    accepts_lazy = True

    def __call__(self, x, target, patterns):
        if isinstance(x, LazyAxpy):
            x = x.materialize()
        if x.shape != target.shape:
            if x.dim() == target.dim() + 1 and x.shape[-1] == 1:
                target = target[..., None]
            else:
                raise ValueError(f"Input and target shapes do not match: {tuple(x.shape)} != {tuple(target.shape)}")

        if target.shape[-1] == 1:  # binary or grayscale target
            loss, loss_patterns = self.eval(x, target, patterns)
        elif target.shape[-1] == 2:  # surface-aware: fractional in/out sub-volumes
            tot = target[..., 0] + target[..., 1]
            w_in = target[..., 0] / tot
            w_out = target[..., 1] / tot
            loss = w_in * self.eval_in(x[..., 0]) + w_out * self.eval_out(x[..., 1])
            loss_patterns = self.eval_sparsity(patterns)
        else:
            raise ValueError(f"[Loss] Received tensors of invalid shape: {tuple(target.shape)}. "
                             "The last dimension should be either 1 or 2.")
        # the two terms have different shapes, hence separate reductions
        return self._reduce_dose(loss) + self._reduce_patterns(loss_patterns)


class L2Loss(Loss):
    def __init__(self, props):
        super().__init__(props)
        self.M = props.get('M', 4)
        self.weight_sparsity = props.get('weight_sparsity', 0)

    def eval_in(self, x):
        return torch.square(x - 1.)

    def eval_out(self, x):
        return torch.square(x)

    def eval(self, x, target, patterns):
        return torch.square(x - target), 0 * patterns

    def eval_sparsity(self, patterns):
        return patterns ** self.M * self.weight_sparsity


class ThresholdedLoss(Loss):
    """Thresholded loss following Wechsler et al. 2024:
    w_obj relu(tu - x)^K + w_lim relu(x - 1)^K inside the object,
    w_void relu(x - tl)^K outside, |p|^M w_sparsity on the patterns."""

    def __init__(self, props):
        super().__init__(props)
        self.K = props.get('K', 2)
        self.M = props.get('M', 4)
        self.tl = props.get('tl', 0.9)
        self.tu = props.get('tu', 0.95)
        self.weight_object = props.get('weight_object', 1)
        self.weight_void = props.get('weight_void', 1)
        self.weight_limit = props.get('weight_limit', 1)
        self.weight_sparsity = props.get('weight_sparsity', 0)
        if self.tl >= self.tu:
            raise ValueError(f"[ThresholdedLoss] Lower threshold ({self.tl}) must be smaller than upper threshold ({self.tu})")

    def eval_in(self, x):
        return self.weight_object * relu(self.tu - x) ** self.K + self.weight_limit * relu(x - 1.) ** self.K

    def eval_out(self, x):
        return self.weight_void * relu(x - self.tl) ** self.K

    def eval_sparsity(self, patterns):
        return torch.abs(patterns) ** self.M * self.weight_sparsity

    def eval(self, x, target, patterns):
        # target is a binary inside/outside mask
        return torch.where(target > 0, self.eval_in(x), self.eval_out(x)), self.eval_sparsity(patterns)

    accepts_lazy = True

    def _fusable(self, x, target):
        if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()):
            return False
        if not (isinstance(target, torch.Tensor) and target.is_cuda and target.dtype == torch.float32
                and target.is_contiguous() and target.numel() == x.numel()):
            return False
        if x.shape != target.shape and not (x.dim() == target.dim() + 1 and x.shape[-1] == 1):
            return False
        return float(self.K).is_integer() and 1 <= int(self.K) <= 8

    @staticmethod
    def _two_channel(x, target):
        return target.shape == x.shape and x.dim() == 4 and x.shape[-1] == 2

    def armijo_sweep(self, x, dx, target, direction, first, count, sums, skip=None):
        """One stage of the Armijo back-tracking search (lbfgs.py:256-266), trials [first, first + count), in
        one sweep of the two volumes: sums[k] (device float64[32]) += sum over the local voxels of the loss
        at x + 2^-k dx, summed over ranks when the dose is z-slab sharded.  Returns (scale, pattern_term) -
        the reference evaluates the pattern term on the DIRECTION (lbfgs.py:260) - or None when this
        volume pair cannot take the fused path.  `skip`: device float, non-zero = search already over."""
        if not (self._fusable(x, target) and self._fusable(dx, target)) or int(self.K) > 4 or self._two_channel(x, target):
            return None   # (surface-aware volumes: the trials go one by one through drtvam_thresholded_loss_sa)
        from . import _lib
        lib = _lib.load()
        _lib.check(lib.drtvam_armijo_losses(
            x.device.index or 0, x.data_ptr(), dx.data_ptr(), target.data_ptr(), x.numel(), float(self.tl), float(self.tu),
            float(self.weight_object), float(self.weight_void), float(self.weight_limit), int(self.K), int(first),
            int(count), sums.data_ptr(), skip.data_ptr() if skip is not None else None,
            torch.cuda.current_stream(x.device).cuda_stream))
        n = x.numel()
        if self.dose_allsum is not None:
            sums[first:first + count] = self.dose_allsum(sums[first:first + count])
            n = self.dose_count if self.dose_count else n
        scale = 1.0 / n if self.reduction is torch.mean else 1.0
        term = None
        if self.weight_sparsity != 0:
            term = self._reduce_patterns(self.eval_sparsity(direction)).to(torch.float64).reshape(1)
        return scale, term

    def __call__(self, x, target, patterns):
        lazy = isinstance(x, LazyAxpy)
        base = x.x if lazy else x
        if not self._fusable(base, target) or (lazy and not self._fusable(x.dx, target)):
            return super().__call__(x.materialize() if lazy else x, target, patterns)
        loss = _FusedThresholded.apply(base, x.dx if lazy else None, x.alpha if lazy else 0.0, target, self)
        if self.weight_sparsity != 0:
            loss = loss + self._reduce_patterns(self.eval_sparsity(patterns))
        return loss


losses = {
    'l2': L2Loss,
    'threshold': ThresholdedLoss,
}
