"""Angle sharding across GPUs (one process per GPU).

``A = sum_k A_k`` over projection angles ``k``: rank ``r`` owns a contiguous
range of angles, traces only those, and one ``all_reduce(SUM)`` of the dose
volume follows each forward projection.  The adjoint needs no collective: the
dose gradient is replicated and pattern gradients stay sharded.  L-BFGS dot
products over the sharded pattern vectors become local dot + scalar
all-reduce (``lbfgs.LinearLBFGS(dot=...)``).
"""
from __future__ import annotations

import torch


def angle_ranges(counts, world_size: int):
    """Split angles [0, n) into `world_size` contiguous ranges balanced by the
    number of active pixels per angle.  Returns a list of (k_lo, k_hi)."""
    n = len(counts)
    total = float(sum(counts))
    bounds = [0]
    acc = 0.0
    k = 0
    for r in range(1, world_size):
        goal = total * r / world_size
        while k < n and acc + counts[k] * 0.5 < goal:
            acc += counts[k]
            k += 1
        bounds.append(k)
    bounds.append(n)
    return [(bounds[i], bounds[i + 1]) for i in range(world_size)]


def shard_bounds(active_pixels: torch.Tensor, n_patterns: int, pixels_per_pattern: int, rank: int, world_size: int):
    """Index range [lo, hi) of `active_pixels` (sorted by pattern) owned by `rank`."""
    if world_size == 1:
        return 0, int(active_pixels.numel())
    k = (active_pixels.to(torch.int64) & 0xFFFFFFFF) // pixels_per_pattern
    if k.numel() > 1 and bool((k[1:] < k[:-1]).any()):
        raise ValueError("active_pixels must be sorted by pattern index to shard by angle")
    counts = torch.bincount(k, minlength=n_patterns).tolist()
    k_lo, k_hi = angle_ranges(counts, world_size)[rank]
    lo = int(sum(counts[:k_lo]))
    hi = int(sum(counts[:k_hi]))
    return lo, hi


def make_dot(group=None, world_size: int = 1):
    """Dot product over angle-sharded vectors: local dot + scalar all-reduce."""
    if world_size == 1:
        return lambda a, b: torch.dot(a.reshape(-1), b.reshape(-1))
    import torch.distributed as dist

    def dot(a, b):
        v = torch.dot(a.reshape(-1), b.reshape(-1))
        dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
        return v

    return dot


def make_sum(group=None, world_size: int = 1):
    """Sum of a scalar tensor over ranks (sparsity term of the loss)."""
    if world_size == 1:
        return lambda v: v
    import torch.distributed as dist

    def allsum(v):
        v = v.clone()
        dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
        return v

    return allsum


def make_consensus(group=None, world_size: int = 1):
    """Rank-mean of a tensor: identical bits on every rank afterwards.  Quantities that every rank computes
    redundantly from the same all-reduced dose (loss sums, the Armijo candidates) can differ in their last
    bits between ranks because their reductions use atomics; a back-tracking decision taken on them has to
    be the same everywhere, or the ranks would step to different patterns."""
    if world_size == 1:
        return lambda v: v
    import torch.distributed as dist

    def consensus(v):
        v = v.clone()
        dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
        return v / world_size

    return consensus


def slab_grid(world_size: int, n_layers: int):
    """(angle groups A, z-slabs S) with A * S = world_size for a planar scene.  The planar kernels amortise one
    2-D walk over the layers of a block (up to 128), so a slab should keep ~100 layers: cutting thinner
    repeats the walk for fewer layers, cutting by angle instead divides the walk but needs an all-reduce
    of the dose inside the angle group.  Measured at C2 on 8 B200 (profiles/r02q_*, r02r_*): 8 slabs of 50
    layers 4.9 ms per step, 8 angle groups 5.3 ms, 2 angle groups x 4 slabs of 100 layers 4.05 ms; on 2 GPUs
    2 slabs of 200 layers (three blocks: 128 + 64 + 8) 14.2 ms, 2 angle groups 11.9 ms."""
    best = (world_size, 1)
    for s in range(2, world_size + 1):
        if world_size % s:
            continue
        if 96 <= n_layers // s and -(-n_layers // s) <= 128:   # a slab is ONE block of layers, mostly full
            best = (world_size // s, s)
    return best


def zslab_shard(scene, slab: int, n_slabs: int, integrator, sensor, k_lo: int = 0, k_hi=None):
    """z-slab / DMD-row partition of a planar scene ("alternative worth
    measuring"): slab s owns the film layers [z0, z1) and, for every projection angle of [k_lo, k_hi), the
    DMD rows whose rays deposit there.  Returns {"window": (z0, z1), "rows": (r0, r1), "index": positions of the
    owned entries in the full active vector} or None when the scene does not allow it (not planar, rows
    that straddle layers, sparse pixel lists, CPU tensors)."""
    p = scene.projector
    if p.active_pixels.device.type != 'cuda' or sensor.film().surface_aware:
        return None
    from . import _lib
    import ctypes as C
    lib = _lib.load()
    n, h, w = p.size()
    if p.active_pixels.numel() != n * h * w:
        return None                                  # filtered / cropped pixel lists: shard by angle
    if any(not scene._same_grid(s_, sensor) for s_ in scene.sensors()):
        return None                                  # e.g. a final_sensor with another grid: shard by angle
    handle = scene.handle(integrator, sensor)
    rz = int(sensor.resolution()[2])
    # balance by layer count (every layer of a planar scene costs the same)
    z0 = rz * slab // n_slabs
    z1 = rz * (slab + 1) // n_slabs
    if z1 <= z0:
        return None
    r0, r1 = C.c_int32(0), C.c_int32(0)
    if lib.drtvam_scene_set_z_window(handle, z0, z1, C.byref(r0), C.byref(r1)) != 0:
        return None
    k_hi = n if k_hi is None else k_hi
    rows = torch.arange(r0.value, r1.value, device=p.active_pixels.device, dtype=torch.int64)
    cols = torch.arange(w, device=rows.device, dtype=torch.int64)
    k = torch.arange(k_lo, k_hi, device=rows.device, dtype=torch.int64)
    index = (k[:, None, None] * (h * w) + rows[None, :, None] * w + cols[None, None, :]).reshape(-1)
    return {"window": (z0, z1), "index": index, "rows": (r0.value, r1.value)}
