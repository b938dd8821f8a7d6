"""Volume integrator: the forward projection ``A p`` and its adjoint ``A^T g``.
Mirrors ``src/drtvam/integrators/volume.py``: ``render`` (:18-56),
``render_forward`` (:58-95), ``render_backward`` (:97-134).  The bounce loop
``sample`` (:136-282) is ``csrc/tracer.cuh`` / ``csrc/planar_kernels.cu``.
"""
from __future__ import annotations

import torch

from .. import _lib
from .common import TVAMIntegrator


def _inv_volume(sensor, scene):
    """(scalar, tensor-or-None): 1 / voxel volume, 0 where the volume is 0 (volume.py:41-42)."""
    vol = sensor.compute_volume(scene)
    if isinstance(vol, torch.Tensor):
        cached = getattr(sensor, '_inv_volumes', None)
        if cached is None or cached[0] is not vol:
            cached = (vol, torch.where(vol != 0, 1.0 / vol, torch.zeros_like(vol)).contiguous())
            sensor._inv_volumes = cached
        return 1.0, cached[1]
    return (1.0 / vol if vol != 0 else 0.0), None


def _sensor_of(scene, sensor):
    if isinstance(sensor, int):
        return scene.sensors()[sensor]
    return sensor


def _check_emitters(scene):
    if len(scene.emitters()) == 0:
        raise Exception("No projector found in the scene")
    if len(scene.emitters()) > 1:
        raise Exception("The scene contains more than one projector. Only one is supported")


class VolumeIntegrator(TVAMIntegrator):
    def to_string(self):
        return ('VolumeIntegrator[\n'
                f'    {self.print_time=},\n'
                f'    {self.transmission_only=},\n'
                f'    {self.sample_time=},\n'
                f'    {self.max_depth=},\n'
                f'    {self.rr_depth=},\n'
                ']')

    __repr__ = to_string

    # CUDA-event timing of the C-ABI calls (the analogue of the reference's
    # KernelHistory timing, optimize.py:291,306,323): set `integrator.events = {}`
    # to collect {'forward' | 'adjoint' | 'allreduce': [(start, end), ...]}.
    events = None

    def _event_pair(self, name):
        if self.events is None:
            return None
        pair = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        pair[0].record()
        self.events.setdefault(name, []).append(pair)
        return pair

    def _project(self, scene, sensor, seed, spp, active_data):
        sensor = _sensor_of(scene, sensor)
        projector = scene.emitters()[0]
        spp, _ = self.prepare(projector, seed, spp)
        handle = scene.handle(self, sensor)
        inv_scalar, inv_tensor = _inv_volume(sensor, scene)
        film = sensor.film()
        film.data = torch.empty(scene.film_shape(sensor), device=projector.active_data.device, dtype=torch.float32)
        data = active_data.detach().to(torch.float32).contiguous()
        pixels = projector.active_pixels
        scene.pin_pixels(handle, pixels)
        if data.numel() != pixels.numel():
            raise ValueError(f"[{projector.__class__.__name__}] active_data and active_pixels must have the same length.")
        lib = _lib.load()
        ev = self._event_pair('forward')
        # per-voxel 1 / sub-voxel volume of a surface-aware film: applied by the library (volume.py:54)
        _lib.check(lib.drtvam_scene_set_inv_volume(handle, inv_tensor.data_ptr() if inv_tensor is not None else None))
        try:
            _lib.check(lib.drtvam_forward(handle, data.data_ptr(), pixels.data_ptr(), data.numel(), spp,
                                          seed & 0xFFFFFFFF, inv_scalar, film.data.data_ptr(), 1,
                                          torch.cuda.current_stream(film.data.device).cuda_stream))
        finally:
            lib.drtvam_scene_set_inv_volume(handle, None)
        if ev is not None:
            ev[1].record()
        ev = self._event_pair('allreduce') if scene.world_size > 1 else None
        scene.reduce_dose(film.data)
        if ev is not None:
            ev[1].record()
        return film.data

    def render(self, scene, sensor=0, seed: int = 0, spp: int = 0, develop: bool = True, evaluate: bool = True,
               host_patterns=None, blocks: int = 4):
        """Primal A p with p = projector.active_data; returns (resz, resy, resx, C).

        `host_patterns` (an addition to the reference signature): the pattern values in PINNED host memory.
        They become projector.active_data, uploaded in `blocks` blocks of whole projection angles on a copy
        stream while the projection of the previous block runs (A = sum_k A_k, so the blocks accumulate into
        the film): the 256 MB upload of C2 hides behind the kernels instead of preceding them."""
        _check_emitters(scene)
        with torch.no_grad():
            if host_patterns is not None:
                return self._project_streamed(scene, sensor, seed, spp, host_patterns, blocks)
            return self._project(scene, sensor, seed, spp, scene.emitters()[0].active_data)

    def _project_streamed(self, scene, sensor, seed, spp, host, blocks):
        sensor = _sensor_of(scene, sensor)
        projector = scene.emitters()[0]
        pixels = projector.active_pixels
        dev = pixels.device
        n = pixels.numel()
        hw = projector.res[0] * projector.res[1]
        rows = getattr(scene, '_z_rows', None)
        per_pattern = hw if rows is None else (rows[1] - rows[0]) * projector.res[0]
        host = host.reshape(-1)
        if host.numel() != n:
            raise ValueError(f"[{projector.__class__.__name__}] active_data and active_pixels must have the same length.")
        inv_scalar, inv_tensor = _inv_volume(sensor, scene)
        nk = n // per_pattern if per_pattern and n % per_pattern == 0 else 0
        import os
        blocks = max(1, min(int(os.environ.get("DRTVAM_STREAM_BLOCKS", blocks)), nk))
        if inv_tensor is not None or nk == 0 or blocks == 1 or not host.is_pinned():
            # per-voxel volumes / ragged lists / pageable memory: one upload, one projection
            projector.active_data = host.to(dev, torch.float32, non_blocking=True)
            return self._project(scene, sensor, seed, spp, projector.active_data)
        spp, _ = self.prepare(projector, seed, spp)
        handle = scene.handle(self, sensor)
        film = sensor.film()
        film.data = torch.empty(scene.film_shape(sensor), device=dev, dtype=torch.float32)
        data = torch.empty(n, device=dev, dtype=torch.float32)
        scene.pin_pixels(handle, pixels)
        main = torch.cuda.current_stream(dev)
        copy = scene.copy_stream()
        copy.wait_stream(main)
        bounds = [per_pattern * (nk * b // blocks) for b in range(blocks + 1)]
        ready = []
        with torch.cuda.stream(copy):
            for b in range(blocks):
                lo, hi = bounds[b], bounds[b + 1]
                data[lo:hi].copy_(host[lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy)
                ready.append(ev)
        data.record_stream(copy)
        lib = _lib.load()
        ev = self._event_pair('forward')
        _lib.check(lib.drtvam_scene_set_inv_volume(handle, None))
        # position of every entry in the wavefront of the whole call: only integrators that draw samples need it
        draws = not (self.regular_sampling and not self.sample_time and projector.kind == 'collimated'
                     and float(scene.medium.get('albedo', 0.0)) == 0.0 and sensor.kind == 'dda')
        base = scene.sampler_base(n) if draws else None
        try:
            for b in range(blocks):
                lo, hi = bounds[b], bounds[b + 1]
                main.wait_event(ready[b])
                if base is not None:
                    _lib.check(lib.drtvam_scene_set_sampler_index(handle, base[lo:hi].data_ptr()))
                _lib.check(lib.drtvam_forward(handle, data[lo:hi].data_ptr(), pixels[lo:hi].data_ptr(), hi - lo, spp,
                                              seed & 0xFFFFFFFF, inv_scalar, film.data.data_ptr(), int(b == 0),
                                              main.cuda_stream))
        finally:
            if base is not None:
                _lib.check(lib.drtvam_scene_set_sampler_index(
                    handle, scene._sampler_index.data_ptr() if scene._sampler_index is not None else None))
        if ev is not None:
            ev[1].record()
        projector.active_data = data
        ev = self._event_pair('allreduce') if scene.world_size > 1 else None
        scene.reduce_dose(film.data)
        if ev is not None:
            ev[1].record()
        return film.data

    def render_forward(self, scene, params, sensor=0, seed: int = 0, spp: int = 0):
        """Forward-mode derivative (volume.py:58-95), same signature as the reference.  There the
        tangent is the gradient attached to the pattern values (`dr.set_grad(params[key], tangent)`,
        picked up by `dr.forward_to(Le)`, volume.py:159-160); here it is the `.grad` of
        `params['projector.active_data']` (or of the projector's own tensor when `params` is None).  A is
        linear in the patterns, so the derivative is A applied to the tangent; without an attached
        tangent it is zero."""
        _check_emitters(scene)
        data = params['projector.active_data'] if params is not None else scene.emitters()[0].active_data
        tangent = getattr(data, 'grad', None)
        with torch.no_grad():
            if tangent is None:
                return torch.zeros(scene.film_shape(_sensor_of(scene, sensor)),
                                   device=scene.emitters()[0].active_data.device, dtype=torch.float32)
            return self._project(scene, sensor, seed, spp, tangent.reshape(-1))

    def render_backward(self, scene, params, grad_in, sensor=0, seed: int = 0, spp: int = 0, host_out=None, blocks: int = 1,
                        wait: bool = True):
        """Adjoint A^T grad_in; returns d loss / d projector.active_data.

        `host_out` (a pinned host tensor of the gradient's length; no counterpart in the reference, whose
        gradients stay Dr.Jit arrays): the gradient is also copied there on the scene's copy stream - with
        `blocks` > 1 in blocks of angles, the copy of a block running under the adjoint of the next one (measured
        at C2: the four smaller adjoint calls cost more than the overlap saves, 21.5 against 20.0 ms per step
        end to end, so one block is the default and the copy overlaps whatever the caller launches next).  The caller's stream
        waits for the copies before it goes on, unless `wait` is false (the caller then waits for
        `scene.copy_stream()` itself before it reads `host_out`)."""
        _check_emitters(scene)
        sensor = _sensor_of(scene, sensor)
        projector = scene.emitters()[0]
        spp, _ = self.prepare(projector, seed, spp)
        handle = scene.handle(self, sensor)
        inv_scalar, inv_tensor = _inv_volume(sensor, scene)
        g = grad_in.detach().to(torch.float32)
        if tuple(g.shape) != tuple(scene.film_shape(sensor)):
            g = g.reshape(scene.film_shape(sensor))
        g = g.contiguous()
        pixels = projector.active_pixels
        scene.pin_pixels(handle, pixels)
        out = torch.empty(pixels.numel(), device=g.device, dtype=torch.float32)
        lib = _lib.load()
        ev = self._event_pair('adjoint')
        _lib.check(lib.drtvam_scene_set_inv_volume(handle, inv_tensor.data_ptr() if inv_tensor is not None else None))
        main = torch.cuda.current_stream(g.device)
        n = pixels.numel()
        hw = projector.res[0] * projector.res[1]
        rows = getattr(scene, '_z_rows', None)
        per_pattern = hw if rows is None else (rows[1] - rows[0]) * projector.res[0]
        nk = n // per_pattern if per_pattern and n % per_pattern == 0 else 0
        # lanes that draw samples are numbered through the whole call: those calls are not cut into blocks
        draws = not (self.regular_sampling and not self.sample_time and projector.kind == 'collimated'
                     and float(scene.medium.get('albedo', 0.0)) == 0.0 and sensor.kind == 'dda')
        stream_out = (host_out is not None and host_out.is_pinned() and host_out.numel() == n and nk > 1
                      and blocks > 1 and not draws)
        blocks = max(1, min(int(blocks), nk)) if stream_out else 1
        bounds = [per_pattern * (nk * b // blocks) for b in range(blocks + 1)] if stream_out else [0, n]
        copy = scene.copy_stream() if host_out is not None else None
        try:
            for b in range(blocks):
                lo, hi = bounds[b], bounds[b + 1]
                _lib.check(lib.drtvam_adjoint(handle, g.data_ptr(), pixels[lo:hi].data_ptr(), hi - lo, spp,
                                              seed & 0xFFFFFFFF, inv_scalar, out[lo:hi].data_ptr(), main.cuda_stream))
                if host_out is not None:
                    done = torch.cuda.Event()
                    done.record(main)
                    with torch.cuda.stream(copy):
                        copy.wait_event(done)
                        host_out.reshape(-1)[lo:hi].copy_(out[lo:hi], non_blocking=True)
        finally:
            lib.drtvam_scene_set_inv_volume(handle, None)
        if host_out is not None:
            out.record_stream(copy)
            if wait:
                main.wait_stream(copy)
        if ev is not None:
            ev[1].record()
        return out
