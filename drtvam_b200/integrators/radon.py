"""Radon pre-filter.  Mirrors ``src/drtvam/integrators/radon.py``: a forward-only integrator
whose image holds, per DMD pixel, the absorbed fraction along the parts of the path that lie
inside the target mesh; ``optimize.py:143-163`` keeps the pixels where it is positive.  The
bounce loop (``sample``, :48-106) is ``radon_kernel`` in ``csrc/filter_kernels.cu``.
"""
from __future__ import annotations

import torch

from .. import _lib
from .common import TVAMIntegrator
from .volume import _check_emitters, _sensor_of


class RadonIntegrator(TVAMIntegrator):
    def to_string(self):
        return ('RadonIntegrator[\n'
                f'    {self.max_depth=},\n'
                f'    {self.rr_depth=},\n'
                ']')

    __repr__ = to_string

    def _call(self, lib, handle, pixels, spp, seed, image, stream):
        _lib.check(lib.drtvam_radon(handle, pixels.data_ptr(), pixels.numel(), spp, seed & 0xFFFFFFFF,
                                    image.data_ptr(), stream))

    def render(self, scene, sensor=0, seed: int = 0, spp: int = 0, develop: bool = True, evaluate: bool = True):
        """Image of shape ``projector.size()`` = (n_patterns, res_y, res_x) (radon.py:13-46)."""
        _check_emitters(scene)
        projector = scene.emitters()[0]
        spp, _ = self.prepare(projector, seed, spp)
        if scene.target is None and isinstance(self, RadonIntegrator):
            raise ValueError("No target shape found in the scene")  # radon.py:57-58
        lib = _lib.load()
        image = torch.empty(projector.size(), device=projector.active_data.device, dtype=torch.float32)
        handle = scene.filter_handle(self, _sensor_of(scene, sensor))
        try:
            self._call(lib, handle, projector.active_pixels, spp, seed, image,
                       torch.cuda.current_stream(image.device).cuda_stream)
            torch.cuda.current_stream(image.device).synchronize()
        finally:
            lib.drtvam_scene_destroy(handle)
        return image

    def _project(self, scene, sensor, seed, spp, active_data):  # the entry drtvam_b200.render uses
        return self.render(scene, sensor, seed, spp)
