"""Corner pre-filter.  Mirrors ``src/drtvam/integrators/filter_corner.py``: one intersection per
lane; a pixel whose first hit lies within ``radius`` of the vertical edge ``|x| = |y| = dist`` of a
square vial renders 0, every other pixel its ray weight (:51-66); ``optimize.py:166-185`` keeps the
positive pixels.  Kernel: ``corner_kernel`` in ``csrc/filter_kernels.cu``.
"""
from __future__ import annotations

from .. import _lib
from .radon import RadonIntegrator


class CornerIntegrator(RadonIntegrator):
    def __init__(self, props):
        super().__init__(props)
        self.dist = props['dist']
        self.radius = props.get('radius', 0.1)

    def to_string(self):
        return 'CornerIntegrator'

    __repr__ = to_string

    def _call(self, lib, handle, pixels, spp, seed, image, stream):
        _lib.check(lib.drtvam_corner_filter(handle, pixels.data_ptr(), pixels.numel(), spp, seed & 0xFFFFFFFF,
                                            float(self.dist), float(self.radius), image.data_ptr(), stream))
