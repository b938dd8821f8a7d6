"""Volumetric film: the dose buffer.  Mirrors ``src/drtvam/film.py:4-43``.

``data`` is a torch tensor ``(resz, resy, resx, C)`` on the GPU, C = 2 when
``surface_aware`` (flat index ``C * (x + y*resx + z*resx*resy) + c``,
``sensor.py:405-409``).  The scatter-add of ``film.write`` (film.py:40-41)
happens inside the forward kernels.
"""
from __future__ import annotations

import torch


class VolumetricFilm:
    def __init__(self, props, device=None):
        resz = props.get('resz', 256)
        # the reference reads the two lateral keys crossed over (film.py:10-11); kept
        resy = props.get('resx', 256)
        resx = props.get('resy', 256)
        self.res = (resx, resy, resz)
        self.surface_aware = props.get('surface_aware', False)
        self.device = torch.device(device) if device is not None else (
            torch.device('cuda', torch.cuda.current_device()) if torch.cuda.is_available() else torch.device('cpu'))
        self.data = None

    @property
    def shape(self):
        resx, resy, resz = self.res
        return (resz, resy, resx, 2 if self.surface_aware else 1)

    def to_string(self):
        return ('VolumetricFilm[\n'
                f'    resolution = {self.shape},\n'
                ']')

    __repr__ = to_string

    def clear(self):
        self.data = torch.zeros(self.shape, device=self.device, dtype=torch.float32)

    def resolution(self):
        return self.res

    def develop(self):
        return self.data
