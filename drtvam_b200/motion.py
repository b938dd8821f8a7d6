"""Projector motion models.  Mirrors ``src/drtvam/motion.py`` of the reference
(``Motion`` base :4-17, ``CircularMotion`` :19-36, registry :39-41).

``eval(time)`` returns the projector's ``to_world`` transform at normalised
time ``time`` in [0, 1].  The kernels consume a per-pattern table of these
(``time = k / n_patterns``), or evaluate the circular motion per lane when
``sample_time`` is on.
"""
from __future__ import annotations

import numpy as np

from .transform import Transform4f


class Motion:
    """For a given normalized time t in [0, 1], returns the projector transform."""

    def __init__(self, props):
        raise NotImplementedError

    def eval(self, time):
        raise NotImplementedError


class CircularMotion(Motion):
    def __init__(self, props):
        self.distance = props['distance']
        self.tilt = props.get('tilt', 0.)  # parsed and unused, as in motion.py:22
        self.clockwise = props.get('clockwise', False)

    def eval(self, time):
        f32 = np.float32
        alpha = f32(2 * np.pi) * f32(time)
        if self.clockwise:
            alpha = -alpha
        origin = f32(self.distance) * np.array([np.cos(alpha, dtype=f32), np.sin(alpha, dtype=f32), 0.], dtype=f32)
        return Transform4f().look_at(origin=origin, target=[0, 0, 0], up=[0, 0, 1])


# List of registered motions
motions = {
    'circular': CircularMotion,
}
