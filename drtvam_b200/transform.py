"""Minimal 4x4 affine transform with the call syntax the reference uses on
``mi.ScalarTransform4f`` / ``mi.Transform4f`` (``optimize.py:49-56``,
``geometry.py:200,208``, ``motion.py:36``).  Host-side, numpy, float32 like the
reference's ``mono`` variants.
"""
from __future__ import annotations

import numpy as np


class Transform4f:
    def __init__(self, matrix=None):
        if matrix is None:
            self.matrix = np.eye(4, dtype=np.float64)
        elif isinstance(matrix, Transform4f):
            self.matrix = matrix.matrix.copy()
        elif np.isscalar(matrix):
            self.matrix = np.eye(4, dtype=np.float64) * float(matrix)
            self.matrix[3, 3] = 1.0
        else:
            self.matrix = np.asarray(matrix, dtype=np.float64).reshape(4, 4).copy()

    # builders (usable as Transform4f().scale(..) like Mitsuba >= 3.6)
    def scale(self, v):
        v = np.broadcast_to(np.asarray(v, np.float64), (3,))
        m = np.eye(4)
        m[0, 0], m[1, 1], m[2, 2] = v
        return Transform4f(self.matrix @ m)

    def translate(self, v):
        v = np.asarray(v, np.float64).reshape(3)
        m = np.eye(4)
        m[:3, 3] = v
        return Transform4f(self.matrix @ m)

    def look_at(self, origin, target, up):
        """Mitsuba's look_at: columns [left, new_up, dir, origin] (motion.py:36)."""
        f32 = np.float32
        origin = np.asarray(origin, f32).reshape(3)
        target = np.asarray(target, f32).reshape(3)
        up = np.asarray(up, f32).reshape(3)
        d = target - origin
        d = d / np.sqrt(np.dot(d, d), dtype=f32)
        left = np.cross(up, d).astype(f32)
        left = left / np.sqrt(np.dot(left, left), dtype=f32)
        new_up = np.cross(d, left).astype(f32)
        m = np.eye(4)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, origin
        return Transform4f(self.matrix @ m)

    def inverse(self):
        return Transform4f(np.linalg.inv(self.matrix))

    def __matmul__(self, other):
        if isinstance(other, Transform4f):
            return Transform4f(self.matrix @ other.matrix)
        p = np.asarray(other, np.float64)
        if p.shape[-1] == 3:  # treated as a point
            q = p @ self.matrix[:3, :3].T + self.matrix[:3, 3]
            return q
        raise TypeError("Transform4f @ expects a Transform4f or a point")

    def transform_vector(self, v):
        return np.asarray(v, np.float64) @ self.matrix[:3, :3].T

    def __repr__(self):
        return f"Transform4f(\n{self.matrix}\n)"


ScalarTransform4f = Transform4f
