"""Projector emitters.  Mirrors ``src/drtvam/projector.py``: ``TVAMProjector``
(:41-165) with its sparse ``(active_data, active_pixels)`` pattern store, crop
window and ``filter_nonzero``; ``CollimatedProjector`` (:167-196),
``TelecentricProjector`` (:199-237), ``LensProjector`` (:241-296);
``load_patterns`` (:8-39).

The ray arithmetic (``get_ray`` / ``sample_ray``) runs inside the CUDA kernels
(``csrc/raygen.cuh``); these classes hold the parameters and the device
buffers (torch tensors) the kernels read.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch

from .motion import Motion, motions


def load_patterns(filepath):
    if os.path.isfile(filepath):
        if filepath.endswith(".npy"):
            patterns = np.load(filepath)
        elif filepath.endswith(".npz"):
            patterns = np.load(filepath)
            if len(patterns.files) != 1:
                raise ValueError(f"Expected a single array in the npz file, but got {len(patterns.files)} arrays.")
            patterns = patterns[patterns.files[0]]
        else:
            raise ValueError(f"Unsupported file format for patterns: {os.path.splitext(filepath)[1]}")
        if len(patterns.shape) != 3:
            raise ValueError(f"Patterns must be 3D, but got a tensor of shape {patterns.shape}.")
        return torch.from_numpy(np.ascontiguousarray(patterns, dtype=np.float32))
    raise ValueError("No patterns found: EXR directories need an OpenEXR reader, which this build does not "
                     "ship; export the patterns as .npy/.npz.")


def _device(props):
    dev = props.get('device', None)
    if dev is None:
        dev = torch.device('cuda', torch.cuda.current_device()) if torch.cuda.is_available() else torch.device('cpu')
    return torch.device(dev)


class TVAMProjector:
    kind = None

    def __init__(self, props):
        self.device = _device(props)
        if 'patterns' in props:
            patterns = props['patterns']
            if isinstance(patterns, str):
                patterns = load_patterns(patterns)
            elif isinstance(patterns, np.ndarray):
                patterns = torch.from_numpy(np.ascontiguousarray(patterns, dtype=np.float32))
            elif not isinstance(patterns, torch.Tensor):
                raise ValueError(f"[{self.__class__.__name__}] patterns must be a tensor, an ndarray or a path")
            if patterns.dim() != 3:
                raise ValueError(f"[{self.__class__.__name__}] Patterns must be 3D, but got a tensor of shape {tuple(patterns.shape)}.")
            n, h, w = patterns.shape
            self.n_patterns = n
            self.res = (w, h)
            self.crop = self.res
            self.crop_offset = (0, 0)
            flat = patterns.to(self.device, torch.float32).reshape(-1)
            if props.get('filter_nonzero', False):
                self.active_pixels = torch.nonzero(flat > 0).reshape(-1).to(torch.int32)
                self.active_data = flat[self.active_pixels.long()].contiguous()
            else:
                self.active_data = flat.contiguous().clone()
                self.active_pixels = torch.arange(n * h * w, device=self.device, dtype=torch.int64).to(torch.int32)
        else:
            self.n_patterns = props.get('n_patterns', 1000)
            resx = props.get('resx', 256)
            resy = props.get('resy', 256)
            self.res = (resx, resy)
            cropx = props.get('cropx', resx)
            cropy = props.get('cropy', resy)
            self.crop = (cropx, cropy)
            if cropx > resx or cropy > resy:
                raise ValueError(f"[{self.__class__.__name__}] Crop resolution ({self.crop}) must be smaller than the base resolution ({self.res}).")
            ox = props.get('crop_offset_x', 0)
            oy = props.get('crop_offset_y', 0)
            self.crop_offset = (ox, oy)
            if ox + cropx > resx or oy + cropy > resy:
                raise ValueError(f"[{self.__class__.__name__}] With the specified crop offset ({self.crop_offset}), the cropped region ({self.crop}) extends beyond the base resolution ({self.res}).")
            n_crop = cropx * cropy
            self.active_data = torch.zeros(self.n_patterns * n_crop, device=self.device, dtype=torch.float32)
            # flat pixel index k*h*w + row*w + col of every pixel of the crop window
            rows = torch.arange(cropy, device=self.device, dtype=torch.int64) + oy
            cols = torch.arange(cropx, device=self.device, dtype=torch.int64) + ox
            window = (rows[:, None] * resx + cols[None, :]).reshape(-1)
            k = torch.arange(self.n_patterns, device=self.device, dtype=torch.int64) * (resx * resy)
            self.active_pixels = (k[:, None] + window[None, :]).reshape(-1).to(torch.int32)

        if "motion" not in props:
            raise ValueError(f"[{self.__class__.__name__}] Missing field 'motion'.")
        if isinstance(props['motion'], Motion):
            self.motion = props['motion']
        elif isinstance(props['motion'], str):
            if props['motion'] not in motions.keys():
                raise ValueError(f"[{self.__class__.__name__}] Invalid motion type: {props['motion']}")
            self.motion = motions[props['motion']](props)
        else:
            raise ValueError(f"[{self.__class__.__name__}] motion must be either a dict or a Motion instance")

        self.pixel_size = (1.0, 1.0)
        self.aperture_radius = 0.0
        self.focus_distance = 1.0

    # -- reference API -----------------------------------------------------
    def active_size(self):
        return int(self.active_data.numel())

    def size(self):
        return (self.n_patterns, self.res[1], self.res[0])

    def patterns(self):
        """Return the full (dense) patterns."""
        n, h, w = self.size()
        out = torch.zeros(n * h * w, device=self.active_data.device, dtype=torch.float32)
        out[self.active_pixels.long() & 0xFFFFFFFF] = self.active_data.detach()
        return out.reshape(n, h, w)

    def parameters_changed(self, keys=None):
        if self.active_data.numel() != self.active_pixels.numel():
            raise ValueError(f"[{self.__class__.__name__}] active_data and active_pixels must have the same length.")

    def motion_table(self):
        """to_world of every pattern at time k / n_patterns: [n, 4, 4]."""
        n = self.n_patterns
        tab = np.empty((n, 4, 4), np.float64)
        for k in range(n):
            t = np.float32(k) / np.float32(n)
            tab[k] = self.motion.eval(t).matrix
        return tab

    def to_string(self):
        return (f'{self.__class__.__name__}[\n'
                f'    pattern count = {self.n_patterns},\n'
                f'    pattern resolution = {self.res},\n'
                ']')

    __repr__ = to_string


def _pixel_size(value, name):
    if isinstance(value, (int, float)):
        return (float(value), float(value))
    if isinstance(value, (tuple, list)) and len(value) == 2:
        return (float(value[0]), float(value[1]))
    raise ValueError(f"[{name}] pixel_size must be a float or a pair of floats")


class CollimatedProjector(TVAMProjector):
    kind = 'collimated'

    def __init__(self, props):
        super().__init__(props)
        self.pixel_size = _pixel_size(props['pixel_size'], self.__class__.__name__)
        self.emitter_size = (self.res[0] * self.pixel_size[0], self.res[1] * self.pixel_size[1])


class TelecentricProjector(TVAMProjector):
    kind = 'telecentric'

    def __init__(self, props):
        super().__init__(props)
        self.pixel_size = _pixel_size(props['pixel_size'], self.__class__.__name__)
        self.aperture_radius = props['aperture_radius']
        self.focus_distance = props['focus_distance']
        self.emitter_size = (self.res[0] * self.pixel_size[0], self.res[1] * self.pixel_size[1])


class LensProjector(TVAMProjector):
    kind = 'lens'

    def __init__(self, props):
        super().__init__(props)
        self.aperture_radius = props['aperture_radius']
        self.focus_distance = props['focus_distance']
        assert not ('fov' in props and 'pixel_size' in props), "Specify either 'fov' or 'pixel_size', not both."
        assert ('fov' in props or 'pixel_size' in props), "Either 'fov' or 'pixel_size' must be specified."
        # pixel_size * N_x = tan(fov / 2) * 2 * focus_distance
        if 'fov' in props:
            self.fov = props['fov']
            ps = math.tan(math.radians(self.fov) / 2) * 2 * self.focus_distance / self.res[0]
        else:
            ps = props['pixel_size']
            self.fov = math.degrees(2 * math.atan(ps * self.res[0] / 2 / self.focus_distance))
        self.pixel_size = (float(ps), float(ps))


projectors = {
    'collimated': CollimatedProjector,
    'lens': LensProjector,
    'telecentric': TelecentricProjector,
}
