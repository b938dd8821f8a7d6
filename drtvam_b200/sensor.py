"""Dose sensors.  Mirrors ``src/drtvam/sensor.py``: ``VolumetricSensor``
(:5-110: bounding box from ``to_world``, ``voxel_size``, ``compute_volume``),
``DDAVolumetricSensor`` (:297-440), ``RatioVolumetricSensor`` (:193-295),
``DeltaVolumetricSensor`` (:112-191).

``accumulate`` -- the voxel walk itself -- is the CUDA hot loop
(``csrc/dda.cuh``, ``csrc/planar_kernels.cu``); these classes carry the
geometry.  The surface-aware
sub-voxel volumes (:53-110) are ``drtvam_compute_volume`` (``csrc/mesh_kernels.cu``).
"""
from __future__ import annotations

import numpy as np

from .film import VolumetricFilm
from .transform import Transform4f


class VolumetricSensor:
    kind = None

    def __init__(self, props):
        film = props.get('film')
        if isinstance(film, dict):
            if film.get('type', 'vfilm') != 'vfilm':
                raise ValueError("Tried to load a VolumetricSensor with a non-volumetric film. The film must be of type VolumetricFilm.")
            film = VolumetricFilm(film, device=props.get('device'))
        if not isinstance(film, VolumetricFilm):
            raise ValueError("Tried to load a VolumetricSensor with a non-volumetric film. The film must be of type VolumetricFilm.")
        self.m_film = film
        self.to_world = Transform4f(props.get('to_world', Transform4f()))
        f32 = np.float32
        lo = (self.to_world @ np.full(3, -0.5)).astype(f32)
        hi = (self.to_world @ np.full(3, 0.5)).astype(f32)
        self.bbox_min, self.bbox_max = np.minimum(lo, hi), np.maximum(lo, hi)
        # size of a voxel along each dimension (sensor.py:19)
        self.voxel_size = (self.bbox_max - self.bbox_min) / np.asarray(self.m_film.resolution(), f32)
        self.volumes = None
        self.id = props.get('id', 'sensor')

    def film(self):
        return self.m_film

    def resolution(self):
        return self.m_film.resolution()

    def compute_volume(self, scene=None, sample_count=2 ** 14):
        """Voxel volume (scalar), or the cached 2-channel in/out sub-volumes when
        the film is surface aware (sensor.py:47-110)."""
        if not self.m_film.surface_aware:
            return float(np.prod(self.voxel_size, dtype=np.float32))
        if self.volumes is not None:
            return self.volumes
        if scene is None or getattr(scene, 'target', None) is None:
            raise ValueError("No target shape found in the scene")
        import torch

        from . import _lib
        lib = _lib.load()
        handle = scene.handle(scene.geometry_integrator(), self)
        dev = scene.projector.active_data.device
        rx, ry, rz = self.resolution()
        vol = torch.empty((rz, ry, rx, 2), device=dev, dtype=torch.float32)
        _lib.check(lib.drtvam_compute_volume(handle, int(sample_count), vol.data_ptr(),
                                             torch.cuda.current_stream(dev).cuda_stream))
        self.volumes = vol
        return self.volumes

    def to_string(self):
        return (f'{self.__class__.__name__}[\n'
                f'    to_world = {self.to_world},\n'
                ']')

    __repr__ = to_string


class DDAVolumetricSensor(VolumetricSensor):
    kind = 'dda'


class RatioVolumetricSensor(VolumetricSensor):
    kind = 'ratio'

    def __init__(self, props):
        super().__init__(props)
        self.majorant = props['majorant']


class DeltaVolumetricSensor(VolumetricSensor):
    """One deposit per medium collision (sensor.py:112-191); scattering media only."""
    kind = 'delta'


sensors = {
    'dda': DDAVolumetricSensor,
    'ratio': RatioVolumetricSensor,
    'delta': DeltaVolumetricSensor,
}
