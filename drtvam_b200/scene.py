"""Scene container, parameter traversal and the differentiable ``render`` entry
point -- the analogue of ``mi.load_dict`` / ``mi.traverse`` / ``mi.render`` as
the reference uses them (``optimize.py:90-91,141,157-159,216,294,328``).

``render`` wraps the C-ABI calls in a ``torch.autograd.Function`` whose forward
is ``drtvam_forward`` and whose backward is ``drtvam_adjoint`` -- the
counterpart of Mitsuba's ``_RenderOp``.

Multi-GPU: when ``scene.shard(rank, world)`` has been called, the projector
holds only this rank's projection angles, every forward projection is followed
by one ``all_reduce(SUM)`` of the dose volume, and the adjoint stays local.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import numpy as np
import torch

from . import _lib
from ._scene import (PHASE_KINDS, PROJ_KINDS, SENSOR_KINDS, SceneDesc, SurfaceDesc, create_handle,
                     surface_from_shape_dict)
from .integrators import integrators as _integrators
from .motion import CircularMotion
from .projector import TVAMProjector, projectors
from .sensor import VolumetricSensor, sensors as _sensors


def sample_tea_32(v0: int, v1: int, rounds: int = 4):
    """Tiny Encryption Algorithm hash Mitsuba uses to derive seed_grad."""
    s = 0
    v0 &= 0xFFFFFFFF
    v1 &= 0xFFFFFFFF
    for _ in range(rounds):
        s = (s + 0x9e3779b9) & 0xFFFFFFFF
        v0 = (v0 + ((((v1 << 4) & 0xFFFFFFFF) + 0xa341316c) ^ ((v1 + s) & 0xFFFFFFFF) ^ ((v1 >> 5) + 0xc8013ea4))) & 0xFFFFFFFF
        v1 = (v1 + ((((v0 << 4) & 0xFFFFFFFF) + 0xad90777d) ^ ((v0 + s) & 0xFFFFFFFF) ^ ((v0 >> 5) + 0x7e95761e))) & 0xFFFFFFFF
    return v0, v1


class Scene:
    def __init__(self, projector: TVAMProjector, sensors: Dict[str, VolumetricSensor],
                 surfaces: List[SurfaceDesc], medium: dict, target: Optional[dict] = None,
                 integrator=None):
        self.projector = projector
        self._sensors = sensors
        self.surfaces = surfaces
        self.medium = medium
        self.target = target
        # The reference keeps the target mesh in the ray-traced scene only for surface-aware runs
        # and otherwise moves it 1e5 away (optimize.py:188-191), so its crossings only exist then.
        main = sensors.get('sensor') or next(iter(sensors.values()))
        self.target_active = bool(target is not None and main.film().surface_aware)
        self.default_integrator = integrator
        self._handles = {}
        self._keepalive = {}
        self._handle_sensor = {}
        self._pinned = {}
        self._sampler_index = None
        self._z_sensor = None
        self.group = None      # torch.distributed process group when angle-sharded
        self.world_size = 1
        self.rank = 0
        self.force_general = False

    # -- Mitsuba-like accessors --------------------------------------------
    def emitters(self):
        return [self.projector]

    def sensors(self):
        return list(self._sensors.values())

    def sensor(self, name):
        return self._sensors[name]

    def integrator(self):
        return self.default_integrator

    # -- lowering ----------------------------------------------------------
    def describe(self, integrator, sensor) -> SceneDesc:
        p = self.projector
        d = SceneDesc()
        d.proj_kind = PROJ_KINDS[p.kind]
        d.n_patterns, d.res_x, d.res_y = p.n_patterns, p.res[0], p.res[1]
        d.pixel_x, d.pixel_y = p.pixel_size
        d.aperture_radius, d.focus_distance = float(p.aperture_radius), float(p.focus_distance)
        circular = isinstance(p.motion, CircularMotion)
        if integrator.sample_time:
            if not circular:
                raise NotImplementedError("sample_time needs an analytic motion; only 'circular' has a device evaluator")
            d.motion_analytic = 1
        if circular:
            d.clockwise = int(bool(p.motion.clockwise))
            d.distance = float(p.motion.distance)
        d.motion_table = p.motion_table()
        d.max_depth, d.rr_depth = int(integrator.max_depth), int(integrator.rr_depth)
        d.transmission_only = int(bool(integrator.transmission_only))
        d.regular_sampling = int(bool(integrator.regular_sampling))
        d.sample_time = int(bool(integrator.sample_time))
        d.print_time = float(integrator.print_time)
        d.sigma_t = float(self.medium['sigma_t'])
        d.albedo = float(self.medium.get('albedo', 0.0))
        phase = self.medium.get('phase') or {'type': 'isotropic'}
        if phase['type'] not in PHASE_KINDS:
            raise NotImplementedError(f"phase function '{phase['type']}' has no kernel")
        d.phase_kind = PHASE_KINDS[phase['type']]
        d.hg_g = float(phase.get('g', 0.0))
        d.surfaces = list(self.surfaces)
        d.bbox_min = tuple(float(x) for x in sensor.bbox_min)
        d.bbox_max = tuple(float(x) for x in sensor.bbox_max)
        d.res = tuple(int(x) for x in sensor.resolution())
        d.surface_aware = int(bool(sensor.film().surface_aware))
        if sensor.kind not in SENSOR_KINDS:
            raise NotImplementedError(f"sensor '{sensor.kind}' has no kernel")
        d.sensor_kind = SENSOR_KINDS[sensor.kind]
        if sensor.kind == 'delta' and not d.albedo > 0:  # volume.py:156-157
            raise ValueError("Tried to render a purely absorptive volume with a delta tracking sensor. "
                             "This is not supported.")
        d.majorant = float(getattr(sensor, 'majorant', 0.0))
        d.target_triangles = self.target_triangles()
        return d

    def target_triangles(self):
        """[n_tri, 3, 3] float32 world-space vertices of the target while it is part of the traced
        scene (surface-aware runs), else None."""
        if not self.target_active:
            return None
        return self.all_target_triangles()

    def all_target_triangles(self):
        t = self.target
        if t is None:
            return None
        return np.ascontiguousarray(np.asarray(t['vertices'], np.float64)[np.asarray(t['triangles'])].astype(np.float32))

    def handle(self, integrator, sensor):
        """C handle for (integrator options, sensor); depth changes reuse it."""
        key = (id(sensor), bool(integrator.transmission_only), bool(integrator.regular_sampling),
               bool(integrator.sample_time), float(integrator.print_time))
        lib = _lib.load()
        h = self._handles.get(key)
        if h is None:
            desc = self.describe(integrator, sensor)
            dev = self.projector.active_data.device
            if dev.type != 'cuda':
                raise RuntimeError("drtvam_b200 has no CPU path: scene tensors must live on a CUDA device")
            h, keep = create_handle(desc, dev.index or 0, desc.target_triangles)
            self._handles[key] = h
            self._keepalive[key] = keep
            self._handle_sensor[h] = sensor
            self._apply_shard(h, sensor)
        _lib.check(lib.drtvam_scene_set_depth(h, int(integrator.max_depth), int(integrator.rr_depth)))
        _lib.check(lib.drtvam_scene_force_general(h, int(self.force_general)))
        return h

    def _apply_shard(self, h, sensor):
        """Shard state of a handle: the z window (z-slab partition; only sensors with the geometry of the one
        the partition was made for) and the positions of the local entries in the unsharded list, which seed
        the PCG32 streams so that a shard draws the samples of the unsharded run."""
        lib = _lib.load()
        if self.partition == "zslab":
            if not self._same_grid(sensor, self._z_sensor):
                raise NotImplementedError("a z-slab sharded scene can only render into sensors with the grid of the "
                                          "sensor it was sharded for; shard by angle (partition='angle') instead")
            r0, r1 = C.c_int32(0), C.c_int32(0)
            _lib.check(lib.drtvam_scene_set_z_window(h, self.z_window[0], self.z_window[1], C.byref(r0), C.byref(r1)))
        if self._sampler_index is not None:
            _lib.check(lib.drtvam_scene_set_sampler_index(h, self._sampler_index.data_ptr()))

    @staticmethod
    def _same_grid(a, b):
        return a is b or (tuple(a.resolution()) == tuple(b.resolution()) and np.array_equal(a.bbox_min, b.bbox_min)
                          and np.array_equal(a.bbox_max, b.bbox_max)
                          and a.film().surface_aware == b.film().surface_aware)

    def film_shape(self, sensor):
        """Shape of the dose buffers of this rank: the film, or its z window when z-slab sharded."""
        shape = tuple(sensor.film().shape)
        if self.partition == "zslab":
            return (self.z_window[1] - self.z_window[0],) + shape[1:]
        return shape

    def copy_stream(self):
        """The stream host <-> device copies of streamed calls run on."""
        if getattr(self, '_copy_stream', None) is None:
            self._copy_stream = torch.cuda.Stream(self.projector.active_pixels.device)
        return self._copy_stream

    def sampler_base(self, n):
        """Positions of the n local entries in the wavefront of the whole (unsharded, unblocked) call, as a
        device int32 tensor - what a call on a sub-range of the list must seed its PCG32 streams with - or
        None when the integrator draws no samples that depend on it."""
        if self._sampler_index is not None:
            return self._sampler_index
        if getattr(self, '_arange', None) is None or self._arange.numel() != n:
            self._arange = torch.arange(n, device=self.projector.active_pixels.device, dtype=torch.int32)
        return self._arange

    def pin_pixels(self, h, pixels):
        """Tells the library that `pixels` (projector.active_pixels, only ever REPLACED by params.update(),
        optimize.py:157-159) stays as it is, so the planar layout check runs once instead of per call."""
        if self._pinned.get(h) is not pixels:
            _lib.check(_lib.load().drtvam_scene_pin_pixels(h, pixels.data_ptr(), pixels.numel()))
            self._pinned[h] = pixels

    def filter_handle(self, integrator, sensor):
        """One-off C handle for the pixel pre-filters (optimize.py:143-185): they run before the target
        is moved away (optimize.py:188-191), so the mesh is attached whether or not the run is
        surface-aware.  The caller destroys it."""
        desc = self.describe(integrator, sensor)
        dev = self.projector.active_data.device
        if dev.type != 'cuda':
            raise RuntimeError("drtvam_b200 has no CPU path: scene tensors must live on a CUDA device")
        h, _ = create_handle(desc, dev.index or 0, self.all_target_triangles())
        return h

    def geometry_integrator(self):
        """Integrator options for calls that only need the geometry (compute_volume)."""
        if self.default_integrator is not None:
            return self.default_integrator
        from .integrators import VolumeIntegrator
        if getattr(self, '_geom_integrator', None) is None:
            self._geom_integrator = VolumeIntegrator({})
        return self._geom_integrator

    def close(self):
        if self._handles:
            lib = _lib.load()
            for h in self._handles.values():
                lib.drtvam_scene_destroy(h)
        self._handles.clear()
        self._keepalive.clear()
        self._handle_sensor.clear()
        self._pinned.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- multi-GPU ---------------------------------------------------------
    def shard(self, rank: int, world_size: int, group=None, partition: str = "angle", integrator=None, sensor=None,
              grid=None):
        """Keep only this rank's part of the projection.

        partition "angle": a contiguous range of projection angles balanced by active-pixel count; every
        forward is followed by one all_reduce(SUM) of the dose.  "zslab": for planar scenes (every ray
        stays in its voxel layer, sensor.py:345) a slab of film layers and the DMD rows that map to it -
        no dose collective for the layers a rank owns alone, the loss is a sharded sum.  The two combine:
        `grid` = (A, S) makes A angle groups x S slabs (rank = a * S + s); the dose of a slab is then
        all-reduced among the A ranks that share it, a tensor A times smaller than the film.  "auto" takes
        the slab grid parallel.slab_grid proposes where the scene allows it, else "angle".
        Returns (lo, hi, partition) with [lo, hi) the owned index range for a pure "angle" partition."""
        from .parallel import shard_bounds, slab_grid, zslab_shard
        p = self.projector
        self.rank, self.world_size, self.group = rank, world_size, group
        n_full = p.active_pixels.numel()
        self._full_active_pixels = p.active_pixels     # the unsharded list: gather_patterns / final pattern export
        # a surface-aware film keeps the angle partition: its tile kernel (csrc/inside_kernels.cu) needs the whole
        # layer range of the handle, and with it the angle shards are 3 x faster than slabs on the per-lane records
        if partition == "auto" and sensor is not None and getattr(sensor.film(), "surface_aware", False):
            partition = "angle"
        if partition in ("auto", "zslab") and world_size > 1 and integrator is not None:
            rz = int(sensor.resolution()[2])
            A, S = grid if grid is not None else (slab_grid(world_size, rz) if partition == "auto" else (1, world_size))
            if A * S != world_size:
                raise ValueError(f"grid {A} x {S} does not match the world size {world_size}")
            a, sl = rank // S, rank % S
            n_pat = p.n_patterns
            k_lo, k_hi = n_pat * a // A, n_pat * (a + 1) // A
            z = zslab_shard(self, sl, S, integrator, sensor, k_lo, k_hi) if S > 1 else None
            if z is not None:
                self.partition = "zslab"
                self.grid = (A, S)
                self._z_sensor = sensor
                self._z_rows = z["rows"]
                self.z_window, self.shard_index = z["window"], z["index"]
                p.active_pixels = p.active_pixels[z["index"]].contiguous()
                p.active_data = p.active_data[z["index"]].contiguous()
                self._make_groups(A, S, a, sl)
                self._shard_handles()
                return 0, n_full, "zslab"
            if partition == "zslab":
                raise ValueError("z-slab sharding needs a planar scene whose DMD rows map to whole voxel layers")
        lo, hi = shard_bounds(p.active_pixels, p.n_patterns, p.res[0] * p.res[1], rank, world_size)
        self.partition = "angle"
        self.grid = (world_size, 1)
        self.shard_index = torch.arange(lo, hi, device=p.active_pixels.device)
        p.active_pixels = p.active_pixels[lo:hi].contiguous()
        p.active_data = p.active_data[lo:hi].contiguous()
        self._shard_handles()
        return lo, hi, "angle"

    _full_active_pixels = None

    def gather_patterns(self, local=None):
        """The pattern vector over the UNSHARDED active-pixel list, assembled from the ranks' shards (every
        pixel is owned by exactly one rank, so a zero-filled scatter + all_reduce(SUM) is a gather); every
        rank gets it.  The reference keeps one vector (optimize.py:331-358 saves it); this is the step that
        makes a multi-GPU optimisation end with the same files."""
        local = (self.projector.active_data if local is None else local).detach()
        if self.world_size == 1 or self._full_active_pixels is None:
            return local
        full = torch.zeros(self._full_active_pixels.numel(), device=local.device, dtype=local.dtype)
        full[self.shard_index.long()] = local
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(full, op=dist.ReduceOp.SUM, group=self.group)
        return full

    def full_patterns(self, local=None):
        """Dense [n_patterns, h, w] pattern images of the whole job (projector.py:100-105 for one rank)."""
        p = self.projector
        full = self.gather_patterns(local)
        pixels = p.active_pixels if self._full_active_pixels is None or self.world_size == 1 else self._full_active_pixels
        n, h, w = p.size()
        out = torch.zeros(n * h * w, device=full.device, dtype=torch.float32)
        out[pixels.long() & 0xFFFFFFFF] = full
        return out.reshape(n, h, w)

    def _make_groups(self, A, S, a, sl):
        """Process groups of an A x S grid: the ranks that share a slab (they all-reduce its dose) and the
        ranks that share an angle group (their slabs make up the film: loss sums, gathers).  Every rank
        creates every group, in the same order, as torch.distributed requires."""
        self.dose_group = self.film_group = None
        self.dose_group_size, self.film_group_size = A, S
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() != A * S:
            return                                    # ranks run one after the other (tests, --emulate-shard)
        for s_ in range(S):
            g = dist.new_group([a_ * S + s_ for a_ in range(A)]) if A > 1 else None
            if s_ == sl:
                self.dose_group = g
        for a_ in range(A):
            g = dist.new_group([a_ * S + s_ for s_ in range(S)]) if A > 1 else None
            if a_ == a:
                self.film_group = g

    grid = (1, 1)
    dose_group = film_group = None
    dose_group_size = film_group_size = 1

    def _shard_handles(self):
        if self.shard_index.device.type == 'cuda':
            self._sampler_index = self.shard_index.to(torch.int64).to(torch.int32).contiguous()
        self._pinned.clear()
        for h in self._handles.values():
            self._apply_shard(h, self._handle_sensor[h])

    partition = "single"
    z_window = None
    shard_index = None

    def reduce_dose(self, dose: torch.Tensor):
        """all_reduce(SUM) of the partial doses of the ranks that project different angles into the same
        layers: everybody for an angle partition, the angle groups of one slab in a grid, nobody when a
        rank owns its slab alone."""
        if self.world_size <= 1:
            return
        import torch.distributed as dist
        if self.partition != "zslab":
            dist.all_reduce(dose, op=dist.ReduceOp.SUM, group=self.group)
        elif self.dose_group_size > 1 and self.dose_group is not None:
            dist.all_reduce(dose, op=dist.ReduceOp.SUM, group=self.dose_group)

    def local_target(self, target: torch.Tensor) -> torch.Tensor:
        """The part of a film-shaped tensor (loss target, dose gradient) this rank's film holds."""
        if self.partition != "zslab":
            return target
        z0, z1 = self.z_window
        return target[z0:z1].contiguous()

    local_dose_grad = local_target

    def gather_dose(self, dose: torch.Tensor) -> torch.Tensor:
        """The full dose volume on every rank (checks, final outputs): identity unless z-slab sharded."""
        if self.partition != "zslab" or self.world_size == 1:
            return dose
        import torch.distributed as dist
        S = self.film_group_size
        z0, z1 = self.z_window
        bounds = [None] * S
        dist.all_gather_object(bounds, (z0, z1), group=self.film_group)
        nz = max(b[1] - b[0] for b in bounds)
        pad = torch.zeros((nz,) + tuple(dose.shape[1:]), device=dose.device, dtype=dose.dtype)
        pad[:z1 - z0] = dose
        parts = [torch.empty_like(pad) for _ in range(S)]
        dist.all_gather(parts, pad, group=self.film_group)
        return torch.cat([parts[r][:bounds[r][1] - bounds[r][0]] for r in range(S)], 0)

    def dose_allsum(self):
        """Sum of a scalar reduced over the local film across the slabs of the film (z-slab sharding): over the
        ranks of this rank's angle group, each of which holds one slab, fully reduced.  Else None."""
        if self.partition != "zslab" or self.world_size == 1:
            return None
        from .parallel import make_sum
        return make_sum(self.film_group, self.film_group_size)


class Params(dict):
    """What ``mi.traverse(scene)`` returns for this path: the keys the
    reference touches (``projector.active_data`` differentiable,
    ``projector.active_pixels``; optimize.py:141,157-159)."""

    def __init__(self, scene: Scene):
        super().__init__()
        self.scene = scene
        super().__setitem__('projector.active_data', scene.projector.active_data)
        super().__setitem__('projector.active_pixels', scene.projector.active_pixels)

    def update(self, values=None):
        if values is not None:
            for k, v in dict(values).items():
                if k in self:
                    self[k] = v
        p = self.scene.projector
        data = self['projector.active_data']
        pixels = self['projector.active_pixels']
        if not isinstance(data, torch.Tensor):
            data = torch.as_tensor(np.asarray(data, np.float32))
        if not isinstance(pixels, torch.Tensor):
            pixels = torch.as_tensor(np.asarray(pixels, np.int64))
        dev = p.active_data.device
        data = data.to(dev, torch.float32)      # no-op (same object, .grad kept) when already there
        p.active_data = data if data.dim() == 1 else data.reshape(-1)
        if pixels.dtype != torch.int32:
            pixels = pixels.to(torch.int64).to(torch.int32)
        p.active_pixels = pixels.to(dev).reshape(-1).contiguous()
        super().__setitem__('projector.active_data', p.active_data)
        super().__setitem__('projector.active_pixels', p.active_pixels)
        p.parameters_changed()
        return []


def traverse(scene: Scene) -> Params:
    return Params(scene)


def load_dict(scene_dict: dict, device=None) -> Scene:
    """Builds a Scene from the dict ``optimize.load_scene`` assembles
    (optimize.py:61-77): one projector, ``sensor`` [+ ``final_sensor``], the
    container shapes, ``printing_medium`` and (optionally) ``target``."""
    if scene_dict.get('type', 'scene') != 'scene':
        raise ValueError("load_dict expects a dict of type 'scene'")
    projector = None
    sensors: Dict[str, VolumetricSensor] = {}
    surfaces: List[SurfaceDesc] = []
    medium = None
    target = None
    integrator = None
    for name, entry in scene_dict.items():
        if not isinstance(entry, dict) and not isinstance(entry, (TVAMProjector, VolumetricSensor)):
            continue
        if isinstance(entry, TVAMProjector):
            projector = entry
            continue
        if isinstance(entry, VolumetricSensor):
            sensors[name] = entry
            continue
        t = entry.get('type')
        if t in projectors:
            if projector is not None:
                raise Exception("The scene contains more than one projector. Only one is supported")
            projector = projectors[t](dict(entry, device=device) if device is not None else entry)
        elif t in _sensors:
            props = dict(entry, id=name)
            if device is not None:
                props['device'] = device
            sensors[name] = _sensors[t](props)
        elif t in _integrators:
            integrator = _integrators[t](entry)
        elif t == 'homogeneous':
            if medium is not None and name != 'printing_medium':
                raise ValueError("There is more than one medium in the scene. Only one is supported")
            medium = entry
        elif t in ('cylinder', 'cube'):
            inline = entry.get('interior') if isinstance(entry.get('interior'), dict) else None
            if inline is not None and inline.get('type') == 'homogeneous':
                if medium is not None:
                    raise ValueError("There is more than one medium in the scene. Only one is supported")
                medium = inline
            surfaces.append(surface_from_shape_dict(name, entry))
        elif t in ('ply', 'obj'):
            if name != 'target':
                # custom vial wall or occluder (geometry.py:55-72,113-134)
                surfaces.append(surface_from_shape_dict(name, entry))
            elif name == 'target':
                from .utils import load_mesh
                from .transform import Transform4f
                verts, tris = load_mesh(entry['filename'])
                tw = entry.get('to_world')
                if tw is not None:
                    verts = Transform4f(tw) @ verts
                target = dict(entry, vertices=np.asarray(verts, np.float64), triangles=tris)
        else:
            raise NotImplementedError(f"scene entry '{name}' of type '{t}' is not supported")
    if projector is None:
        raise Exception("No projector found in the scene")
    if medium is None:
        raise ValueError("No printing medium found in the scene")
    if not sensors:
        raise ValueError("No sensor found in the scene")
    return Scene(projector, sensors, surfaces, medium, target, integrator)


class _RenderOp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, active_data, scene, integrator, sensor, spp, spp_grad, seed, seed_grad):
        ctx.args = (scene, integrator, sensor, spp_grad, seed_grad)
        return integrator._project(scene, sensor, seed, spp, active_data)

    @staticmethod
    def backward(ctx, grad_out):
        scene, integrator, sensor, spp_grad, seed_grad = ctx.args
        g = integrator.render_backward(scene, None, grad_out, sensor, seed_grad, spp_grad)
        return g, None, None, None, None, None, None, None


def render(scene: Scene, params: Optional[Params] = None, sensor=0, integrator=None,
           seed: int = 0, seed_grad: int = 0, spp: int = 0, spp_grad: int = 0) -> torch.Tensor:
    """Differentiable forward projection; same signature as ``mi.render``."""
    if integrator is None:
        integrator = scene.integrator()
        if integrator is None:
            raise ValueError("No integrator specified")
    if isinstance(sensor, int):
        sensor = scene.sensors()[sensor]
    if spp_grad == 0:
        spp_grad = spp
    if seed_grad == 0:
        # de-correlate primal and adjoint samples like mi.render (App. A.1)
        seed_grad = sample_tea_32(seed, 1)[0]
    data = params['projector.active_data'] if params is not None else scene.projector.active_data
    if params is not None and data is not scene.projector.active_data:
        params.update()
        data = params['projector.active_data']
    if torch.is_grad_enabled() and data.requires_grad:
        return _RenderOp.apply(data, scene, integrator, sensor, spp, spp_grad, seed, seed_grad)
    with torch.no_grad():
        return integrator._project(scene, sensor, seed, spp, data)
