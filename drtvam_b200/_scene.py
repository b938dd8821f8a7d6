"""Flat scene description: what ``mi.load_dict(scene_dict)`` builds for the hot
path in the reference (``optimize.py:61-73``, ``geometry.py`` ``to_dict()``),
reduced to the PODs the CUDA kernels (and, in tests, the CPU oracle) consume.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field, asdict
from typing import List, Optional

import numpy as np

from . import _lib

SURF_CYLINDER, SURF_BOX, SURF_MESH = 0, 1, 2
BSDF_NULL, BSDF_DIELECTRIC, BSDF_ABSORBER = 0, 1, 2
PROJ_KINDS = {"collimated": 0, "telecentric": 1, "lens": 2}
PHASE_KINDS = {"isotropic": 0, "rayleigh": 1, "hg": 2}
SENSOR_KINDS = {"dda": 0, "ratio": 1, "delta": 2}

# Mitsuba's named indices of refraction (dielectric plugin table); "air" is the
# default ext_ior the square vial's outer wall relies on (geometry.py:201-204).
IOR_TABLE = {
    "vacuum": 1.0, "helium": 1.000036, "hydrogen": 1.000132, "air": 1.000277,
    "carbon dioxide": 1.00045, "water": 1.3330, "acetone": 1.36, "ethanol": 1.361,
    "carbon tetrachloride": 1.461, "glycerol": 1.4729, "benzene": 1.501,
    "silicone oil": 1.52045, "bromine": 1.661, "water ice": 1.31, "fused quartz": 1.458,
    "pyrex": 1.470, "acrylic glass": 1.49, "polypropylene": 1.49, "bk7": 1.5046,
    "sodium chloride": 1.544, "amber": 1.55, "pet": 1.5750, "diamond": 2.419,
}


def lookup_ior(value, default: str) -> float:
    if value is None:
        value = default
    if isinstance(value, str):
        if value not in IOR_TABLE:
            raise ValueError(f"Unknown material name '{value}' for an index of refraction")
        return IOR_TABLE[value]
    return float(value)


@dataclass
class SurfaceDesc:
    kind: int = SURF_CYLINDER
    bsdf: int = BSDF_NULL
    interior_medium: int = 0
    exterior_medium: int = 0
    radius: float = 0.0
    z0: float = 0.0
    z1: float = 0.0
    hx: float = 0.0
    hy: float = 0.0
    hz: float = 0.0
    eta: float = 1.0
    name: str = ""
    # SURF_MESH: [n_tri, 3, 3] float32 world-space vertices (uploaded with
    # drtvam_scene_add_mesh_surface, not part of the C struct)
    triangles: object = None


@dataclass
class SceneDesc:
    # projector
    proj_kind: int = 0
    n_patterns: int = 1
    res_x: int = 1
    res_y: int = 1
    pixel_x: float = 1.0
    pixel_y: float = 1.0
    aperture_radius: float = 0.0
    focus_distance: float = 1.0
    # motion
    motion_analytic: int = 0
    clockwise: int = 0
    distance: float = 0.0
    motion_table: Optional[np.ndarray] = None  # [n_patterns, 4, 4] float64
    # integrator
    max_depth: int = 6
    rr_depth: int = 6
    transmission_only: int = 1
    regular_sampling: int = 0
    sample_time: int = 0
    print_time: float = 1.0
    # medium
    sigma_t: float = 1.0
    albedo: float = 0.0
    phase_kind: int = 0
    hg_g: float = 0.0
    # surfaces
    surfaces: List[SurfaceDesc] = field(default_factory=list)
    # sensor / film
    bbox_min: tuple = (-0.5, -0.5, -0.5)
    bbox_max: tuple = (0.5, 0.5, 0.5)
    res: tuple = (1, 1, 1)
    surface_aware: int = 0
    sensor_kind: int = 0
    majorant: float = 0.0
    # [n_tri, 3, 3] world-space target mesh of a surface-aware run (uploaded with
    # drtvam_scene_set_target, not part of the C struct), else None
    target_triangles: object = None

    def film_shape(self):
        rx, ry, rz = self.res
        return (rz, ry, rx, 2 if self.surface_aware else 1)

    def voxel_volume(self) -> float:
        # sensor.py:19,51 in fp32 like the reference
        ext = (np.asarray(self.bbox_max, np.float32) - np.asarray(self.bbox_min, np.float32))
        vs = ext / np.asarray(self.res, np.float32)
        return float(np.prod(vs, dtype=np.float32))

    def as_dict(self) -> dict:
        tri, self.target_triangles = self.target_triangles, None
        d = asdict(self)
        self.target_triangles = tri
        d["target_triangles"] = tri
        d["surfaces"] = [asdict(s) for s in self.ordered_surfaces()]
        return d

    def ordered_surfaces(self):
        """Closed-form surfaces first, then the meshes: the order they have on the device (the
        meshes are appended to the handle after drtvam_scene_create)."""
        return [s for s in self.surfaces if s.kind != SURF_MESH] + [s for s in self.surfaces if s.kind == SURF_MESH]

    def to_c(self):
        """Returns (DrtvamSceneDesc, keepalive) for drtvam_scene_create."""
        if len(self.surfaces) > _lib.MAX_SURFACES:
            raise ValueError(f"too many vial surfaces ({len(self.surfaces)} > {_lib.MAX_SURFACES})")
        d = _lib.DrtvamSceneDesc()
        d.struct_size = C.sizeof(_lib.DrtvamSceneDesc)
        d.abi_version = _lib.ABI_VERSION
        for k in ("proj_kind", "n_patterns", "res_x", "res_y", "pixel_x", "pixel_y", "aperture_radius",
                  "focus_distance", "motion_analytic", "clockwise", "distance", "max_depth", "rr_depth",
                  "transmission_only", "regular_sampling", "sample_time", "print_time", "sigma_t", "albedo",
                  "phase_kind", "hg_g", "surface_aware", "sensor_kind", "majorant"):
            setattr(d, k, getattr(self, k))
        keep = None
        if self.motion_table is not None:
            tab = np.ascontiguousarray(np.asarray(self.motion_table, np.float64)[:, :3, :].astype(np.float32))
            tab64 = np.ascontiguousarray(np.asarray(self.motion_table, np.float64)[:, :3, :])
            keep = (tab, tab64)
            d.motion_table_host = tab.ctypes.data_as(C.POINTER(C.c_float))
            d.motion_table_host_f64 = tab64.ctypes.data_as(C.POINTER(C.c_double))
        closed = [s for s in self.surfaces if s.kind != SURF_MESH]
        d.n_surf = len(closed)
        for i, s in enumerate(closed):
            c = d.surf[i]
            for k in ("kind", "bsdf", "interior_medium", "exterior_medium", "radius", "z0", "z1", "hx", "hy", "hz",
                      "eta"):
                setattr(c, k, getattr(s, k))
        for a in range(3):
            d.bbox_min[a] = self.bbox_min[a]
            d.bbox_max[a] = self.bbox_max[a]
            d.res[a] = int(self.res[a])
        return d, keep


def _bsdf_from_dict(b: Optional[dict]):
    """Returns (bsdf kind, eta).  Mitsuba defaults: no bsdf -> diffuse."""
    if b is None:
        return BSDF_ABSORBER, 1.0
    t = b.get("type")
    if t == "null":
        return BSDF_NULL, 1.0
    if t == "dielectric":
        int_ior = lookup_ior(b.get("int_ior"), "bk7")
        ext_ior = lookup_ior(b.get("ext_ior"), "air")
        return BSDF_DIELECTRIC, int_ior / ext_ior
    if t == "diffuse":
        return BSDF_ABSORBER, 1.0
    raise NotImplementedError(f"BSDF type '{t}' has no closed-form interface kernel")


def _is_medium_ref(v) -> bool:
    if v is None:
        return False
    if isinstance(v, dict):
        return v.get("type") == "homogeneous" or (v.get("type") == "ref" and v.get("id") == "printing_medium")
    return False


def surface_from_shape_dict(name: str, sd: dict) -> SurfaceDesc:
    """One Mitsuba shape dict of a container (geometry.py to_dict()) -> SurfaceDesc."""
    bsdf, eta = _bsdf_from_dict(sd.get("bsdf"))
    s = SurfaceDesc(bsdf=bsdf, eta=eta, name=name,
                    interior_medium=int(_is_medium_ref(sd.get("interior"))),
                    exterior_medium=int(_is_medium_ref(sd.get("exterior"))))
    t = sd["type"]
    if t == "cylinder":
        p0 = np.asarray(sd.get("p0", [0, 0, 0]), np.float64)
        p1 = np.asarray(sd.get("p1", [0, 0, 1]), np.float64)
        if abs(p0[0]) + abs(p0[1]) + abs(p1[0]) + abs(p1[1]) > 0:
            raise NotImplementedError("only cylinders about the z axis have a closed-form kernel")
        s.kind = SURF_CYLINDER
        s.radius = float(sd.get("radius", 1.0))
        s.z0, s.z1 = float(min(p0[2], p1[2])), float(max(p0[2], p1[2]))
    elif t == "cube":
        tw = sd.get("to_world")
        m = np.eye(4) if tw is None else np.asarray(getattr(tw, "matrix", tw), np.float64).reshape(4, 4)
        off = m.copy()
        off[0, 0] = off[1, 1] = off[2, 2] = 0
        off[3, 3] = 0
        if np.abs(off).max() > 0:
            raise NotImplementedError("only origin-centred axis-aligned boxes have a closed-form kernel")
        s.kind = SURF_BOX
        s.hx, s.hy, s.hz = float(m[0, 0]), float(m[1, 1]), float(m[2, 2])
    elif t in ("ply", "obj"):
        # custom vial walls (geometry.py:113-134) and occluders (geometry.py:55-72): a triangle
        # soup with face normals, walked through a BVH on the device
        from .transform import Transform4f
        from .utils import load_mesh
        verts, tris = load_mesh(sd["filename"])
        tw = sd.get("to_world")
        if tw is not None:
            verts = Transform4f(tw) @ verts
        s.kind = SURF_MESH
        s.triangles = np.ascontiguousarray(np.asarray(verts, np.float64)[np.asarray(tris)].astype(np.float32))
    else:
        raise NotImplementedError(f"shape type '{t}' is not supported")
    return s


def create_handle(desc: SceneDesc, device_index: int, target_triangles=None):
    """drtvam_scene_create + the mesh interfaces + (optionally) the target mesh.  Returns
    (handle, keepalive); the caller destroys the handle."""
    lib = _lib.load()
    cdesc, keep = desc.to_c()
    out = C.c_void_p()
    _lib.check(lib.drtvam_scene_create(C.byref(cdesc), device_index, C.byref(out)))
    h = out.value
    try:
        for s in desc.ordered_surfaces():
            if s.kind != SURF_MESH:
                continue
            tri = np.ascontiguousarray(s.triangles, np.float32)
            _lib.check(lib.drtvam_scene_add_mesh_surface(h, tri.ctypes.data, tri.shape[0], int(s.bsdf), float(s.eta),
                                                         int(s.interior_medium), int(s.exterior_medium)))
        if target_triangles is not None:
            tri = np.ascontiguousarray(target_triangles, np.float32)
            _lib.check(lib.drtvam_scene_set_target(h, tri.ctypes.data, tri.shape[0]))
    except Exception:
        lib.drtvam_scene_destroy(h)
        raise
    return h, keep
