"""ctypes wrapper of the CPU oracle (``oracle/drtvam_oracle.c``).

TEST INFRASTRUCTURE ONLY -- see the header of drtvam_oracle.c.  Imported by
tests/, ``__graft_entry__.smoke()`` and bench.py's CPU-baseline legs; never by
the product package ``drtvam_b200``.  PARITY UNPINNED (no runnable Mitsuba /
Dr.Jit here; the reference's tests hold no golden volumes).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "_build")
MAX_SURF = 8


class OrcSurface(C.Structure):
    _fields_ = [("kind", C.c_int32), ("bsdf", C.c_int32), ("interior_medium", C.c_int32),
                ("exterior_medium", C.c_int32), ("radius", C.c_double), ("z0", C.c_double), ("z1", C.c_double),
                ("hx", C.c_double), ("hy", C.c_double), ("hz", C.c_double), ("eta", C.c_double),
                ("n_tri", C.c_int32), ("_pad_tri", C.c_int32), ("tri", C.POINTER(C.c_double))]


class OrcScene(C.Structure):
    _fields_ = [
        ("proj_kind", C.c_int32), ("n_patterns", C.c_int32), ("res_x", C.c_int32), ("res_y", C.c_int32),
        ("pixel_x", C.c_double), ("pixel_y", C.c_double), ("aperture_radius", C.c_double),
        ("focus_distance", C.c_double),
        ("motion_analytic", C.c_int32), ("clockwise", C.c_int32), ("distance", C.c_double),
        ("motion_table", C.POINTER(C.c_double)),
        ("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("transmission_only", C.c_int32),
        ("regular_sampling", C.c_int32), ("sample_time", C.c_int32), ("print_time", C.c_double),
        ("sigma_t", C.c_double), ("albedo", C.c_double), ("phase_kind", C.c_int32), ("hg_g", C.c_double),
        ("n_surf", C.c_int32), ("surf", OrcSurface * MAX_SURF),
        ("bbox_min", C.c_double * 3), ("bbox_max", C.c_double * 3), ("res", C.c_int32 * 3),
        ("surface_aware", C.c_int32), ("sensor_kind", C.c_int32), ("majorant", C.c_double),
        ("n_tri", C.c_int32), ("_pad_tri", C.c_int32), ("tri", C.POINTER(C.c_double)),
    ]


def build(force: bool = False):
    """Compile the two oracle builds (f32 voxel-exact, f64 truth) with make."""
    libs = [os.path.join(BUILD, f"liboracle_{p}.so") for p in ("f32", "f64")]
    src = os.path.join(HERE, "drtvam_oracle.c")
    stale = force or any(not os.path.exists(p) or os.path.getmtime(p) < os.path.getmtime(src) for p in libs)
    if stale:
        subprocess.run(["make", "-C", HERE, "-B" if force else "-s"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    return libs


_libs = {}


def _lib(precision: str):
    if precision not in _libs:
        path = os.path.join(BUILD, f"liboracle_{precision}.so")
        if not os.path.exists(path):
            build()
        lib = C.CDLL(path)
        real = C.c_float if precision == "f32" else C.c_double
        lib.orc_forward.restype = C.c_int
        lib.orc_forward.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                    C.c_void_p, C.POINTER(C.c_uint64), C.c_int]
        lib.orc_adjoint.restype = C.c_int
        lib.orc_adjoint.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                    C.c_void_p, C.c_int]
        lib.orc_raygen.restype = C.c_int
        lib.orc_raygen.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p]
        lib.orc_radon.restype = C.c_int
        lib.orc_radon.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p]
        lib.orc_corner.restype = C.c_int
        lib.orc_corner.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_double,
                                   C.c_double, C.c_void_p]
        lib.orc_compute_volume.restype = C.c_int
        lib.orc_compute_volume.argtypes = [C.POINTER(OrcScene), C.c_uint32, C.c_void_p, C.c_int]
        lib.orc_discretize.restype = C.c_int
        lib.orc_discretize.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_int]
        lib.orc_rng_stream.restype = C.c_int
        lib.orc_rng_stream.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]
        assert lib.orc_sizeof_real() == C.sizeof(real)
        _libs[precision] = (lib, np.float32 if precision == "f32" else np.float64)
    return _libs[precision]


class Oracle:
    """One scene, evaluated on the CPU.  `desc` is SceneDesc.as_dict() (or any
    dict with the same keys)."""

    def __init__(self, desc: dict, precision: str = "f64", threads: int = 0):
        self.lib, self.real = _lib(precision)
        self.threads = threads
        s = OrcScene()
        for k in ("proj_kind", "n_patterns", "res_x", "res_y", "pixel_x", "pixel_y", "aperture_radius",
                  "focus_distance", "motion_analytic", "clockwise", "distance", "max_depth", "rr_depth",
                  "transmission_only", "regular_sampling", "sample_time", "print_time", "sigma_t", "albedo",
                  "phase_kind", "hg_g", "surface_aware", "sensor_kind", "majorant"):
            setattr(s, k, desc[k])
        if desc["sensor_kind"] == 2 and not desc["albedo"] > 0:  # volume.py:156-157
            raise ValueError("Tried to render a purely absorptive volume with a delta tracking sensor. "
                             "This is not supported.")
        self._table = None
        if desc.get("motion_table") is not None:
            self._table = np.ascontiguousarray(np.asarray(desc["motion_table"], np.float64).reshape(-1, 16))
            s.motion_table = self._table.ctypes.data_as(C.POINTER(C.c_double))
        surfs = desc["surfaces"]
        s.n_surf = len(surfs)
        self._surf_tri = []
        for i, sd in enumerate(surfs):
            for k in ("kind", "bsdf", "interior_medium", "exterior_medium", "radius", "z0", "z1", "hx", "hy", "hz",
                      "eta"):
                setattr(s.surf[i], k, sd[k])
            if sd["kind"] == 2:  # mesh interface: [n_tri][3][3] world-space vertices
                tri = np.ascontiguousarray(np.asarray(sd["triangles"], np.float64).reshape(-1, 9))
                self._surf_tri.append(tri)
                s.surf[i].n_tri = tri.shape[0]
                s.surf[i].tri = tri.ctypes.data_as(C.POINTER(C.c_double))
        for a in range(3):
            s.bbox_min[a] = desc["bbox_min"][a]
            s.bbox_max[a] = desc["bbox_max"][a]
            s.res[a] = int(desc["res"][a])
        # target mesh: [n_tri][3][3] world-space vertices (only passed for surface-aware runs)
        self._tri = None
        if desc.get("target_triangles") is not None:
            self._tri = np.ascontiguousarray(np.asarray(desc["target_triangles"], np.float64).reshape(-1, 9))
            s.n_tri = self._tri.shape[0]
            s.tri = self._tri.ctypes.data_as(C.POINTER(C.c_double))
        self.scene = s
        rx, ry, rz = (int(v) for v in desc["res"])
        self.film_shape = (rz, ry, rx, 2 if desc["surface_aware"] else 1)
        self.regular = bool(desc["regular_sampling"])
        self.last_visits = 0

    def forward(self, active_data, active_pixels, spp=1, seed=0):
        """Unnormalised film (before the 1/voxel-volume scale), shape (rz, ry, rx, C)."""
        data = np.ascontiguousarray(active_data, np.float32)
        pix = np.ascontiguousarray(active_pixels, np.uint32)
        film = np.zeros(self.film_shape, self.real)
        visits = C.c_uint64(0)
        rc = self.lib.orc_forward(C.byref(self.scene), data.ctypes.data, pix.ctypes.data, data.size, spp, seed,
                                  film.ctypes.data, C.byref(visits), self.threads)
        assert rc == 0
        self.last_visits = int(visits.value)
        return film

    def adjoint(self, dL, active_pixels, spp=1, seed=0):
        """d/d active_data of <film, dL>; dL already carries the 1/volume scale."""
        g = np.ascontiguousarray(np.asarray(dL).reshape(self.film_shape), self.real)
        pix = np.ascontiguousarray(active_pixels, np.uint32)
        out = np.zeros(pix.size, self.real)
        rc = self.lib.orc_adjoint(C.byref(self.scene), g.ctypes.data, pix.ctypes.data, pix.size, spp, seed,
                                  out.ctypes.data, self.threads)
        assert rc == 0
        return out

    def radon(self, active_pixels, spp=1, seed=0):
        """RadonIntegrator.render (integrators/radon.py:13-46): image (n_patterns, res_y, res_x)."""
        pix = np.ascontiguousarray(active_pixels, np.uint32)
        img = np.zeros((self.scene.n_patterns, self.scene.res_y, self.scene.res_x), self.real)
        assert self.lib.orc_radon(C.byref(self.scene), pix.ctypes.data, pix.size, spp, seed, img.ctypes.data) == 0
        return img

    def corner(self, active_pixels, dist, radius=0.1, spp=1, seed=0):
        """CornerIntegrator.render (integrators/filter_corner.py:17-66)."""
        pix = np.ascontiguousarray(active_pixels, np.uint32)
        img = np.zeros((self.scene.n_patterns, self.scene.res_y, self.scene.res_x), self.real)
        assert self.lib.orc_corner(C.byref(self.scene), pix.ctypes.data, pix.size, spp, seed, float(dist),
                                   float(radius), img.ctypes.data) == 0
        return img

    def compute_volume(self, sample_count=2 ** 14):
        """Surface-aware sub-voxel volumes (sensor.py:47-110), shape (rz, ry, rx, 2) = (in, out)."""
        rz, ry, rx, _ = self.film_shape
        out = np.zeros((rz, ry, rx, 2), self.real)
        rc = self.lib.orc_compute_volume(C.byref(self.scene), int(sample_count), out.ctypes.data, self.threads)
        assert rc == 0, "compute_volume needs a target mesh"
        return out

    def discretize(self):
        """Binary occupancy of the target at the voxel centres (utils.py:83-128), shape (rz, ry, rx, 1)."""
        rz, ry, rx, _ = self.film_shape
        out = np.zeros((rz, ry, rx, 1), self.real)
        rc = self.lib.orc_discretize(C.byref(self.scene), out.ctypes.data, self.threads)
        assert rc == 0, "discretize needs a target mesh"
        return out

    def raygen(self, active_pixels, spp=1, seed=0):
        pix = np.ascontiguousarray(active_pixels, np.uint32)
        n = pix.size * (1 if self.regular else spp)
        out = np.zeros((n, 6), self.real)
        rc = self.lib.orc_raygen(C.byref(self.scene), pix.ctypes.data, pix.size, spp, seed, out.ctypes.data)
        assert rc == 0
        return out


def rng_stream(seed: int, lane: int, n: int, precision: str = "f64"):
    lib, _ = _lib(precision)
    out = np.zeros(n, np.float32)
    lib.orc_rng_stream(seed, lane, n, out.ctypes.data)
    return out
