"""Synthetic benchmark / test scenes in the reference's config.json schema
(SURVEY.md section 8(d), configs C1-C5).  The Benchy asset is not available
(`.MISSING_LARGE_BLOBS`), so targets are procedural meshes of the same
bounding size: an axis-aligned cube, or a box with a cylindrical hole (the
shape of the reference's ``tests/files/box_hole.ply`` fixture, regenerated
here rather than copied).
"""
from __future__ import annotations

import os

import numpy as np

from .utils import write_ply


def cube_mesh(side=1.0):
    h = 0.5 * side
    v = np.array([[x, y, z] for x in (-h, h) for y in (-h, h) for z in (-h, h)], np.float64)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = []
    for a, b, c, d in quads:
        tris += [[a, b, c], [a, c, d]]
    return v, np.asarray(tris)


def box_mesh(sx, sy, sz, caps=True):
    """Axis-aligned box centred on the origin, outward-facing; `caps=False` leaves the +-z faces out
    (the open cuvette walls of tests/files/cuvette_*.ply)."""
    v, t = cube_mesh(1.0)
    v = v * np.array([sx, sy, sz], np.float64)
    if not caps:
        n = np.cross(v[t[:, 1]] - v[t[:, 0]], v[t[:, 2]] - v[t[:, 0]])
        t = t[np.abs(n[:, 2]) < 1e-12 * np.abs(n).max()]
    return v, t


def box_hole_mesh(sx=80.0, sy=80.0, sz=40.0, hole_r=20.0, hole_cx=-10.0, hole_cy=0.0, segments=48):
    """Box [-sx/2,sx/2]x[-sy/2,sy/2]x[-sz/2,sz/2] with a through hole along z
    (outward-facing, watertight).  Defaults reproduce the analytic reference
    volume of tests/test_optimization.py:130-144 (in voxel units)."""
    hx, hy, hz = 0.5 * sx, 0.5 * sy, 0.5 * sz
    ang = np.linspace(0, 2 * np.pi, segments, endpoint=False)
    ring = np.stack([hole_cx + hole_r * np.cos(ang), hole_cy + hole_r * np.sin(ang)], 1)
    corners = np.array([[-hx, -hy], [hx, -hy], [hx, hy], [-hx, hy]])
    verts, tris = [], []

    def add(p):
        verts.append(p)
        return len(verts) - 1

    # outer walls
    for i in range(4):
        a, b = corners[i], corners[(i + 1) % 4]
        i0, i1 = add([a[0], a[1], -hz]), add([b[0], b[1], -hz])
        i2, i3 = add([b[0], b[1], hz]), add([a[0], a[1], hz])
        tris += [[i0, i1, i2], [i0, i2, i3]]
    # hole wall (normals point into the hole = out of the solid)
    for i in range(segments):
        a, b = ring[i], ring[(i + 1) % segments]
        i0, i1 = add([a[0], a[1], -hz]), add([b[0], b[1], -hz])
        i2, i3 = add([b[0], b[1], hz]), add([a[0], a[1], hz])
        tris += [[i0, i2, i1], [i0, i3, i2]]
    # caps: the region between the ring and the rectangle is star-shaped about
    # the hole centre, so ring vertices are pushed radially onto the rectangle
    def to_rect(p):
        d = p - np.array([hole_cx, hole_cy])
        ts = []
        for axis, lim in ((0, hx), (0, -hx), (1, hy), (1, -hy)):
            if d[axis] != 0:
                t = (lim - (hole_cx, hole_cy)[axis]) / d[axis]
                if t > 0:
                    ts.append(t)
        return np.array([hole_cx, hole_cy]) + min(ts) * d

    outer = np.array([to_rect(p) for p in ring])
    outer_ang = np.arctan2(outer[:, 1] - hole_cy, outer[:, 0] - hole_cx) % (2 * np.pi)
    corner_ang = np.arctan2(corners[:, 1] - hole_cy, corners[:, 0] - hole_cx) % (2 * np.pi)
    for z, up in ((hz, True), (-hz, False)):
        ri = [add([p[0], p[1], z]) for p in ring]
        oi = [add([p[0], p[1], z]) for p in outer]
        ci = [add([c[0], c[1], z]) for c in corners]

        def emit(t):
            tris.append(list(t) if up else list(t)[::-1])

        for i in range(segments):
            j = (i + 1) % segments
            emit((ri[i], oi[i], oi[j]))
            emit((ri[i], oi[j], ri[j]))
            a0, a1 = outer_ang[i], outer_ang[j]
            for k in range(4):
                ca = corner_ang[k]
                inside = (a0 < ca < a1) if a0 < a1 else (ca > a0 or ca < a1)
                if inside:
                    emit((oi[i], ci[k], oi[j]))
    return np.asarray(verts, np.float64), np.asarray(tris)


def write_target(path, kind="box_hole"):
    if kind == "cube":
        v, t = cube_mesh(1.0)
    elif kind == "box_hole":
        v, t = box_hole_mesh()
    else:
        raise ValueError(kind)
    write_ply(path, v, t)
    return path


MEDIUM_ABSORPTIVE = {"ior": 1.347, "phase": {"type": "rayleigh"}, "extinction": 0.03, "albedo": 0.0}


def make_config(name: str, workdir: str, n=None, n_patterns=None, spp=None, regular_sampling=True,
                n_steps=None, grid=None, surface_aware=False):
    """Config dict for synthetic config `name` in {C1..C5}; `n` overrides the
    DMD / grid edge (default: the full-size value), `n_patterns` the angle count."""
    name = name.upper()
    full = {"C1": 128, "C2": 400, "C3": 400, "C4": 400, "C5": 400}[name]
    n = n or full
    grid = grid or n
    n_patterns = n_patterns or n
    extent = 6.4 if name == "C1" else 10.0
    pixel = extent / n
    os.makedirs(workdir, exist_ok=True)
    target = os.path.join(workdir, "target_cube.ply" if name == "C1" else "target_box_hole.ply")
    if not os.path.exists(target):
        write_target(target, "cube" if name == "C1" else "box_hole")
    cfg = {
        "projector": {"type": "collimated", "n_patterns": n_patterns, "resx": n, "resy": n,
                      "pixel_size": pixel, "motion": "circular", "distance": 20},
        "sensor": {"type": "dda", "scalex": extent, "scaley": extent, "scalez": extent,
                   "film": {"type": "vfilm", "resx": grid, "resy": grid, "resz": grid}},
        "target": {"filename": target, "size": 3.2 if name == "C1" else 7.0},
        "loss": {"type": "threshold"},
        "spp": 1, "spp_ref": 1, "max_depth": 6, "rr_depth": 6,
        "regular_sampling": regular_sampling,
        "n_steps": n_steps or (5 if name == "C1" else 40),
        "output": os.path.join(workdir, f"out_{name}"),
    }
    if name == "C1":
        cfg["vial"] = {"type": "index_matched", "r": 5.0, "medium": dict(MEDIUM_ABSORPTIVE)}
        cfg["progressive"] = True
    elif name == "C2":
        cfg["vial"] = {"type": "index_matched", "r": 7.5,
                       "medium": {"ior": 1.347, "extinction": 0.1, "albedo": 0.0}}
    elif name in ("C3", "C5"):
        cfg["vial"] = {"type": "cylindrical", "r_int": 7.5, "r_ext": 8.0, "ior": 1.54,
                       "medium": {"ior": 1.49, "extinction": 0.1, "albedo": 0.0}}
        if name == "C5":
            cfg["vial"]["medium"].update({"albedo": 0.9, "phase": {"type": "hg", "g": 0.95}})
            cfg["spp"] = cfg["spp_ref"] = cfg["spp_grad"] = spp or 16
            cfg["regular_sampling"] = False
    elif name == "C4":
        cfg["vial"] = {"type": "square", "w_int": 10.191, "w_ext": 12.408, "ior": 1.54,
                       "medium": dict(MEDIUM_ABSORPTIVE)}
    if spp is not None:
        cfg["spp"] = spp
    if surface_aware:
        # surface-aware discretization (film.py:22, sensor.py:53-110) needs a plain final_sensor
        # for the final render (optimize.py:113-116)
        import copy
        cfg["final_sensor"] = copy.deepcopy(cfg["sensor"])
        cfg["sensor"]["film"]["surface_aware"] = True
    return cfg
