"""Shared test helpers: synthetic scenes lowered to the flat description both
the CUDA library and the CPU oracle consume."""
import numpy as np
import torch

from drtvam_b200.integrators import VolumeIntegrator
from drtvam_b200.optimize import load_scene
from drtvam_b200.scene import load_dict
from drtvam_b200.synthetic import make_config


def build(name, workdir, n, device="cpu", integrator=None, **kw):
    """(scene, integrator, sensor, SceneDesc) for synthetic config `name` at edge `n`."""
    cfg = make_config(name, workdir, n=n, **kw)
    scene = load_dict(load_scene(cfg), device=device)
    props = {"max_depth": cfg["max_depth"], "rr_depth": cfg["rr_depth"],
             "regular_sampling": cfg["regular_sampling"], "print_time": 1.0}
    props.update(integrator or {})
    integ = VolumeIntegrator(props)
    sensor = scene.sensor("sensor")
    return scene, integ, sensor, scene.describe(integ, sensor), cfg


def rel_l2(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def random_inputs(desc, n_active, seed=0):
    rng = np.random.default_rng(seed)
    p = rng.random(n_active).astype(np.float32)
    g = rng.standard_normal(desc.film_shape()).astype(np.float32)
    return p, g


def pixels_np(scene):
    return (scene.projector.active_pixels.cpu().numpy().astype(np.int64) & 0xFFFFFFFF).astype(np.uint32)
