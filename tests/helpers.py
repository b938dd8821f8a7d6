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


def flip_report(a, b, thresh=1e-3):
    """(flagged, rel_l2_rest, rel_l2_all) of a JITTERED comparison.

    A jittered ray that starts within an ulp of a voxel boundary is a knife edge of fp32 rounding:
    which voxel it is truncated into, and whether the reference's `dtmax < 0 -> inf` rule
    (sensor.py:359) freezes an axis, can differ between two compilations of the same arithmetic (CUDA
    kernels with different inlining contexts, gcc for the oracle).  Such a FLIPPED ray moves its whole
    deposit (its whole gradient entry) while every other ray agrees to ~1e-6.  Entries whose
    difference exceeds `thresh` x rms(b) are flagged - for a pattern gradient one entry is one ray, for
    a dose volume one flipped ray flags at most the voxels of its two chords - and the caller asserts
    their COUNT, the bound on everything else, and the plain norm."""
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    diff = np.abs(a - b)
    rms = float(np.sqrt(np.mean(b * b)))
    flagged = diff > thresh * max(rms, 1e-300)
    rest = ~flagged
    rel_rest = float(np.linalg.norm(diff[rest]) / max(np.linalg.norm(b[rest]), 1e-300))
    rel_all = float(np.linalg.norm(diff) / max(np.linalg.norm(b), 1e-300))
    return int(flagged.sum()), rel_rest, rel_all


def assert_jitter_parity(a, b, max_flipped_rays, voxels_per_ray=1, tol=1e-4, band=3e-2, what=""):
    """Asserts a jittered comparison: at most `max_flipped_rays` knife-edge rays (each flags up to
    `voxels_per_ray` entries), everything else inside `tol`, the plain relative L2 inside `band`."""
    flagged, rel_rest, rel_all = flip_report(a, b)
    msg = f"{what}: flagged {flagged} entries (allowed {max_flipped_rays} rays x {voxels_per_ray}), rest {rel_rest:.2e}, all {rel_all:.2e}"
    assert flagged <= max_flipped_rays * voxels_per_ray, msg
    assert rel_rest < tol, msg
    assert rel_all < band, msg
    return flagged


def record(test, **numbers):
    """Appends the measured parity numbers of a GPU test to gpurun_out/parity_numbers.jsonl (evidence for
    profiles/; a no-op when the directory is not writable)."""
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        os.makedirs(os.path.join(root, "gpurun_out"), exist_ok=True)
        with open(os.path.join(root, "gpurun_out", "parity_numbers.jsonl"), "a") as f:
            f.write(json.dumps({"test": test, **{k: (float(v) if isinstance(v, (int, float, np.floating, np.integer)) else v)
                                                  for k, v in numbers.items()}}) + "\n")
    except OSError:
        pass
