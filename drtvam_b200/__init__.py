"""drtvam_b200 -- B200-native forward projection + adjoint for Dr.TVAM.

Drop-in for the data-parallel hot path of ``drtvam`` (the reference's
``mi.render`` / ``render_backward`` on its ``volume`` integrator): same config
schema, projector / sensor / film / integrator / loss / L-BFGS classes and
registry functions (``src/drtvam/__init__.py:32-51``), backed by hand-written
sm_100a CUDA kernels behind the C-ABI of ``include/drtvam_b200.h``.

There is no CPU fallback: the CUDA library must be present (built in-tree by
``python -m drtvam_b200.build``) for anything that traces rays.
"""
from . import geometry, loss, motion  # noqa: F401
from ._lib import DrtvamError, load as load_library  # noqa: F401
from .scene import Scene, load_dict, render, traverse  # noqa: F401

__version__ = "0.1.0"


def register_geometry(name, cls):
    if name in geometry.geometries:
        raise ValueError(f"Geometry '{name}' is already registered.")
    if not issubclass(cls, geometry.Container):
        raise ValueError(f"Class '{cls}' is not a subclass of 'geometry.Container'.")
    geometry.geometries[name] = cls


def register_motion(name, cls):
    if name in motion.motions:
        raise ValueError(f"Motion '{name}' is already registered.")
    if not issubclass(cls, motion.Motion):
        raise ValueError(f"Class '{cls}' is not a subclass of 'motion.Motion'.")
    motion.motions[name] = cls


def register_loss(name, cls):
    if name in loss.losses:
        raise ValueError(f"Loss '{name}' is already registered.")
    if not issubclass(cls, loss.Loss):
        raise ValueError(f"Class '{cls}' is not a subclass of 'loss.Loss'.")
    loss.losses[name] = cls
