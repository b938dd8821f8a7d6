"""Debug: surface-aware planar path against the general tracer, per channel."""
import os, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from helpers import build, pixels_np, random_inputs, rel_l2
from test_surface_aware import _handle
from drtvam_b200 import _lib

for name in ("C2", "C4"):
    scene, integ, sensor, d, _ = build(name, tempfile.mkdtemp(), 40, surface_aware=True)
    pix = pixels_np(scene)
    p, g = random_inputs(d, pix.size, 5)
    h = _handle(d)
    print(name, "plan", h.lib.drtvam_scene_plan(h.h, pix.size, 1))
    f = h.forward(p, pix)
    a = h.adjoint(g, pix)
    _lib.check(h.lib.drtvam_scene_force_general(h.h, 1))
    f0 = h.forward(p, pix)
    a0 = h.adjoint(g, pix)
    for c in (0, 1):
        print(" ch", c, "sum planar", f[..., c].sum(), "general", f0[..., c].sum(), "rel", rel_l2(f[..., c], f0[..., c]))
    print(" total rel", rel_l2(f.sum(-1), f0.sum(-1)), "adjoint rel", rel_l2(a, a0))
    dd = np.abs(f - f0)
    i = np.unravel_index(dd.argmax(), dd.shape)
    print(" worst", i, f[i], f0[i])

# which lanes are wrong: one angle at a time, then one DMD row/column at a time (C2)
scene, integ, sensor, d, _ = build("C2", tempfile.mkdtemp(), 40, surface_aware=True)
pix = pixels_np(scene)
p, g = random_inputs(d, pix.size, 5)
h = _handle(d)
hg = _handle(d)
_lib.check(hg.lib.drtvam_scene_force_general(hg.h, 1))
hw = d.res_x * d.res_y
errs = []
for k in range(d.n_patterns):
    q = np.zeros_like(p); q[k * hw:(k + 1) * hw] = p[k * hw:(k + 1) * hw]
    errs.append(rel_l2(h.forward(q, pix)[..., 1], hg.forward(q, pix)[..., 1]))
print("per-angle ch1 err", np.round(errs, 4))
k = int(np.argmax(errs))
q = np.zeros_like(p); q[k * hw:(k + 1) * hw] = 1.0
f, f0 = h.forward(q, pix), hg.forward(q, pix)
print("angle", k, "err", rel_l2(f[..., 1], f0[..., 1]))
z = d.res_y // 2 + 3
np.set_printoptions(linewidth=250, precision=2, suppress=True)
print("layer", z, "planar-SA ch1 * 1e3\n", (f[z, :, :, 1] * 1e3)[::2, ::2])
print("general ch1 * 1e3\n", (f0[z, :, :, 1] * 1e3)[::2, ::2])
