import sys, os, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
import oracle as orc
from helpers import build, pixels_np, random_inputs, rel_l2
from test_gpu_parity import Handle
work = tempfile.mkdtemp()
scene, integ, sensor, d, _ = build("C2", work, 144, regular_sampling=False, n_patterns=6)
pix = pixels_np(scene)
p, g = random_inputs(d, pix.size, 17)
p = (p - 0.3).astype(np.float32)
fast, slow = Handle(d), Handle(d, force_general=True)
o32 = orc.Oracle(d.as_dict(), "f32")
spp, seed = 3, 1
ff, fs, fo = fast.forward(p, pix, spp, seed), slow.forward(p, pix, spp, seed), o32.forward(p, pix, spp, seed)
af, as_, ao = fast.adjoint(g, pix, spp, seed), slow.adjoint(g, pix, spp, seed), o32.adjoint(g, pix, spp, seed)
print("fwd fast-slow", rel_l2(ff, fs), "fast-orc", rel_l2(ff, fo), "slow-orc", rel_l2(fs, fo))
print("adj fast-slow", rel_l2(af, as_), "fast-orc", rel_l2(af, ao), "slow-orc", rel_l2(as_, ao))
for nm, f, a in (("fast", ff, af), ("slow", fs, as_), ("orc", fo, ao)):
    print(nm, "dot", float((f.astype(np.float64) * g).sum()), float((a.astype(np.float64) * p).sum()))
# which pixel feeds voxel (48,4,74) differently: one-hot gradient
g1 = np.zeros_like(g); g1[48, 0:5, 74, 0] = 1.0; g1[48, 4, 71, 0] = -1.0; g1[48, 2, 72, 0] = -1; g1[48, 0, 73, 0] = -1
a1f, a1s = fast.adjoint(g1, pix, spp, seed), slow.adjoint(g1, pix, spp, seed)
hw = d.res_x * d.res_y
for j in np.argsort(np.abs(a1f - a1s))[-5:]:
    print("pixel", j, j // hw, (j % hw) // d.res_x, j % d.res_x, a1f[j], a1s[j], "p", p[j])
# run the forward twice: is the general forward deterministic?
fs2 = slow.forward(p, pix, spp, seed); ff2 = fast.forward(p, pix, spp, seed)
print("repeat slow", rel_l2(fs2, fs), "repeat fast", rel_l2(ff2, ff))
