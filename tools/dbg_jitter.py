import sys, os, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
from helpers import build, pixels_np, random_inputs, rel_l2
from test_gpu_parity import Handle
name = sys.argv[1] if len(sys.argv) > 1 else "C1"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
work = tempfile.mkdtemp()
npat = int(sys.argv[3]) if len(sys.argv) > 3 else None
scene, integ, sensor, d, _ = build(name, work, n, regular_sampling=False, n_patterns=npat)
pix = pixels_np(scene)
p, g = random_inputs(d, pix.size, 17)
p = (p - 0.3).astype(np.float32)
fast, slow = Handle(d), Handle(d, force_general=True)
for spp, seed in ((1, 0), (4, 11), (1, 3), (2, 8), (3, 1)):
    f, s = fast.forward(p, pix, spp, seed), slow.forward(p, pix, spp, seed)
    print("fwd", spp, seed, rel_l2(f, s), "sum", f.sum(), s.sum())
    diff = np.abs(f - s)[..., 0]
    zz, yy, xx = np.unravel_index(np.argsort(diff.ravel())[-8:], diff.shape)
    for z, y, x in zip(zz, yy, xx):
        print("   voxel z,y,x", z, y, x, f[z, y, x, 0], s[z, y, x, 0])
    perlayer = np.abs(f - s)[..., 0].sum(axis=(1, 2))
    print("   layers with diff:", np.nonzero(perlayer > 1e-6 * np.abs(s).sum() / s.shape[0])[0][:20])
    a, b = fast.adjoint(g, pix, spp, seed), slow.adjoint(g, pix, spp, seed)
    print("adj", rel_l2(a, b))
    da = np.abs(a - b)
    bad = np.argsort(da)[-6:]
    hw = d.res_x * d.res_y
    for j in bad:
        print("   pixel k,row,col", j // hw, (j % hw) // d.res_x, j % d.res_x, a[j], b[j])
