import sys, os, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
from helpers import build, pixels_np, random_inputs, rel_l2
from test_gpu_parity import Handle
name, n, k, row, col = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
work = tempfile.mkdtemp()
scene, integ, sensor, d, _ = build(name, work, n, regular_sampling=False)
pix = pixels_np(scene)
p = np.zeros(pix.size, np.float32)
j = (k * n + row) * n + col
p[j] = 1.0
fast, slow = Handle(d), Handle(d, force_general=True)
f, s = fast.forward(p, pix, 1, 0)[..., 0], slow.forward(p, pix, 1, 0)[..., 0]
rays = slow.raygen(pix, 1, 0)
print("ray", rays[j])
for nm, v in (("planar", f), ("general", s)):
    zz, yy, xx = np.nonzero(v)
    print(nm, "nnz", zz.size, "sum", v.sum())
    for z, y, x in list(zip(zz, yy, xx))[:60]:
        print("   ", z, y, x, v[z, y, x])
