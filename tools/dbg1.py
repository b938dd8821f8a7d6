import sys, tempfile
sys.path[:0]=['/root/repo','/root/repo/oracle','/root/repo/tests']
import numpy as np, torch
import oracle as orc
from helpers import build, pixels_np, random_inputs, rel_l2
from test_gpu_parity import Handle, _plan
w=tempfile.mkdtemp()
for name in ["C1","C2","C3","C4"]:
    for n,npat in ((160,24),(48,48)):
        scene, integ, sensor, d, _ = build(name, w, n, n_patterns=npat)
        pix=pixels_np(scene); p,g=random_inputs(d,pix.size,11); p=(p-0.4).astype(np.float32)
        fast,slow=Handle(d),Handle(d,force_general=True)
        ff,fs=fast.forward(p,pix),slow.forward(p,pix)
        af,as_=fast.adjoint(g,pix),slow.adjoint(g,pix)
        o=orc.Oracle(d.as_dict(),"f64")
        fo=o.forward(p,pix); ao=o.adjoint(g,pix)
        print(name,n,"plan",_plan(fast,pix.size),"fwd fast/slow",rel_l2(ff,fs),"fast/orc",rel_l2(ff,fo),"slow/orc",rel_l2(fs,fo),
              "adj fast/slow",rel_l2(af,as_),"fast/orc",rel_l2(af,ao),"slow/orc",rel_l2(as_,ao), flush=True)
        if rel_l2(ff,fs)>2e-5:
            dlt=np.abs(ff-fs)[...,0]; idx=np.unravel_index(np.argsort(dlt.ravel())[-5:],dlt.shape)
            print("  worst voxels (z,y,x):",list(zip(*idx)),dlt[idx], ff[...,0][idx], fs[...,0][idx])
            print("  sum fast",ff.sum(),"slow",fs.sum(),"orc",fo.sum())
scene, integ, sensor, d, _ = build("C5", w, 32, spp=4)
pix=pixels_np(scene); p,g=random_inputs(d,pix.size,5)
o=orc.Oracle(d.as_dict(),"f64"); h=Handle(d)
f_gpu,f_cpu=h.forward(p,pix,4,2),o.forward(p,pix,4,2)
print("C5 fwd",rel_l2(f_gpu,f_cpu),f_gpu.sum(),f_cpu.sum())
print("C5 adj",rel_l2(h.adjoint(g,pix,4,2),o.adjoint(g,pix,4,2)))
o32=orc.Oracle(d.as_dict(),"f32")
print("C5 f32 vs f64 oracle fwd",rel_l2(o32.forward(p,pix,4,2),f_cpu),"adj",rel_l2(o32.adjoint(g,pix,4,2),o.adjoint(g,pix,4,2)))
