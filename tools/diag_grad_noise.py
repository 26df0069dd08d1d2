#!/usr/bin/env python
"""Gradient error study on one case: reference-vs-reference run-to-run noise (float atomics), ours vs reference, and
both against the mean of several reference runs. usage: tools/diag_grad_noise.py [kind P W H]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "3dgs-mcmc_b200"), ROOT]
import torch
import helpers
from diff_gaussian_rasterization import _native as C
from oracle import ref_cuda

kind, P, W, H = (sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else ("garden", 200000, 640, 360)
dev = torch.device("cuda:0")
scene, cam = helpers.make_case(kind, P, W, H, seed=0, cam_radius=6.0 if kind == "garden" else 4.03, fovx=1.2 if kind == "garden" else 0.6911)
s, c = helpers.torch_inputs(scene, cam, dev)
args = helpers.native_forward_args(s, c, 3)
g = torch.Generator(device="cpu").manual_seed(1)
dL = torch.randn((3, H, W), generator=g).to(dev)
e = torch.empty(0)
names = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")

def ours():
    R, color, radii, geom, binning, img, used = C.rasterize_gaussians(*args)
    out = C.rasterize_gaussians_backward(c["bg"], s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"],
                                         c["tanfovx"], c["tanfovy"], dL, s["shs"], 3, c["campos"], geom, R, binning, img, False)
    torch.cuda.synchronize()
    return dict(zip(names, [t.double() for t in out]))

def ref():
    Rr, color_r, radii_r, used_r = ref_cuda.forward(*args)
    out = ref_cuda.backward(c["bg"], s["means3D"], radii_r, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"],
                            c["tanfovx"], c["tanfovy"], dL, s["shs"], 3, c["campos"], Rr)
    torch.cuda.synchronize()
    return dict(zip(names, [out[n].double() for n in names]))

refs = [ref() for _ in range(6)]
mean = {n: sum(r[n] for r in refs) / len(refs) for n in names}
o1, o2 = ours(), ours()
rl = helpers.rel_l2
print(f"{'grad':10s} {'ref-ref':>10s} {'ours-ref':>10s} {'ours-ours':>10s} {'ref-mean':>10s} {'ours-mean':>10s}")
for n in names:
    print(f"{n:10s} {rl(refs[0][n], refs[1][n]):10.2e} {rl(o1[n], refs[0][n]):10.2e} {rl(o1[n], o2[n]):10.2e} {rl(refs[0][n], mean[n]):10.2e} {rl(o1[n], mean[n]):10.2e}")
