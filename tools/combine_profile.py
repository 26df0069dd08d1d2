#!/usr/bin/env python
"""Single-GPU stand-in for the V-view colour-gradient exchange so that its kernels can be timed and captured by ncu without
peers: V camera views of the bench scene are rendered forward + backward on ONE GPU, each view's masked dL/dRGB is pushed
into its slot of a local gather buffer (gs_dp_color_grad_push, dense or sparse layout) and the summed SH gradient is rebuilt
(gs_dp_sh_grad_combine). Prints the times of both kernels, the non-zero fraction, and whether the two layouts agree bit for
bit.   usage: combine_profile.py [V=8] [P=3000000] [W=1920] [H=1080]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "3dgs-mcmc_b200"))
import torch  # noqa: E402
import gs_synth  # noqa: E402
from diff_gaussian_rasterization import _native as C  # noqa: E402

dev = torch.device("cuda:0")
argv = [int(a) for a in sys.argv[1:]] + [None] * 4
V, P, W, H = argv[0] or 8, argv[1] or 3_000_000, argv[2] or 1920, argv[3] or 1080
s = gs_synth.to_torch(gs_synth.scene_garden(P, 0), dev)
cams = [gs_synth.to_torch(c, dev) for c in gs_synth.orbit_cameras(16, 6.0, W, H, 1.2, 20.0)][:V]
bg, e = torch.zeros(3, device=dev), torch.empty(0)
campos = torch.stack([c["campos"] for c in cams])
gen = torch.Generator(device="cpu").manual_seed(5)
out = {"V": V, "P": P}
results = {}


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for layout in ("dense", "sparse"):
    stride = (C.dp_slot_floats(P, layout) + 63) // 64 * 64
    gathered = torch.zeros((V, stride), device=dev)
    nz = []
    for v, c in enumerate(cams):
        R, color, radii, geom, binning, img, used = C.rasterize_gaussians(
            bg, s["means3D"], e, s["opacities"], s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"],
            c["tanfovx"], c["tanfovy"], H, W, s["shs"], 3, c["campos"], False, False)
        dL = torch.randn((3, H, W), generator=gen).to(dev)
        scratch = C.backward_render(bg, P, R, dL, geom, binning, img)
        C.dp_color_grad_push(P, radii, geom, scratch, [gathered.data_ptr()], v * stride, False, layout=layout)
        if v == V - 1:
            out[f"push_ms_{layout}"] = timed(lambda: C.dp_color_grad_push(P, radii, geom, scratch, [gathered.data_ptr()], v * stride, False, layout=layout))
        if layout == "dense":
            nz.append(float((gathered[v, : 3 * P].view(P, 3) != 0).any(1).float().mean().item()))
            out.setdefault("visible_fraction", []).append(round(float((radii > 0).float().mean().item()), 4))
    if nz:
        out["nonzero_fraction_per_view"] = [round(x, 4) for x in nz]
    sh = torch.empty((P, 16, 3), device=dev)
    out[f"combine_ms_{layout}"] = timed(lambda: C.dp_sh_grad_combine(s["means3D"], campos, gathered, stride, 3, 16, out=sh, layout=layout))
    results[layout] = sh
gen2 = None
out["layouts_bit_identical"] = bool(torch.equal(results["dense"].view(torch.int32), results["sparse"].view(torch.int32)))
print(json.dumps(out))
