#!/usr/bin/env python
"""Single-GPU stand-in for the 8-view SH-gradient rebuild (gs_dp_sh_grad_combine) so that ncu can capture it."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "3dgs-mcmc_b200"))
import torch  # noqa: E402
import gs_synth  # noqa: E402
from diff_gaussian_rasterization import _native as C  # noqa: E402

dev = torch.device("cuda:0")
P, V = 3_000_000, int(sys.argv[1]) if len(sys.argv) > 1 else 8
means = torch.from_numpy(gs_synth.scene_garden(P, 0)["means3D"]).to(dev)
stride = (3 * P + 63) // 64 * 64
g = torch.randn((V, stride), device=dev)
g[:, : 3 * P].view(V, P, 3)[torch.rand((V, P), device=dev) < 0.25] = 0.0     # ~25 % culled per view
campos = torch.randn((V, 3), device=dev) * 6
out = torch.empty((P, 16, 3), device=dev)
for _ in range(3):
    C.dp_sh_grad_combine(means, campos, g, stride, 3, 16, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    C.dp_sh_grad_combine(means, campos, g, stride, 3, 16, out=out)
e1.record()
torch.cuda.synchronize()
print("combine ms", e0.elapsed_time(e1) / 10, "V", V)
