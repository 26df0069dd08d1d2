#!/usr/bin/env python
"""Runs the SURVEY.md §8 f1 kernels (fused Adam, fused L1 + SSIM loss) a few times at the bench sizes so that ncu can
capture them:  ncu --set full -k regex:'adam|ssim|loss' python tools/next_rows_profile.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "3dgs-mcmc_b200"))

import torch  # noqa: E402

from diff_gaussian_rasterization import loss as fused_loss  # noqa: E402
from diff_gaussian_rasterization.optim import FusedAdam  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    P, H, W = 3_000_000, 1080, 1920
    shapes = [(P, 3), (P, 16, 3), (P, 1), (P, 3), (P, 4)]
    ps = [torch.nn.Parameter(torch.randn(s, device=dev)) for s in shapes]
    opt = FusedAdam([{"params": [p], "lr": 1e-3} for p in ps], lr=0.0, eps=1e-15)
    for p in ps:
        p.grad = torch.randn_like(p)
    gt = torch.rand((3, H, W), device=dev)
    img = (gt + 0.05 * torch.randn_like(gt)).clamp(0, 1)
    for _ in range(3):
        opt.step()
        x = img.detach().requires_grad_(True)
        fused_loss.photometric_loss(x, gt, 0.2).backward()
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
