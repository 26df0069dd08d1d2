#!/usr/bin/env python
"""Golden vectors for the fused sliding-window gather, generated with the REFERENCE's own utils/tempo_utils.py
(rigid_deform, rotvec2quat, rotvec2mat, quat_mul: :4-84) imported from /root/reference in this container (CPU), composed
exactly as SwinGaussianModel.get_basic_para_at does (scene/gaussian_model.py:598-649: index both sets, cat, age-based
rigid_deform, F.normalize / sigmoid / exp — the model class itself hard-codes .cuda() and cannot be instantiated here).
Gradients by torch autograd for a fixed random upstream gradient. Run once:

    python tests/golden/make_golden_swin.py        # writes tests/golden/swin_gather.npz
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.tempo_utils import rigid_deform  # noqa: E402

KEYS = ("xyz", "rotation", "features_dc", "features_rest", "opacity", "scaling", "rigid_v", "rigid_rotvec", "rigid_rotcen",
        "frame_start")


def make_set(N, g, M=16):
    s = {"xyz": torch.randn((N, 3), generator=g), "rotation": torch.randn((N, 4), generator=g),
         "features_dc": torch.randn((N, 1, 3), generator=g), "features_rest": torch.randn((N, M - 1, 3), generator=g) * 0.1,
         "opacity": torch.randn((N, 1), generator=g), "scaling": torch.randn((N, 3), generator=g) - 3,
         "rigid_v": torch.randn((N, 3), generator=g) * 0.05, "rigid_rotvec": torch.randn((N, 3), generator=g) * 0.2,
         "rigid_rotcen": torch.randn((N, 3), generator=g), "frame_start": torch.randint(0, 10, (N,), generator=g).float()}
    s["rigid_rotvec"][::7] = 0.0          # no rotation: angle 0, F.normalize of the zero vector
    return s


def reference_gather(frame, imm, mat, ii, mi, deform):
    """scene/gaussian_model.py:598-649 with the model's tensors passed in."""
    age = torch.cat((frame - imm["frame_start"][ii], frame - mat["frame_start"][mi]), dim=0)
    cat = lambda k: torch.cat((imm[k][ii], mat[k][mi]), dim=0)  # noqa: E731
    pos, rot = rigid_deform(cat("xyz"), cat("rotation"), cat("rigid_v"), cat("rigid_rotvec"), cat("rigid_rotcen"), age,
                            skip=not deform)
    feature = torch.cat((torch.cat((imm["features_dc"][ii], imm["features_rest"][ii]), dim=1),
                         torch.cat((mat["features_dc"][mi], mat["features_rest"][mi]), dim=1)), dim=0)
    return {"xyz": pos, "rotation": torch.nn.functional.normalize(rot), "feature": feature,
            "opacity": torch.sigmoid(cat("opacity")), "scaling": torch.exp(cat("scaling"))}


def main():
    g = torch.Generator().manual_seed(21)
    out = {}
    for name, (Ni, Nm, frame, deform) in {"deform": (300, 200, 6.0, True), "skip": (120, 80, 4.0, False),
                                          "immature_only": (150, 1, 9.0, True)}.items():
        imm, mat = make_set(Ni, g), make_set(Nm, g)
        imm["frame_start"][:5] = frame        # age 0: identity motion
        ii = torch.nonzero(imm["frame_start"] <= frame, as_tuple=True)[0]
        mi = torch.nonzero(mat["frame_start"] <= frame, as_tuple=True)[0]
        if name == "immature_only":
            mi = mi[:0]
        for k in KEYS[:9]:
            imm[k].requires_grad_(True)
        res = reference_gather(frame, imm, mat, ii, mi, deform)
        ups = {k: torch.randn(v.shape, generator=g) for k, v in res.items()}
        total = sum((res[k] * ups[k]).sum() for k in res)
        grads = torch.autograd.grad(total, [imm[k] for k in KEYS[:9]], allow_unused=True)
        out[f"{name}/frame"] = np.float32(frame)
        out[f"{name}/deform"] = np.int32(deform)
        out[f"{name}/imm_idx"] = ii.numpy()
        out[f"{name}/mat_idx"] = mi.numpy()
        for k in KEYS:
            out[f"{name}/imm/{k}"] = imm[k].detach().numpy()
            out[f"{name}/mat/{k}"] = mat[k].detach().numpy()
        for k in res:
            out[f"{name}/out/{k}"] = res[k].detach().numpy()
            out[f"{name}/up/{k}"] = ups[k].numpy()
        for k, gr in zip(KEYS[:9], grads):
            out[f"{name}/grad/{k}"] = (torch.zeros_like(imm[k]) if gr is None else gr).numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "swin_gather.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
