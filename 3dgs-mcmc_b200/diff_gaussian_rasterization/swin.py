"""Fused sliding-window gather, backed by csrc/swin.cu.

``gather_active(frame, immature, matured, imm_idx, mat_idx, deform=True)`` returns the dict
``{"xyz", "rotation", "feature", "opacity", "scaling"}`` that ``SwinGaussianModel.get_basic_para_at(frame)``
(scene/gaussian_model.py:598-649) builds with 12 index ops, 11 ``torch.cat``, ``rigid_deform``
(utils/tempo_utils.py:49-84) and the activations — here one forward and one backward kernel. ``immature`` / ``matured``
are dicts with the model's tensors under the keys of ``SET_KEYS`` (the model's attribute names without the leading
underscore / ``_matured_`` prefix); ``imm_idx`` / ``mat_idx`` are what ``derive_idx_of_active(frame)`` returns.
Gradients flow to the immature set's nine parameter tensors; the matured set is detached, as in the reference.
"""
import torch

from . import _native as _C

SET_KEYS = _C.SWIN_SET_KEYS
__all__ = ["gather_active", "SET_KEYS"]


class _SwinGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, frame, deform, imm_idx, mat_idx, mat_set, *imm_set):
        outs = _C.swin_gather_forward(frame, deform, imm_idx, mat_idx, imm_set, mat_set)
        ctx.frame, ctx.deform = float(frame), bool(deform)
        ctx.save_for_backward(imm_idx, *imm_set)
        return outs

    @staticmethod
    def backward(ctx, *out_grads):
        imm_idx, *imm_set = ctx.saved_tensors
        Ki = int(imm_idx.numel())
        # only the first Ki rows of the outputs come from the immature set (kernel reads rows [0, Ki) of each gradient)
        d = _C.swin_gather_backward(ctx.frame, ctx.deform, imm_idx, imm_set, out_grads)
        del Ki
        return (None, None, None, None, None, *d, None)


def gather_active(frame, immature, matured, imm_idx, mat_idx, deform=True):
    imm_set = tuple(immature[k] for k in SET_KEYS)
    mat_set = tuple(matured[k] for k in SET_KEYS) if matured is not None and int(mat_idx.numel()) else None
    xyz, rotation, feature, opacity, scaling = _SwinGather.apply(frame, deform, imm_idx, mat_idx, mat_set, *imm_set)
    return {"xyz": xyz, "rotation": rotation, "feature": feature, "opacity": opacity, "scaling": scaling}
