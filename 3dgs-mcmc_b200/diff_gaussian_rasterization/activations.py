"""Fused parameter activations + regularisers, backed by csrc/activations.cu.

``activate(raw_opacity, raw_scaling, raw_rotation, opacity_reg=0, scale_reg=0)`` returns
``(opacity, scaling, rotation, reg)`` with

* ``opacity = torch.sigmoid(raw_opacity)``, ``scaling = torch.exp(raw_scaling)``,
  ``rotation = torch.nn.functional.normalize(raw_rotation)`` — the reference's ``get_opacity / get_scaling /
  get_rotation`` (scene/gaussian_model_static.py:43-56, :126-146), and
* ``reg = opacity_reg * |opacity|.mean() + scale_reg * |scaling|.mean()`` — the two regulariser terms the MCMC trainer
  adds to the loss (train_static.py:99-100), for which the reference evaluates the activations a second time.

One forward kernel and one backward kernel replace ~25 elementwise / reduction launches per iteration. No CPU path.
"""
import torch

from . import _native as _C

__all__ = ["activate"]


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw_opacity, raw_scaling, raw_rotation, opacity_reg, scale_reg):
        opacity, scaling, rotation, means = _C.activate_forward(raw_opacity, raw_scaling, raw_rotation)
        ctx.save_for_backward(raw_opacity, raw_scaling, raw_rotation)
        ctx.weights = (float(opacity_reg), float(scale_reg))
        reg = opacity_reg * means[0] + scale_reg * means[1]
        return opacity, scaling, rotation, reg

    @staticmethod
    def backward(ctx, g_opacity, g_scaling, g_rotation, g_reg):
        raw_opacity, raw_scaling, raw_rotation = ctx.saved_tensors
        w_o, w_s = ctx.weights
        P = max(int(raw_opacity.size(0)), 1)
        d_o, d_s, d_q = _C.activate_backward(raw_opacity, raw_scaling, raw_rotation, g_opacity, g_scaling, g_rotation,
                                             w_o / P, w_s / (3.0 * P), g_reg)
        return d_o, d_s, d_q, None, None


def activate(raw_opacity, raw_scaling, raw_rotation, opacity_reg=0.0, scale_reg=0.0):
    return _Activate.apply(raw_opacity, raw_scaling, raw_rotation, opacity_reg, scale_reg)
