"""Fused L1 + SSIM training loss, backed by csrc/loss.cu.

Mirrors the reference's utils/loss_utils.py for the functions the trainers use (train_static.py:94-97):

* ``ssim(img1, img2, window_size=11, size_average=True)``  (loss_utils.py:40-67) — same signature and value; the
  gradient flows to ``img1`` only (the ground truth needs none).
* ``l1_loss(network_output, gt)``                            (loss_utils.py:21-22) — ``gs_l1_loss_forward / _backward``.
* ``photometric_loss(image, gt, lambda_dssim)`` = ``(1 - lambda) * l1_loss + lambda * (1 - ssim)`` — the composition
  train_static.py:97 writes out, as ONE forward kernel (+ a 1-block reduction) and ONE backward kernel instead of
  five ``conv2d`` calls and ~20 elementwise kernels each way. There is no CPU path.
"""
import torch

from . import _native as _C

__all__ = ["ssim", "l1_loss", "photometric_loss"]


class _PhotometricLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, w_l1, w_ssim_term, bias):
        # value = w_l1 * mean|image - gt| + w_ssim_term * mean(ssim_map) + bias
        means, ws = _C.photometric_loss_forward(image, gt)
        ctx.save_for_backward(image, gt, ws)
        ctx.weights = (w_l1, w_ssim_term)
        return w_l1 * means[0] + w_ssim_term * means[1] + bias

    @staticmethod
    def backward(ctx, grad_out):
        image, gt, ws = ctx.saved_tensors
        w_l1, w_ssim_term = ctx.weights
        n = float(image.numel())
        grad = _C.photometric_loss_backward(image, gt, ws, w_l1 / n, w_ssim_term / n, grad_scale=grad_out)
        return grad, None, None, None, None


def photometric_loss(image, gt, lambda_dssim=0.2):
    """(1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt)) (train_static.py:95-97)."""
    lam = float(lambda_dssim)
    return _PhotometricLoss.apply(image, gt, 1.0 - lam, -lam, lam)


def ssim(img1, img2, window_size=11, size_average=True):
    if window_size != 11 or not size_average:
        raise NotImplementedError("the fused kernel implements the trainers' call: window_size=11, size_average=True")
    return _PhotometricLoss.apply(img1, img2, 0.0, 1.0, 0.0)


class _L1Loss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, network_output, gt):
        ctx.save_for_backward(network_output, gt)
        return _C.l1_loss_forward(network_output, gt)[0]

    @staticmethod
    def backward(ctx, grad_out):
        network_output, gt = ctx.saved_tensors
        grad = _C.l1_loss_backward(network_output, gt, grad_scale=grad_out)
        # the ground truth needs no gradient in the trainers; if asked for, it is the negated map
        return grad, (-grad if ctx.needs_input_grad[1] else None)


def l1_loss(network_output, gt):
    """torch.abs(network_output - gt).mean() (utils/loss_utils.py:21-22): one streaming kernel + a fixed-order reduction
    forward, one streaming kernel backward."""
    return _L1Loss.apply(network_output, gt)
