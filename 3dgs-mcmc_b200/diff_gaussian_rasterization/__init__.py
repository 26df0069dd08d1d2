"""Drop-in replacement for the reference's ``diff_gaussian_rasterization`` package, backed by
hand-written sm_100a CUDA kernels (lib3dgs_mcmc_b200.so, C ABI in include/gs_b200.h).

Public surface and calling conventions mirror
submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py of NeutrinoLiu/3dgs-mcmc:

* ``GaussianRasterizationSettings``  (reference :157-169) — same fields, same order.
* ``GaussianRasterizer``             (reference :171-220) — ``forward(means3D, means2D, opacities, shs,
  colors_precomp, scales, rotations, cov3D_precomp) -> (color[3,H,W], radii[P], is_used[P])`` and
  ``markVisible(positions)``.
* ``rasterize_gaussians`` / ``_RasterizeGaussians`` (reference :21-155) — autograd function with the same
  argument order, saved tensors and gradient tuple.
* ``compute_relocation``             (reference :224-226).
* ``add_noise`` — additive: the fused form of the inline perturbation in train_static.py:132-140.

so ``gaussian_renderer/__init__.py``, ``utils/reloc_utils.py`` and the trainers run unchanged when this
directory's parent (``3dgs-mcmc_b200/``) is on ``sys.path``.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _native as _C
from ._native import add_noise  # noqa: F401  (additive API)

__all__ = [
    "GaussianRasterizationSettings",
    "GaussianRasterizer",
    "rasterize_gaussians",
    "compute_relocation",
    "add_noise",
]


def _snapshot(args):
    """CPU copies of every tensor argument, taken before the call so a crash cannot corrupt them."""
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _call_with_snapshot(fn, args, debug, dump_name, what):
    """debug=True: on failure, save the inputs to <dump_name> and re-raise (reference :83-90, :132-139)."""
    if not debug:
        return fn(*args)
    saved = _snapshot(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print(f"\nAn error occured in {what}. Please forward {dump_name} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        s = raster_settings
        native_args = (
            s.bg, means3D, colors_precomp, opacities, scales, rotations, s.scale_modifier, cov3Ds_precomp,
            s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy, s.image_height, s.image_width, sh, s.sh_degree,
            s.campos, s.prefiltered, s.debug,
        )
        num_rendered, color, radii, geom_buffer, binning_buffer, img_buffer, is_used = _call_with_snapshot(
            _C.rasterize_gaussians, native_args, s.debug, "snapshot_fw.dump", "forward")

        ctx.raster_settings = s
        ctx.num_rendered = num_rendered
        # opacities is not saved: the backward reads it from the packed splat stream, as the reference reads
        # it from conic_opacity.w.
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom_buffer, binning_buffer, img_buffer)
        return color, radii, is_used

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, _grad_is_used):
        s = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
         geom_buffer, binning_buffer, img_buffer) = ctx.saved_tensors
        native_args = (
            s.bg, means3D, radii, colors_precomp, scales, rotations, s.scale_modifier, cov3Ds_precomp,
            s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy, grad_out_color, sh, s.sh_degree, s.campos,
            geom_buffer, ctx.num_rendered, binning_buffer, img_buffer, s.debug,
        )
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = _call_with_snapshot(
            _C.rasterize_gaussians_backward, native_args, s.debug, "snapshot_bw.dump", "backward")
        # One gradient per forward input, None for the settings tuple.
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None)


class _RasterizeGaussiansSplitSH(torch.autograd.Function):
    """Additive: the SH coefficients as the model stores them, features_dc [P,1,3] and features_rest [P,15,3]
    (scene/gaussian_model_static.py:139-144), instead of their concatenation. Same kernels and arithmetic as
    _RasterizeGaussians - bit-identical image and gradients - without the 576 MB copy (P = 3M) of get_features on the way
    in and autograd's split of the gradient on the way out."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh_dc, sh_rest, opacities, scales, rotations, raster_settings):
        s = raster_settings
        absent = torch.Tensor([])
        num_rendered, color, radii, geom_buffer, binning_buffer, img_buffer, is_used = _C.rasterize_gaussians(
            s.bg, means3D, absent, opacities, scales, rotations, s.scale_modifier, absent, s.viewmatrix, s.projmatrix,
            s.tanfovx, s.tanfovy, s.image_height, s.image_width, (sh_dc, sh_rest), s.sh_degree, s.campos, s.prefiltered,
            s.debug)
        ctx.raster_settings = s
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(means3D, scales, rotations, radii, sh_dc, sh_rest, geom_buffer, binning_buffer, img_buffer)
        ctx.mark_non_differentiable(radii, is_used)
        return color, radii, is_used

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, _grad_is_used):
        s = ctx.raster_settings
        means3D, scales, rotations, radii, sh_dc, sh_rest, geom_buffer, binning_buffer, img_buffer = ctx.saved_tensors
        absent = torch.Tensor([])
        (grad_means2D, _grad_colors, grad_opacities, grad_means3D, _grad_cov3D, grad_sh, grad_scales,
         grad_rotations) = _C.rasterize_gaussians_backward(
            s.bg, means3D, radii, absent, scales, rotations, s.scale_modifier, absent, s.viewmatrix, s.projmatrix,
            s.tanfovx, s.tanfovy, grad_out_color, (sh_dc, sh_rest), s.sh_degree, s.campos, geom_buffer, ctx.num_rendered,
            binning_buffer, img_buffer, s.debug)
        return (grad_means3D, grad_means2D, grad_sh[0], grad_sh[1], grad_opacities, grad_scales, grad_rotations, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    if isinstance(sh, (tuple, list)):
        if colors_precomp.numel() or cov3Ds_precomp.numel():
            raise Exception('Split SH tensors (features_dc, features_rest) go with scales / rotations only!')
        return _RasterizeGaussiansSplitSH.apply(means3D, means2D, sh[0], sh[1], opacities, scales, rotations, raster_settings)
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of the points whose view-space depth passes the near cull for this camera."""
        with torch.no_grad():
            s = self.raster_settings
            return _C.mark_visible(positions, s.viewmatrix, s.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_scale_rot = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (has_scale_rot and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        # Absent optionals travel as empty tensors (null pointers at the C ABI), like the reference.
        absent = torch.Tensor([])
        shs = absent if shs is None else shs
        colors_precomp = absent if colors_precomp is None else colors_precomp
        scales = absent if scales is None else scales
        rotations = absent if rotations is None else rotations
        cov3D_precomp = absent if cov3D_precomp is None else cov3D_precomp

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings)


def compute_relocation(opacity_old, scale_old, N, binoms, n_max):
    new_opacity, new_scale = _C.compute_relocation(opacity_old, scale_old, N.int(), binoms, n_max)
    return new_opacity, new_scale
