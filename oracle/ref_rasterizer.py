"""`GaussianRasterizer` bound to the reference's OWN CUDA rasterizer (oracle/_ref/libref_dgr.so via oracle.ref_cuda):
the same autograd contract as diff_gaussian_rasterization/__init__.py of the reference (:21-155), so that the
reference's Python callers can be run once on this package and once on the golden implementation. Test infrastructure.

The shim keeps ONE set of state buffers (the last forward's), so a backward must follow its own forward before the next
forward — which is how the callers under test use it.
"""
import contextlib

import torch
import torch.nn as nn

from oracle import ref_cuda


class _RefRasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, s):
        c = lambda t: t.contiguous() if t.numel() else t  # noqa: E731
        args = [c(t.detach()) for t in (means3D, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, sh)]
        means3D_, colors_, opac_, scales_, rots_, cov_, sh_ = args
        torch.cuda.synchronize()     # the reference library runs on the legacy default stream
        R, color, radii, is_used = ref_cuda.forward(s.bg, means3D_, colors_, opac_, scales_, rots_, s.scale_modifier, cov_,
                                                    s.viewmatrix.contiguous(), s.projmatrix.contiguous(), s.tanfovx, s.tanfovy,
                                                    s.image_height, s.image_width, sh_, s.sh_degree, s.campos.contiguous())
        torch.cuda.synchronize()
        ctx.s, ctx.R = s, R
        ctx.save_for_backward(means3D_, colors_, scales_, rots_, cov_, sh_, radii)
        ctx.mark_non_differentiable(radii, is_used)
        return color, radii, is_used

    @staticmethod
    def backward(ctx, grad_color, _gr, _gu):
        s = ctx.s
        means3D, colors, scales, rots, cov, sh, radii = ctx.saved_tensors
        torch.cuda.synchronize()
        g = ref_cuda.backward(s.bg, means3D, radii, colors, scales, rots, s.scale_modifier, cov, s.viewmatrix.contiguous(),
                              s.projmatrix.contiguous(), s.tanfovx, s.tanfovy, grad_color.contiguous(), sh, s.sh_degree,
                              s.campos.contiguous(), ctx.R)
        torch.cuda.synchronize()
        none_if_absent = lambda t, like: t if like.numel() else None  # noqa: E731
        return (g["means3D"], g["means2D"], none_if_absent(g["sh"], sh), none_if_absent(g["colors"], colors), g["opacity"],
                none_if_absent(g["scales"], scales), none_if_absent(g["rotations"], rots), none_if_absent(g["cov3D"], cov), None)


class RefGaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None):
        absent = torch.Tensor([])
        shs = absent if shs is None else shs
        colors_precomp = absent if colors_precomp is None else colors_precomp
        scales = absent if scales is None else scales
        rotations = absent if rotations is None else rotations
        cov3D_precomp = absent if cov3D_precomp is None else cov3D_precomp
        return _RefRasterize.apply(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   self.raster_settings)


@contextlib.contextmanager
def bound_to(module):
    """Inside the block, `module.GaussianRasterizer` (e.g. the reference's gaussian_renderer) is the reference CUDA one."""
    saved = module.GaussianRasterizer
    module.GaussianRasterizer = RefGaussianRasterizer
    try:
        yield
    finally:
        module.GaussianRasterizer = saved
