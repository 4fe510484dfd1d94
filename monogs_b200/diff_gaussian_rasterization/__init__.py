"""Drop-in replacement for the `diff_gaussian_rasterization` package of
rmurai0610/diff-gaussian-rasterization-w-pose as vendored by MonoGS
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py).

Public surface kept exactly: `GaussianRasterizationSettings` (13-field NamedTuple,
ref :173-186), `GaussianRasterizer` (nn.Module with `forward` and `markVisible`,
ref :188-244) and `rasterize_gaussians` (ref :21-46).  `forward` returns
`(color, radii, depth, opacity, n_touched)`; gradients flow to means3D, means2D
(screen-space, used for densification), sh, colors_precomp, opacities, scales,
rotations, cov3D_precomp and to the zero-valued pose leaves `theta` / `rho`.
The kernels behind it are the sm_100a library of this repository.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _snapshot(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


class _RasterizeGaussians(torch.autograd.Function):
    """Autograd node (ref :48-171). Inputs: means3D, means2D, sh, colors_precomp, opacities, scales,
    rotations, cov3Ds_precomp, theta, rho, raster_settings."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                raster_settings):
        s = raster_settings
        args = (s.bg, means3D, colors_precomp, opacities, scales, rotations, s.scale_modifier, cov3Ds_precomp,
                s.viewmatrix, s.projmatrix, s.projmatrix_raw, s.tanfovx, s.tanfovy, s.image_height, s.image_width,
                sh, s.sh_degree, s.campos, s.prefiltered, s.debug)
        if s.debug:
            saved = _snapshot(args)  # copy first: a faulting kernel may corrupt the inputs
            try:
                res = _C.rasterize_gaussians(*args)
            except Exception:
                torch.save(saved, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            res = _C.rasterize_gaussians(*args)
        num_rendered, color, radii, geom, point_list, img, depth, opacity, n_touched = res

        ctx.raster_settings = s
        ctx.num_rendered = num_rendered
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, point_list,
                              img)
        ctx.mark_non_differentiable(radii, n_touched)
        return color, radii, depth, opacity, n_touched

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_opacity, grad_n_touched):
        # grad_opacity / grad_radii / grad_n_touched are ignored exactly like the reference (ref :116-138).
        s = ctx.raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, point_list, img = ctx.saved_tensors
        P = means3D.shape[0]
        if grad_color is None and grad_depth is None:
            grad_color = torch.zeros(3, s.image_height, s.image_width, dtype=torch.float32, device=means3D.device)
        args = (s.bg, means3D, radii, colors_precomp, scales, rotations, s.scale_modifier, cov3Ds_precomp, s.viewmatrix,
                s.projmatrix, s.projmatrix_raw, s.tanfovx, s.tanfovy, grad_color, grad_depth, sh, s.sh_degree,
                s.campos, geom, ctx.num_rendered, point_list, img, s.debug)
        need = ctx.needs_input_grad
        want = dict(means3D=need[0], means2D=need[1], sh=need[2], colors=need[3], opacity=need[4], scales=need[5],
                    rotations=need[6], cov3D=need[7])
        if s.debug:
            saved = _snapshot(args)
            try:
                out, tau = _C.rasterize_gaussians_backward_fused(*args, want=want)
            except Exception:
                torch.save(saved, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            out, tau = _C.rasterize_gaussians_backward_fused(*args, want=want)
        grad_rho = tau[:3].view(1, -1)
        grad_theta = tau[3:].view(1, -1)
        return (out["means3D"], out["means2D"], out["sh"], out["colors"], out["opacity"], out["scales"],
                out["rotations"], out["cov3D"], grad_theta if need[8] else None, grad_rho if need[9] else None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, theta, rho, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of the points in front of the near plane of this camera (ref :193-202)."""
        with torch.no_grad():
            s = self.raster_settings
            return _C.mark_visible(positions, s.viewmatrix, s.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, theta=None, rho=None):
        # NOTE the positional order (means3D, means2D, opacities, shs, ...) -- same as the reference (ref :204).
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (has_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = lambda: torch.Tensor([])
        shs = empty() if shs is None else shs
        colors_precomp = empty() if colors_precomp is None else colors_precomp
        scales = empty() if scales is None else scales
        rotations = empty() if rotations is None else rotations
        cov3D_precomp = empty() if cov3D_precomp is None else cov3D_precomp
        theta = empty() if theta is None else theta
        rho = empty() if rho is None else rho
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   theta, rho, self.raster_settings)
