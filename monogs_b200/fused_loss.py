"""Render + L1 photometric/depth loss in one autograd node (north star (d)).

The unchanged MonoGS callers compute `get_loss_mapping` / `get_loss_tracking` (utils/slam_utils.py:58-146) with 6-10
small framework kernels and boolean-mask indexing, then autograd hands dL/dcolor, dL/ddepth images to the
rasterizer's backward.  This ADDITIONAL entry point (the reference API is untouched) evaluates the same losses with
one reduction kernel right after the forward and lets the backward tile renderer form dL/dpixel on the fly
(gsr_l1_loss_forward / gsr_backward_l1 of include/gsr_b200.h); no gradient image is materialised.

    loss, color, radii, depth, opacity, n_touched = render_l1_loss(
        raster_settings, means3D, means2D, opacities, colors_precomp, scales, rotations, theta, rho,
        gt_rgb, gt_depth, mask=viewpoint.mask, grad_mask=viewpoint.grad_mask,
        exposure_a=viewpoint.exposure_a, exposure_b=viewpoint.exposure_b, kind="tracking")
    loss.backward()

`color`, `depth`, `opacity` are returned detached (the loss already carries their gradient).  Supported: precomputed
colours or SHs, scales+rotations or cov3D_precomp, the non-inverted-depth branches of both losses.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _abi
from .diff_gaussian_rasterization import _C

KINDS = {"mapping": _abi.LOSS_MAPPING, "tracking": _abi.LOSS_TRACKING}


def _u8(m, name, H, W):
    if m is None:
        return None
    if m.numel() != H * W:
        raise RuntimeError(f"{name}: expected {H}x{W} elements")
    m = m.contiguous()
    return m.view(torch.uint8) if m.dtype == torch.bool else (m != 0).view(torch.uint8)


def _scalar(t):
    if t is None:
        return None
    return t.detach().reshape(-1)[:1].to(torch.float32).contiguous()


def l1_loss_images(kind, color, depth, opacity, gt_rgb, gt_depth=None, mask=None, grad_mask=None, exposure_a=None,
                   exposure_b=None, lambda_rgb=0.9, upstream=None, want_grads=True):
    """The callers' loss on given images, without the rasterizer: -> (terms f32[8], dL_dcolor [3,H,W], dL_ddepth
    [1,H,W]).  terms = {loss, l1_rgb, l1_depth, k_rgb, k_depth, dL/da, dL/db, mean opacity} (gsr_l1_loss_forward);
    the two images are upstream * d loss / d(color, depth) exactly as the fused backward forms them per pixel
    (gsr_l1_loss_grad) -- i.e. what autograd would hand the rasterizer after get_loss_mapping / get_loss_tracking
    (utils/slam_utils.py:58-146)."""
    if kind not in KINDS:
        raise ValueError(f"kind must be one of {sorted(KINDS)}")
    lib = _abi.load()
    color = _C._f32(color, "color")
    H, W = int(color.shape[-2]), int(color.shape[-1])
    dev = color.device
    depth_c = None if depth is None else _C._f32(depth, "depth")
    opac_c = None if opacity is None else _C._f32(opacity, "opacity")
    gt_rgb_c = _C._f32(gt_rgb, "gt_rgb")
    gt_depth_c = None if gt_depth is None else _C._f32(gt_depth, "gt_depth")
    mask_c, gmask_c = _u8(mask, "mask", H, W), _u8(grad_mask, "grad_mask", H, W)
    ea, eb = _scalar(exposure_a), _scalar(exposure_b)
    up = None if upstream is None else upstream.detach().reshape(1).to(torch.float32).contiguous()
    out = torch.empty(8, dtype=torch.float32, device=dev)
    acc = torch.empty(8, dtype=torch.float64, device=dev)
    p = _C._ptr
    with _C._on_device(dev):
        rc = lib.gsr_l1_loss_forward(_C._stream(), KINDS[kind], W, H, color.data_ptr(), p(depth_c), p(opac_c),
                                     p(gt_rgb_c), p(gt_depth_c), p(mask_c), p(gmask_c), p(ea), p(eb),
                                     float(lambda_rgb), acc.data_ptr(), out.data_ptr())
        _abi.check(rc, "gsr_l1_loss_forward")
        if not want_grads:
            return out, None, None
        spec = _abi.L1LossSpec(KINDS[kind], color.data_ptr(), p(depth_c), p(opac_c), p(gt_rgb_c), p(gt_depth_c),
                               p(mask_c), p(gmask_c), p(ea), p(eb), out.data_ptr(), p(up))
        d_color = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        d_depth = torch.empty(1, H, W, dtype=torch.float32, device=dev)
        rc = lib.gsr_l1_loss_grad(_C._stream(), W, H, C.byref(spec), d_color.data_ptr(), d_depth.data_ptr())
    _abi.check(rc, "gsr_l1_loss_grad")
    return out, d_color, d_depth


class _RenderL1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                exposure_a, exposure_b, s, gt_rgb, gt_depth, mask, grad_mask, kind, lambda_rgb):
        lib = _abi.load()
        res = _C.rasterize_gaussians(s.bg, means3D, colors_precomp, opacities, scales, rotations, s.scale_modifier,
                                     cov3Ds_precomp, s.viewmatrix, s.projmatrix, s.projmatrix_raw, s.tanfovx,
                                     s.tanfovy, s.image_height, s.image_width, sh, s.sh_degree, s.campos,
                                     s.prefiltered, s.debug)
        R, color, radii, geom, point_list, img, depth, opacity, n_touched = res
        H, W = int(s.image_height), int(s.image_width)
        dev = color.device
        gt_rgb_c = _C._f32(gt_rgb, "gt_rgb")
        gt_depth_c = _C._f32(gt_depth, "gt_depth")
        mask_c, gmask_c = _u8(mask, "mask", H, W), _u8(grad_mask, "grad_mask", H, W)
        ea, eb = _scalar(exposure_a), _scalar(exposure_b)
        out = torch.empty(8, dtype=torch.float32, device=dev)
        acc = torch.empty(8, dtype=torch.float64, device=dev)
        with _C._on_device(dev):
            rc = lib.gsr_l1_loss_forward(_C._stream(), KINDS[kind], W, H, color.data_ptr(), depth.data_ptr(),
                                         opacity.data_ptr(), _C._ptr(gt_rgb_c), _C._ptr(gt_depth_c),
                                         _C._ptr(mask_c), _C._ptr(gmask_c), _C._ptr(ea), _C._ptr(eb),
                                         float(lambda_rgb), acc.data_ptr(), out.data_ptr())
        _abi.check(rc, "gsr_l1_loss_forward")
        ctx.s, ctx.R, ctx.kind = s, R, kind
        ctx.exp_shapes = (tuple(exposure_a.shape), tuple(exposure_b.shape))
        ctx.keep = (gt_rgb_c, gt_depth_c, mask_c, gmask_c, ea, eb)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, point_list,
                              img, color, depth, opacity, out)
        ctx.mark_non_differentiable(color, radii, depth, opacity, n_touched, out)
        return out[0].clone(), color, radii, depth, opacity, n_touched, out

    @staticmethod
    def backward(ctx, g_loss, *_unused):
        lib = _abi.load()
        s = ctx.s
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, point_list, img, color, depth,
         opacity, out) = ctx.saved_tensors
        gt_rgb_c, gt_depth_c, mask_c, gmask_c, ea, eb = ctx.keep
        P = int(means3D.size(0))
        dev = means3D.device
        need = ctx.needs_input_grad
        fopt = dict(dtype=torch.float32, device=dev)
        M = int(sh.size(1)) if sh.numel() else 0
        has_sr = scales.numel() != 0
        shapes = dict(means3D=(P, 3), means2D=(P, 3), sh=(P, M, 3), colors=(P, 3), opacity=(P, 1), scales=(P, 3),
                      rotations=(P, 4), cov3D=(P, 6))
        idx = dict(means3D=0, means2D=1, sh=2, colors=3, opacity=4, scales=5, rotations=6, cov3D=7)
        g = {}
        for n, i in idx.items():
            if not need[i]:
                g[n] = None
            elif P == 0 or (n == "sh" and M == 0) or (n in ("scales", "rotations") and not has_sr):
                g[n] = torch.zeros(*shapes[n], **fopt)
            else:
                g[n] = torch.empty(*shapes[n], **fopt)
        tau = torch.zeros(6, **fopt)
        up = g_loss.detach().reshape(1).to(torch.float32).contiguous()
        if P > 0:
            spec = _abi.L1LossSpec(KINDS[ctx.kind], color.data_ptr(), depth.data_ptr(), opacity.data_ptr(),
                                   _C._ptr(gt_rgb_c), _C._ptr(gt_depth_c), _C._ptr(mask_c), _C._ptr(gmask_c),
                                   _C._ptr(ea), _C._ptr(eb), out.data_ptr(), up.data_ptr())
            f = lambda t, n: _C._f32(t, n)
            with _C._on_device(dev):
                scratch = torch.empty(lib.gsr_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
                keep = [f(s.bg, "bg"), f(means3D, "means3D"), f(sh, "sh"), f(colors_precomp, "colors"),
                        f(scales, "scales"), f(rotations, "rotations"), f(cov3Ds_precomp, "cov3D"),
                        f(s.viewmatrix, "viewmatrix"), f(s.projmatrix, "projmatrix"),
                        f(s.projmatrix_raw, "projmatrix_raw"), f(s.campos, "campos")]
                bg, m3, shc, col, sc, ro, cv, vm, pm, pr, cp = keep
                p = _C._ptr
                rc = lib.gsr_backward_l1(
                    _C._stream(), P, int(s.sh_degree), M, int(s.image_width), int(s.image_height), p(bg), p(m3), p(shc),
                    p(col), p(sc), float(s.scale_modifier), p(ro), p(cv), p(vm), p(pm), p(pr), p(cp), float(s.tanfovx),
                    float(s.tanfovy), radii.data_ptr(), geom.data_ptr(), img.data_ptr(), p(point_list), int(ctx.R),
                    C.byref(spec), scratch.data_ptr(), p(g["means2D"]), p(g["colors"]), p(g["opacity"]),
                    p(g["means3D"]), p(g["cov3D"]), p(g["sh"]), p(g["scales"]), p(g["rotations"]), None,
                    tau.data_ptr(), int(bool(s.debug)))
            _abi.check(rc, "gsr_backward_l1")
        g_a = (g_loss * out[5]).reshape(ctx.exp_shapes[0]) if need[10] else None
        g_b = (g_loss * out[6]).reshape(ctx.exp_shapes[1]) if need[11] else None
        return (g["means3D"], g["means2D"], g["sh"], g["colors"], g["opacity"], g["scales"], g["rotations"],
                g["cov3D"], tau[3:].view(1, -1) if need[8] else None, tau[:3].view(1, -1) if need[9] else None,
                g_a, g_b, None, None, None, None, None, None, None)


def render_l1_loss(raster_settings, means3D, means2D, opacities, colors_precomp=None, scales=None, rotations=None,
                   theta=None, rho=None, gt_rgb=None, gt_depth=None, mask=None, grad_mask=None, exposure_a=None,
                   exposure_b=None, kind="mapping", lambda_rgb=0.9, shs=None, cov3D_precomp=None,
                   return_terms=False):
    """-> (loss, color, radii, depth, opacity, n_touched[, terms]); terms = f32[8] {loss, l1_rgb, l1_depth, k_rgb,
    k_depth, dL/da, dL/db, mean opacity}. See the module docstring."""
    if kind not in KINDS:
        raise ValueError(f"kind must be one of {sorted(KINDS)}")
    if gt_rgb is None:
        raise ValueError("gt_rgb is required")
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if ((scales is None or rotations is None) and cov3D_precomp is None) or \
            ((scales is not None or rotations is not None) and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
    dev = means3D.device
    e = lambda: torch.empty(0, device=dev)
    z = lambda t: e() if t is None else t
    ea = exposure_a if exposure_a is not None else e()
    eb = exposure_b if exposure_b is not None else e()
    out = _RenderL1.apply(means3D, means2D, z(shs), z(colors_precomp), opacities, z(scales), z(rotations),
                          z(cov3D_precomp), z(theta), z(rho), ea, eb, raster_settings, gt_rgb, gt_depth, mask,
                          grad_mask, kind, lambda_rgb)
    loss, color, radii, depth, opacity, n_touched, terms = out
    if return_terms:
        return loss, color, radii, depth, opacity, n_touched, terms.detach()
    return loss, color, radii, depth, opacity, n_touched
