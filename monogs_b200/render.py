"""Drop-in for MonoGS's render() wrapper with a cached-camera fast path.

Reference: gaussian_splatting/gaussian_renderer/__init__.py:26-170 (same signature, same returned dict) on top of
utils/camera_utils.py:30-49 (FoVx / FoVy / projection_matrix properties), :171-178 (world_view_transform,
camera_center) and :224-231 (get_full_proj_transform).  Per call the reference pays two `.cpu().item()` stream
synchronisations (FoVx, FoVy), a rebuild of the projection matrix out of ~20 one-element tensor kernels, a 4x4
assembly, a general 4x4 `inverse()` and a `bmm`; at > 1 k rasterizer iterations/s that host work is the tracking
loop's bottleneck.  Here
  * everything that depends only on the intrinsics (tan(FoV/2), the transposed projection matrix) is computed ONCE per
    version of `fx` / `fy` (tensor `_version` counters) with the reference's own arithmetic, so the numbers are
    bit-identical to what its properties return;
  * the per-pose matrices (view, full projection, camera centre) come from ONE kernel, `gsr_camera_setup`, with no
    host synchronisation.
The reference file is not edited: callers `from monogs_b200.render import render` (or `install_aliases`).
"""
from __future__ import annotations

import math
import weakref

import torch

from . import _abi
from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from .diff_gaussian_rasterization import _C

ZNEAR, ZFAR = 0.01, 100.0  # camera_utils.py:41-42

_intrinsics_cache: "weakref.WeakKeyDictionary" = weakref.WeakKeyDictionary()


def projection_matrix_T(fx, fy, cx, cy, W, H, znear=ZNEAR, zfar=ZFAR):
    """Transposed OpenCV-intrinsics projection matrix, evaluated with the operation order of
    graphics_utils.getProjectionMatrix (:68-89) on `fx`, `fy` one-element device tensors (so every element carries
    the same fp32 roundings as `CameraIntrinsics.projection_matrix`)."""
    ndc_x, ndc_y = (2 * cx - W) / W, (2 * cy - H) / H
    kx, ky = znear / fx, znear / fy
    x_lo, x_hi = kx * ((ndc_x - 1.0) * W / 2.0), kx * ((ndc_x + 1.0) * W / 2.0)
    y_hi, y_lo = ky * ((ndc_y + 1.0) * H / 2.0), ky * ((ndc_y - 1.0) * H / 2.0)
    P = torch.zeros(4, 4, device=fx.device)
    P[0, 0] = 2.0 * znear / (x_hi - x_lo)
    P[1, 1] = 2.0 * znear / (y_hi - y_lo)
    P[0, 2] = (x_hi + x_lo) / (x_hi - x_lo)
    P[1, 2] = (y_hi + y_lo) / (y_hi - y_lo)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P.transpose(0, 1).contiguous()


def camera_constants(cam_intrinsics):
    """(tanfovx, tanfovy, proj_T) of a CameraIntrinsics-like object (fx, fy: one-element tensors; cx, cy, width,
    height: numbers). Cached per object and per in-place version of fx / fy; not cached while they require grad."""
    fx, fy = cam_intrinsics.fx, cam_intrinsics.fy
    key = (fx._version, fy._version, fx.data_ptr(), fy.data_ptr(), float(cam_intrinsics.cx), float(cam_intrinsics.cy),
           int(cam_intrinsics.width), int(cam_intrinsics.height))
    cacheable = not (fx.requires_grad or fy.requires_grad)
    if cacheable:
        hit = _intrinsics_cache.get(cam_intrinsics)
        if hit is not None and hit[0] == key:
            return hit[1], hit[2], hit[3]
    with torch.no_grad():
        # FoV = 2 * atan(size / (2 f)) in fp32 on the device like the properties; ONE host read for both
        half = torch.cat([torch.atan(cam_intrinsics.width / (2 * fx.detach())),
                          torch.atan(cam_intrinsics.height / (2 * fy.detach()))]).cpu()
        tanfovx = math.tan((2 * half[0].item()) * 0.5)
        tanfovy = math.tan((2 * half[1].item()) * 0.5)
    if cacheable:
        with torch.no_grad():
            proj_T = projection_matrix_T(fx.detach(), fy.detach(), cam_intrinsics.cx, cam_intrinsics.cy,
                                         cam_intrinsics.width, cam_intrinsics.height)
        _intrinsics_cache[cam_intrinsics] = (key, tanfovx, tanfovy, proj_T)
    else:
        proj_T = projection_matrix_T(fx, fy, cam_intrinsics.cx, cam_intrinsics.cy, cam_intrinsics.width,
                                     cam_intrinsics.height)
    return tanfovx, tanfovy, proj_T


def camera_matrices(R: torch.Tensor, T: torch.Tensor, proj_T: torch.Tensor):
    """(viewmatrix[4,4], projmatrix[4,4], campos[3]) for the world->camera rotation R[3,3] and translation T[3]
    (any strides, fp32, CUDA) in one launch: gsr_camera_setup."""
    if not R.is_cuda:
        raise RuntimeError("monogs_b200.render: CUDA tensors required (there is no CPU path)")
    if R.dtype != torch.float32 or T.dtype != torch.float32:
        R, T = R.float(), T.float()
    lib = _abi.load()
    buf = torch.empty(35, dtype=torch.float32, device=R.device)
    p = proj_T.detach()
    if not p.is_contiguous():
        p = p.contiguous()
    with _C._on_device(R.device):
        rc = lib.gsr_camera_setup(_C._stream(), R.data_ptr(), int(R.stride(0)), int(R.stride(1)), T.data_ptr(),
                                  int(T.stride(0)), p.data_ptr(), buf.data_ptr(), buf.data_ptr() + 64,
                                  buf.data_ptr() + 128)
    _abi.check(rc, "gsr_camera_setup")
    return buf[:16].view(4, 4), buf[16:32].view(4, 4), buf[32:35]


def render(viewpoint_camera, cam_intrinsics, means, rotations, scales, opacity, features, bg_color,
           scaling_modifier=1.0, override_color=None, mask=None):
    """Same contract as the reference's render(): returns None for an empty map, else the dict
    {render, viewspace_points, visibility_filter, radii, depth, opacity, n_touched}. `override_color` replaces
    `features` when given; `mask` selects a subset of the Gaussians (n_touched then refers to the subset)."""
    if means.shape[0] == 0:
        return None
    screenspace_points = torch.zeros_like(means, requires_grad=True)  # leaf: receives dL/dmean2D in .grad
    tanfovx, tanfovy, proj_T = camera_constants(cam_intrinsics)
    viewmatrix, projmatrix, campos = camera_matrices(viewpoint_camera.R, viewpoint_camera.T, proj_T)
    settings = GaussianRasterizationSettings(
        image_height=int(cam_intrinsics.height), image_width=int(cam_intrinsics.width), tanfovx=tanfovx,
        tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier, viewmatrix=viewmatrix, projmatrix=projmatrix,
        projmatrix_raw=proj_T, sh_degree=0, campos=campos, prefiltered=False, debug=False)
    if scales.shape[-1] == 1:  # isotropic map: one scale per Gaussian
        scales = scales.expand(-1, 3)
    colors = features if override_color is None else override_color
    means2D = screenspace_points
    if mask is not None:
        means, means2D, colors, opacity, scales, rotations = (means[mask], means2D[mask], colors[mask], opacity[mask],
                                                              scales[mask], rotations[mask])
    image, radii, depth, alpha, n_touched = GaussianRasterizer(raster_settings=settings)(
        means3D=means, means2D=means2D, shs=None, colors_precomp=colors, opacities=opacity, scales=scales,
        rotations=rotations, cov3D_precomp=None, theta=viewpoint_camera.cam_rot_delta,
        rho=viewpoint_camera.cam_trans_delta)
    return {"render": image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii,
            "depth": depth, "opacity": alpha, "n_touched": n_touched}
