"""Gaussian insertion from an RGB-D keyframe on the device.

`create_viewpoint_pcd` has the arguments and the returned 6-tuple of the reference's
`GaussianModel.create_viewpoint_pcd` (gaussian_splatting/scene/gaussian_model.py:121-319) -- the caller of
`distCUDA2` -- so `GaussianModel.extend_from_viewpoint` can call it unchanged:

    points_3d [K,3], features [K,3], scales [K,1 or 3] (log), rots [K,4], opacities [K,1] (logit), points_ids [K]

The reference assembles it from ~40 framework ops, two of them sort-based `torch.median` calls over the whole image
and four boolean-mask indexings; here the device side is 4 radix-select launches per median, a mask + stable
compaction pair, one gather/unproject kernel, `distCUDA2` (csrc/knn.cu) and one finishing kernel, all through the
C ABI (gsr_select_kth_f32, gsr_insert_select, gsr_insert_points, gsr_knn_dist2, gsr_insert_finish).  What stays on
the host is what the reference also does there: reading the number of selected pixels (a boolean-mask indexing in
the reference, :242-245) and `torch.randperm` on the host RNG (:256), so a seeded run keeps the same pixels
(`rng="device"` draws on the GPU instead).  The intrinsics inverse and [R|T]^-1 (:163-165) are formed inside the
gather kernel.  No CPU fallback: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _abi

DOWNSAMPLE_INIT = 32   # gaussian_model.py:169  (pcd_downsample_init)
DOWNSAMPLE = 64        # gaussian_model.py:171  (pcd_downsample)


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (there is no CPU path)")
    return t.detach().to(torch.float32).contiguous()


def median_lower(values: torch.Tensor, other: Optional[torch.Tensor] = None) -> torch.Tensor:
    """torch.median semantics (the lower median) of `values`, or of |values - other|, as a device scalar; radix
    select (gsr_select_kth_f32), no sort, no host synchronisation."""
    lib = _abi.load()
    a = _f32(values, "values").view(-1)
    b = _f32(other, "other").view(-1) if other is not None else None
    n = a.numel()
    if n == 0:
        raise RuntimeError("median of an empty tensor")
    if b is not None and b.numel() != n:
        raise RuntimeError("median_lower: shapes differ")
    out = torch.empty(1, dtype=torch.float32, device=a.device)
    ws = torch.empty(8192, dtype=torch.uint8, device=a.device)
    with torch.cuda.device(a.device):
        rc = lib.gsr_select_kth_f32(_stream(a.device), n, a.data_ptr(), b.data_ptr() if b is not None else None,
                                    (n - 1) // 2, out.data_ptr(), ws.data_ptr())
    _abi.check(rc, "gsr_select_kth_f32")
    return out[0]


def w2c_from_RT(R: torch.Tensor, T: torch.Tensor) -> torch.Tensor:
    """getWorld2View (utils/camera_utils.py / graphics_utils.py): [[R, T], [0, 1]]."""
    m = torch.zeros(4, 4, dtype=torch.float32, device=R.device)
    m[:3, :3] = R
    m[:3, 3] = T
    m[3, 3] = 1.0
    return m


def create_viewpoint_pcd(viewpoint, cam_intrinsics, render_depth=None, render_opacity=None, init=False,
                         isotropic=True, random_indices: Optional[torch.Tensor] = None, rng: str = "host"):
    """See the module docstring. `viewpoint` needs .rgb [3,H,W], .depth [H,W], .segmentation [H,W], .R, .T and,
    unless init, .exposure_a / .exposure_b; `cam_intrinsics` needs fx, fy, cx, cy. `render_depth` /
    `render_opacity` are the rasterizer's [1,H,W] outputs. `random_indices` (optional, int64) replaces the
    torch.randperm draw -- parity tests pass the reference's draw; rng="device" draws it on the GPU instead."""
    lib = _abi.load()
    rgb = _f32(viewpoint.rgb, "viewpoint.rgb")
    dev = rgb.device
    H, W = int(rgb.shape[1]), int(rgb.shape[2])
    gt_depth = _f32(viewpoint.depth, "viewpoint.depth").view(H, W)
    seg = viewpoint.segmentation
    seg_dtype = seg.dtype
    seg32 = seg.contiguous().view(torch.int32) if seg.element_size() == 4 else seg.to(torch.int32).contiguous()
    fnum = lambda v: float(v.item()) if torch.is_tensor(v) else float(v)
    fx, fy, cx, cy = (fnum(getattr(cam_intrinsics, k)) for k in ("fx", "fy", "cx", "cy"))
    R_w2c = viewpoint.R.to(device=dev, dtype=torch.float32).contiguous()      # K^-1 and [R|T]^-1 (:163-165) are
    T_w2c = viewpoint.T.to(device=dev, dtype=torch.float32).contiguous()      # formed inside gsr_insert_points

    with torch.cuda.device(dev):
        st = _stream(dev)
        ws = torch.empty(int(lib.gsr_insert_ws_bytes(W, H)), dtype=torch.uint8, device=dev)
        med_depth = torch.empty(1, dtype=torch.float32, device=dev)
        n = H * W
        _abi.check(lib.gsr_select_kth_f32(st, n, gt_depth.data_ptr(), None, (n - 1) // 2, med_depth.data_ptr(),
                                          ws.data_ptr()), "gsr_select_kth_f32")                     # :177
        med_err, rd, ro = None, None, None
        if not init:
            if render_depth is None:
                raise RuntimeError("render_depth is required unless init=True")
            rd = _f32(render_depth, "render_depth").view(H, W)
            ro = _f32(render_opacity, "render_opacity").view(H, W) if render_opacity is not None else None
            med_err = torch.empty(1, dtype=torch.float32, device=dev)
            _abi.check(lib.gsr_select_kth_f32(st, n, gt_depth.data_ptr(), rd.data_ptr(), (n - 1) // 2,
                                              med_err.data_ptr(), ws.data_ptr()), "gsr_select_kth_f32")  # :218
        sel = torch.empty(n, dtype=torch.int32, device=dev)
        count = torch.zeros(1, dtype=torch.int32, device=dev)
        _abi.check(lib.gsr_insert_select(st, W, H, gt_depth.data_ptr(), rd.data_ptr() if rd is not None else None,
                                         ro.data_ptr() if ro is not None else None, int(bool(init)),
                                         med_err.data_ptr() if med_err is not None else None, sel.data_ptr(),
                                         count.data_ptr(), ws.data_ptr()), "gsr_insert_select")
        num_points = int(count.item())                                       # the reference's mask indexing syncs too
        factor = DOWNSAMPLE_INIT if init else DOWNSAMPLE
        keep = int(num_points * (1.0 / factor))                              # :249-254
        if random_indices is None:
            if rng == "device":   # same distribution, drawn on the GPU (no 1-5 ms host permutation + upload)
                random_indices = torch.randperm(num_points, device=dev)[:keep]
            else:                 # host RNG, like the reference (:256): a seeded run keeps the same pixels
                random_indices = torch.randperm(num_points)[:keep]
        else:
            ri = torch.as_tensor(random_indices)
            if ri.numel() and (int(ri.max()) >= num_points or int(ri.min()) < 0):
                raise RuntimeError("random_indices out of range")
        pick = torch.as_tensor(random_indices).to(device=dev, dtype=torch.int64, non_blocking=True).contiguous()
        K = int(pick.numel())
        xyz = torch.empty(K, 3, dtype=torch.float32, device=dev)
        feat = torch.empty(K, 3, dtype=torch.float32, device=dev)
        ids = torch.empty(K, dtype=torch.int32, device=dev)
        cols = 1 if isotropic else 3
        scales = torch.empty(K, cols, dtype=torch.float32, device=dev)
        rots = torch.empty(K, 4, dtype=torch.float32, device=dev)
        opac = torch.empty(K, 1, dtype=torch.float32, device=dev)
        if K:
            ea = eb = None
            if not init:
                ea = _f32(viewpoint.exposure_a, "exposure_a").view(-1)
                eb = _f32(viewpoint.exposure_b, "exposure_b").view(-1)
            _abi.check(lib.gsr_insert_points(st, K, sel.data_ptr(), pick.data_ptr(), W, H, rgb.data_ptr(),
                                             ea.data_ptr() if ea is not None else None,
                                             eb.data_ptr() if eb is not None else None, gt_depth.data_ptr(),
                                             seg32.data_ptr(), R_w2c.data_ptr(), T_w2c.data_ptr(), fx, fy, cx, cy,
                                             xyz.data_ptr(), feat.data_ptr(), ids.data_ptr()), "gsr_insert_points")
            dist2 = torch.empty(K, dtype=torch.float32, device=dev)
            knn_ws = torch.empty(int(lib.gsr_knn_bytes(K)), dtype=torch.uint8, device=dev)
            _abi.check(lib.gsr_knn_dist2(st, K, xyz.data_ptr(), dist2.data_ptr(), knn_ws.data_ptr()),
                       "gsr_knn_dist2")                                      # distCUDA2, :296
            _abi.check(lib.gsr_insert_finish(st, K, dist2.data_ptr(), med_depth.data_ptr(), cols, scales.data_ptr(),
                                             rots.data_ptr(), opac.data_ptr()), "gsr_insert_finish")
    points_ids = ids.view(seg_dtype) if seg.element_size() == 4 else ids.to(seg_dtype)
    return xyz, feat, scales, rots, opac, points_ids
