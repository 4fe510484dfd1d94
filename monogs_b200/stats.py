"""Mapper bookkeeping next to the rasterizer's outputs, fused on the device.

* add_densification_stats  -- utils/slam_mapper.py:452-460 + gaussian_splatting/scene/gaussian_model.py:888-892:
  `max_radii_2d[vis] = max(max_radii_2d[vis], radii[vis])`, `xyz_gradient_accum[vis] += ||grad[vis, :2]||`,
  `denom[vis] += 1` with vis = radii > 0, in ONE kernel and without the boolean-mask gathers (nonzero + sync).
* visibility_prune_mask    -- utils/slam_mapper.py:399-440: `nr_obs = sum_k (n_touched_k > 0)`,
  `to_prune = (nr_obs <= prune_coviz) & (kf_idx >= min_kf)`; the reference keeps `nr_obs` / `kf_idx` on the CPU and
  copies every visibility vector there, here everything stays on the GPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _abi
from .diff_gaussian_rasterization import _C

MAX_KEYFRAMES = 16


def _i32(t, name):
    if t.dtype != torch.int32 or not t.is_cuda:
        raise RuntimeError(f"{name}: expected an int32 CUDA tensor")
    return t.contiguous()


def _inplace_f32(t, name, P):
    if t is None:
        return None
    if t.dtype != torch.float32 or not t.is_cuda or not t.is_contiguous() or t.numel() != P:
        raise RuntimeError(f"{name}: expected a contiguous float32 CUDA tensor with {P} elements (updated in place)")
    return t


def add_densification_stats(radii: torch.Tensor, viewspace_grad: Optional[torch.Tensor],
                            max_radii_2d: Optional[torch.Tensor], xyz_gradient_accum: Optional[torch.Tensor],
                            denom: Optional[torch.Tensor]) -> None:
    """In-place update of the three statistics for the Gaussians with radii > 0 (see module docstring).
    radii i32[P]; viewspace_grad f32[P,3] (the `.grad` of render()'s `viewspace_points`) or None."""
    P = int(radii.numel())
    if P == 0:
        return
    lib = _abi.load()
    radii_c = _i32(radii, "radii")
    g = None if viewspace_grad is None else _C._f32(viewspace_grad, "viewspace_grad")
    if g is not None and (g.dim() != 2 or g.shape[0] != P or g.shape[1] != 3):
        raise RuntimeError("viewspace_grad: expected [P, 3]")
    mr = _inplace_f32(max_radii_2d, "max_radii_2d", P)
    ga, dn = _inplace_f32(xyz_gradient_accum, "xyz_gradient_accum", P), _inplace_f32(denom, "denom", P)
    with _C._on_device(radii.device):
        rc = lib.gsr_densify_stats(_C._stream(), P, radii_c.data_ptr(), _C._ptr(g), _C._ptr(mr), _C._ptr(ga),
                                   _C._ptr(dn))
    _abi.check(rc, "gsr_densify_stats")


def visibility_prune_mask(n_touched: Sequence[torch.Tensor], kf_idx: torch.Tensor, min_kf: int, coviz: int = 3):
    """-> (nr_obs i32[P], to_prune bool[P]) for the window's n_touched outputs (<= 16 keyframes per call)."""
    K = len(n_touched)
    if K == 0 or K > MAX_KEYFRAMES:
        raise ValueError(f"between 1 and {MAX_KEYFRAMES} keyframes per call")
    P = int(kf_idx.numel())
    dev = kf_idx.device
    nr_obs = torch.empty(P, dtype=torch.int32, device=dev)
    to_prune = torch.empty(P, dtype=torch.uint8, device=dev)
    if P == 0:
        return nr_obs, to_prune.bool()
    lib = _abi.load()
    keep = [_i32(t, f"n_touched[{k}]") for k, t in enumerate(n_touched)]
    if any(t.numel() != P for t in keep):
        raise RuntimeError("n_touched: every array must have kf_idx's length")
    ptrs = (C.c_void_p * K)(*[t.data_ptr() for t in keep])
    kf = _i32(kf_idx, "kf_idx")
    with _C._on_device(dev):
        rc = lib.gsr_visibility_prune(_C._stream(), P, K, ptrs, kf.data_ptr(), int(min_kf), int(coviz),
                                      nr_obs.data_ptr(), to_prune.data_ptr())
    _abi.check(rc, "gsr_visibility_prune")
    return nr_obs, to_prune.view(torch.bool)
