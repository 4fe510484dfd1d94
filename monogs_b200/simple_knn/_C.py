"""Mirror of simple_knn._C (submodules/simple-knn/ext.cpp:15-17, spatial.cu:15-26)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _abi


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    """points f32[P,3] (CUDA) -> f32[P]: mean squared distance to the 3 nearest other points."""
    lib = _abi.load()
    P = int(points.size(0))
    out = torch.zeros(P, dtype=torch.float32, device=points.device)
    if P == 0:
        return out
    if points.dtype != torch.float32 or not points.is_cuda:
        raise RuntimeError("distCUDA2: expected a float32 CUDA tensor")
    pts = points.contiguous()
    with torch.cuda.device(points.device):
        ws = torch.empty(lib.gsr_knn_bytes(P), dtype=torch.uint8, device=points.device)
        rc = lib.gsr_knn_dist2(C.c_void_p(torch.cuda.current_stream().cuda_stream), P, pts.data_ptr(),
                               out.data_ptr(), ws.data_ptr())
        _abi.check(rc, "gsr_knn_dist2")
    return out
