"""oracle/ref_cuda.py -- TEST INFRASTRUCTURE (never imported by the product path).

ctypes wrapper around oracle/_ref/libref_dgr.so = the UNMODIFIED reference CUDA rasterizer +
simple-knn compiled for sm_100a by oracle/Makefile from /root/reference (see oracle/ref_shim.cu).
It exposes the same three functions as the reference's pybind module
(submodules/diff-gaussian-rasterization/ext.cpp:15-19) plus `distCUDA2`, operating on torch CUDA
tensors, so parity tests and the `--impl reference` benchmark arm can run the real reference on the
GPU box (where /root/reference itself does not exist).
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

_LIB = Path(__file__).resolve().parent / "_ref" / "libref_dgr.so"
_lib = None
vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
# Saved-state slots of the shim (oracle/ref_shim.cu): with a ring of n > 1 every forward takes the next slot and
# returns its number in the `geomBuffer` position, the place where the reference keeps that state
# (diff_gaussian_rasterization/__init__.py:104); the backward then consumes the slot it is handed. n = 1 (default)
# keeps the one-state behaviour: backward consumes the last forward.
_ring = 1
_next_slot = 0


def set_ring(n: int) -> None:
    global _ring, _next_slot
    if not 1 <= n <= 16:
        raise ValueError("ring size must be 1..16")
    _ring, _next_slot = n, 0
    load().ref_set_slot(0)


def available() -> bool:
    return _LIB.exists()


def load():
    global _lib
    if _lib is None:
        if not _LIB.exists():
            raise RuntimeError(f"{_LIB} missing: run `make -C oracle ref` where /root/reference exists")
        lib = C.CDLL(str(_LIB))
        lib.ref_forward.restype = i32
        lib.ref_forward.argtypes = [i32, i32, i32, vp, i32, i32, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, f32, f32,
                                    i32, vp, vp, vp, vp, vp, i32]
        lib.ref_backward.restype = None
        lib.ref_backward.argtypes = [i32, i32, i32, i32, vp, i32, i32, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, vp,
                                     f32, f32, vp, vp, vp] + [vp] * 11 + [i32]
        lib.ref_mark_visible.restype = None
        lib.ref_mark_visible.argtypes = [i32, vp, vp, vp, vp]
        lib.ref_last_state.restype = None
        lib.ref_last_state.argtypes = [C.POINTER(vp)] * 11
        lib.ref_knn.restype = None
        lib.ref_knn.argtypes = [i32, vp, vp]
        lib.ref_sync.restype = i32
        lib.ref_set_slot.restype = i32
        lib.ref_set_slot.argtypes = [i32]
        _lib = lib
    return _lib


def _p(t):
    return None if (t is None or t.numel() == 0) else t.data_ptr()


def _c(t):
    return None if (t is None or t.numel() == 0) else t.contiguous()


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy, image_height, image_width, sh,
                        degree, campos, prefiltered, debug):
    """Same contract as RasterizeGaussiansCUDA (rasterize_points.cu:35-122); the three state buffers live inside
    the shim, so the returned buffer slots are None. Uses the legacy default stream like the reference."""
    lib = load()
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    dev = means3D.device
    fo = dict(dtype=torch.float32, device=dev)
    color = torch.zeros(3, H, W, **fo)
    depth = torch.zeros(1, H, W, **fo)
    opac = torch.zeros(1, H, W, **fo)
    radii = torch.zeros(P, dtype=torch.int32, device=dev)
    n_touched = torch.zeros(P, dtype=torch.int32, device=dev)
    R = 0
    slot_t = None
    if _ring > 1:
        global _next_slot
        slot = _next_slot
        _next_slot = (_next_slot + 1) % _ring
        lib.ref_set_slot(slot)
        slot_t = torch.tensor([slot], dtype=torch.int32)
    if P != 0:
        M = int(sh.size(1)) if (sh is not None and sh.numel()) else 0
        keep = [_c(t) for t in (background, means3D, sh, colors, opacity, scales, rotations, cov3D_precomp,
                                viewmatrix, projmatrix, campos)]
        bg, m3, shc, col, op, sc, ro, cv, vm, pm, cp = keep
        R = lib.ref_forward(P, int(degree), M, _p(bg), W, H, _p(m3), _p(shc), _p(col), _p(op), _p(sc),
                            float(scale_modifier), _p(ro), _p(cv), _p(vm), _p(pm), _p(cp), float(tan_fovx),
                            float(tan_fovy), int(bool(prefiltered)), color.data_ptr(), depth.data_ptr(),
                            opac.data_ptr(), radii.data_ptr(), n_touched.data_ptr(), int(bool(debug)))
    return R, color, radii, slot_t, None, None, depth, opac, n_touched


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_depth, sh, degree, campos, geomBuffer, R, binningBuffer,
                                 imageBuffer, debug):
    """Same contract as RasterizeGaussiansBackwardCUDA (rasterize_points.cu:124-211). Must follow the
    rasterize_gaussians call whose state it consumes (the shim keeps exactly one)."""
    lib = load()
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    M = int(sh.size(1)) if (sh is not None and sh.numel()) else 0
    fo = dict(dtype=torch.float32, device=means3D.device)
    z = lambda *s: torch.zeros(*s, **fo)
    dmeans3D, dmeans2D, dcolors, ddepths = z(P, 3), z(P, 3), z(P, 3), z(P, 1)
    dconic, dopacity, dcov3D, dsh = z(P, 2, 2), z(P, 1), z(P, 6), z(P, M, 3)
    dscales, drots, dtau = z(P, 3), z(P, 4), z(P, 6)
    if isinstance(geomBuffer, torch.Tensor) and geomBuffer.numel() == 1:
        lib.ref_set_slot(int(geomBuffer.item()))
    if P != 0:
        keep = [_c(t) for t in (background, means3D, sh, colors, scales, rotations, cov3D_precomp, viewmatrix,
                                projmatrix, projmatrix_raw, campos, radii, dL_dout_color, dL_dout_depth)]
        bg, m3, shc, col, sc, ro, cv, vm, pm, pr, cp, rd, gc, gd = keep
        lib.ref_backward(P, int(degree), M, int(R), _p(bg), W, H, _p(m3), _p(shc), _p(col), _p(sc),
                         float(scale_modifier), _p(ro), _p(cv), _p(vm), _p(pm), _p(pr), _p(cp), float(tan_fovx),
                         float(tan_fovy), _p(rd), _p(gc), _p(gd), dmeans2D.data_ptr(), dconic.data_ptr(),
                         dopacity.data_ptr(), dcolors.data_ptr(), ddepths.data_ptr(), dmeans3D.data_ptr(),
                         dcov3D.data_ptr(), _p(dsh), dscales.data_ptr(), drots.data_ptr(), dtau.data_ptr(),
                         int(bool(debug)))
    return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drots, dtau


def mark_visible(means3D, viewmatrix, projmatrix):
    lib = load()
    P = int(means3D.size(0))
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    if P:
        m, v, p = means3D.contiguous(), viewmatrix.contiguous(), projmatrix.contiguous()
        lib.ref_mark_visible(P, m.data_ptr(), v.data_ptr(), p.data_ptr(), present.data_ptr())
    return present


def last_state(P: int, R: int, W: int, H: int, device):
    """Views (copies) of the reference's internal state after the last forward: dict of torch tensors."""
    lib = load()
    ptrs = [vp() for _ in range(11)]
    lib.ref_last_state(*[C.byref(x) for x in ptrs])
    names = ["depths", "means2D", "cov3D", "conic_opacity", "tiles_touched", "point_offsets", "point_list_keys",
             "point_list", "ranges", "n_contrib", "accum_alpha"]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    spec = dict(depths=(torch.float32, P), means2D=(torch.float32, 2 * P), cov3D=(torch.float32, 6 * P),
                conic_opacity=(torch.float32, 4 * P), tiles_touched=(torch.int32, P), point_offsets=(torch.int32, P),
                point_list_keys=(torch.int64, R), point_list=(torch.int32, R), ranges=(torch.int32, 2 * tiles),
                n_contrib=(torch.int32, W * H), accum_alpha=(torch.float32, W * H))
    out = {}
    for n, ptr in zip(names, ptrs):
        dt, cnt = spec[n]
        t = torch.empty(cnt, dtype=dt, device=device)
        if cnt:
            nbytes = cnt * t.element_size()
            torch.cuda.synchronize()
            _memcpy_d2d(t.data_ptr(), ptr.value, nbytes)
        out[n] = t
    return out


def _memcpy_d2d(dst, src, nbytes):
    # tiny helper through torch: wrap the raw pointer with a uint8 tensor via cuda array interface
    class _Raw:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = dict(shape=(n,), typestr="|u1", data=(ptr, False), version=2)
    s = torch.as_tensor(_Raw(src, nbytes), device="cuda")
    d = torch.as_tensor(_Raw(dst, nbytes), device="cuda")
    d.copy_(s)
    torch.cuda.synchronize()


def distCUDA2(points):
    lib = load()
    P = int(points.size(0))
    out = torch.zeros(P, dtype=torch.float32, device=points.device)
    if P:
        pts = points.contiguous()
        torch.cuda.synchronize()
        lib.ref_knn(P, pts.data_ptr(), out.data_ptr())
        torch.cuda.synchronize()
    return out


def sync():
    return load().ref_sync()
