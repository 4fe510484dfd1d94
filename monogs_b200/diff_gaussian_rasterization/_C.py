"""Host-side mirror of the reference's pybind module `diff_gaussian_rasterization._C`
(submodules/diff-gaussian-rasterization/ext.cpp:15-19), implemented on top of the
C ABI in include/gsr_b200.h via ctypes.  Same function names, argument order and
return tuples as RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA /
markVisible (rasterize_points.cu:35-232).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _abi

NUM_CHANNELS = 3
_capacity = {}  # (P, W, H, device) -> instance capacity for the bubble-free gsr_forward path


def _ptr(t):
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32(t, name):
    """contiguous fp32 CUDA tensor or None for an absent (empty) one."""
    if t is None or t.numel() == 0:
        return None
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name}: expected scalar type Float but found {t.dtype}")
    if not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor")
    return t.contiguous()


def _stream():
    """The caller's current CUDA stream (raw handle; the C call is ~10x cheaper than building a Stream object)."""
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch.cuda.current_device()))


class _on_device:
    """`with _on_device(dev)` that costs nothing when `dev` already is the current device (the usual case:
    one process per GPU); the generic context manager is ~15 us of Python per use."""
    __slots__ = ("ctx",)

    def __init__(self, dev):
        self.ctx = None if dev.index is None or dev.index == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy, image_height, image_width,
                        sh, degree, campos, prefiltered, debug):
    """-> (num_rendered, color[3,H,W], radii[P] i32, geomBuffer u8, binningBuffer u8, imgBuffer u8,
           depth[1,H,W], opacity[1,H,W], n_touched[P] i32)   (rasterize_points.cu:35-122)"""
    lib = _abi.load()
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:58-60
    if not means3D.is_cuda:
        raise RuntimeError("means3D: expected a CUDA tensor (this rasterizer has no CPU path)")
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    dev = means3D.device
    fopt = dict(dtype=torch.float32, device=dev)
    if P == 0:
        z = lambda *s: torch.zeros(*s, **fopt)
        e = torch.empty(0, dtype=torch.uint8, device=dev)
        zi = torch.zeros(0, dtype=torch.int32, device=dev)
        return 0, z(NUM_CHANNELS, H, W), zi, e, e.clone(), e.clone(), z(1, H, W), z(1, H, W), zi.clone()

    with _on_device(dev):
        means3D_c = _f32(means3D, "means3D")
        colors_c, opacity_c = _f32(colors, "colors"), _f32(opacity, "opacity")
        scales_c, rot_c, cov_c = _f32(scales, "scales"), _f32(rotations, "rotations"), _f32(cov3D_precomp, "cov3D_precomp")
        sh_c, campos_c = _f32(sh, "sh"), _f32(campos, "campos")
        view_c, proj_c, bg_c = _f32(viewmatrix, "viewmatrix"), _f32(projmatrix, "projmatrix"), _f32(background, "bg")
        M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0

        out_color = torch.empty(NUM_CHANNELS, H, W, **fopt)
        out_depth = torch.empty(1, H, W, **fopt)
        out_opacity = torch.empty(1, H, W, **fopt)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        n_touched = torch.zeros(P, dtype=torch.int32, device=dev)
        geom = torch.empty(lib.gsr_geom_bytes(P, W, H), dtype=torch.uint8, device=dev)
        img = torch.empty(lib.gsr_image_bytes(W, H), dtype=torch.uint8, device=dev)

        R = C.c_int64(0)
        st = _stream()
        plan_args = (st, P, int(degree), M, W, H)
        in_args = (_ptr(means3D_c), _ptr(sh_c), _ptr(colors_c), _ptr(opacity_c), _ptr(scales_c),
                   float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(view_c), _ptr(proj_c), _ptr(campos_c),
                   float(tan_fovx), float(tan_fovy), int(bool(prefiltered)))
        out_args = (out_color.data_ptr(), out_depth.data_ptr(), out_opacity.data_ptr(), n_touched.data_ptr())
        key = (P, W, H, dev.index)
        cap = _capacity.get(key)
        # the instance-sort workspace exists only where the radix binning runs (images of > 4096 tiles)
        bin_bytes = lib.gsr_binning_bytes if lib.gsr_binning_needed(W, H) else (lambda n: 0)
        done = False
        if cap is not None and not debug:
            # single call, no GPU bubble: the instance-level kernels are enqueued for a capacity remembered from
            # the last call of this shape; the host wait for num_rendered overlaps them
            point_list = torch.empty(lib.gsr_point_list_bytes(cap), dtype=torch.uint8, device=dev)
            binning = torch.empty(bin_bytes(cap), dtype=torch.uint8, device=dev)
            rc = lib.gsr_forward(*plan_args, _ptr(bg_c), *in_args, radii.data_ptr(), geom.data_ptr(), img.data_ptr(),
                                 _ptr(binning), point_list.data_ptr(), cap, *out_args, C.byref(R), 0)
            if rc != _abi.GSR_ECAPACITY:
                _abi.check(rc, "gsr_forward")
                done = True
            else:
                n_touched.zero_()
        else:
            rc = lib.gsr_forward_plan(*plan_args, *in_args, radii.data_ptr(), geom.data_ptr(), C.byref(R),
                                      int(bool(debug)))
            _abi.check(rc, "gsr_forward_plan")
        R = int(R.value)
        if not done:
            point_list = torch.empty(lib.gsr_point_list_bytes(R), dtype=torch.uint8, device=dev)
            binning = torch.empty(bin_bytes(R) if R > 0 else 0, dtype=torch.uint8, device=dev)
            rc = lib.gsr_forward_render(st, P, W, H, _ptr(bg_c), radii.data_ptr(), geom.data_ptr(), img.data_ptr(),
                                        _ptr(binning), point_list.data_ptr(), R, *out_args, int(bool(debug)))
            _abi.check(rc, "gsr_forward_render")
        _capacity[key] = int(R * 1.25) + 65536
        if debug:
            rasterize_gaussians.last_binning = binning  # test hook: sorted keys live in here
    return R, out_color, radii, geom, point_list, img, out_depth, out_opacity, n_touched


def _backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                   projmatrix, projmatrix_raw, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth, sh, degree, campos,
                   geomBuffer, R, binningBuffer, imageBuffer, debug, want, per_gaussian_tau):
    lib = _abi.load()
    P = int(means3D.size(0))
    dev = means3D.device
    fopt = dict(dtype=torch.float32, device=dev)
    M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
    H, W = None, None
    for g in (dL_dout_color, dL_dout_depth):
        if g is not None:
            H, W = int(g.size(1)), int(g.size(2))
    if H is None:
        raise RuntimeError("backward needs at least one of dL_dout_color / dL_dout_depth")
    names = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")
    shapes = dict(means2D=(P, 3), colors=(P, NUM_CHANNELS), opacity=(P, 1), means3D=(P, 3), cov3D=(P, 6),
                  sh=(P, M, 3), scales=(P, 3), rotations=(P, 4))
    out = {}
    for n in names:
        if not want.get(n, True):
            out[n] = None
        elif P == 0 or (n == "sh" and M == 0):
            out[n] = torch.zeros(*shapes[n], **fopt)
        elif n in ("scales", "rotations") and (scales is None or scales.numel() == 0):
            out[n] = torch.zeros(*shapes[n], **fopt)  # cov3D_precomp path: reference leaves them zero
        else:
            out[n] = torch.empty(*shapes[n], **fopt)
    tau_pg = (torch.zeros(P, 6, **fopt) if P == 0 else torch.empty(P, 6, **fopt)) if per_gaussian_tau else None
    tau = torch.zeros(6, **fopt)
    if P == 0:
        return out, tau_pg, tau
    with _on_device(dev):
        scratch = torch.empty(lib.gsr_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
        gc, gd = _f32(dL_dout_color, "dL_dout_color"), _f32(dL_dout_depth, "dL_dout_depth")
        args = [_f32(background, "bg"), _f32(means3D, "means3D"), _f32(sh, "sh"), _f32(colors, "colors"),
                _f32(scales, "scales")]
        rot_c, cov_c = _f32(rotations, "rotations"), _f32(cov3D_precomp, "cov3D_precomp")
        cam = [_f32(viewmatrix, "viewmatrix"), _f32(projmatrix, "projmatrix"), _f32(projmatrix_raw, "projmatrix_raw"),
               _f32(campos, "campos")]
        radii_c = radii.contiguous()
        rc = lib.gsr_backward(
            _stream(), P, int(degree), M, W, H, _ptr(args[0]), _ptr(args[1]), _ptr(args[2]), _ptr(args[3]),
            _ptr(args[4]), float(scale_modifier), _ptr(rot_c), _ptr(cov_c), _ptr(cam[0]), _ptr(cam[1]), _ptr(cam[2]),
            _ptr(cam[3]), float(tan_fovx), float(tan_fovy), radii_c.data_ptr(), geomBuffer.data_ptr(),
            imageBuffer.data_ptr(), _ptr(binningBuffer), int(R), _ptr(gc), _ptr(gd), scratch.data_ptr(),
            _ptr(out["means2D"]), _ptr(out["colors"]), _ptr(out["opacity"]), _ptr(out["means3D"]),
            _ptr(out["cov3D"]), _ptr(out["sh"]), _ptr(out["scales"]), _ptr(out["rotations"]), _ptr(tau_pg),
            tau.data_ptr(), int(bool(debug)))
        _abi.check(rc, "gsr_backward")
    return out, tau_pg, tau


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_depth, sh, degree, campos, geomBuffer, R, binningBuffer,
                                 imageBuffer, debug):
    """Reference-compatible raw backward (rasterize_points.cu:124-211):
    -> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, dL_dtau[P,6])"""
    out, tau_pg, _ = _backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                    cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy,
                                    dL_dout_color, dL_dout_depth, sh, degree, campos, geomBuffer, R, binningBuffer,
                                    imageBuffer, debug, {}, True)
    return (out["means2D"], out["colors"], out["opacity"], out["means3D"], out["cov3D"], out["sh"], out["scales"],
            out["rotations"], tau_pg)


def rasterize_gaussians_backward_fused(*args, want=None):
    """Same inputs as rasterize_gaussians_backward; returns (dict of wanted grads, dL_dtau[6]) with the
    pose gradient already summed over Gaussians in-kernel (replaces torch.sum at __init__.py:152)."""
    out, _, tau = _backward_impl(*args, want or {}, False)
    return out, tau


def debug_sorted_keys(geom, binning, point_list, P, R, W, H):
    """Test hook: the reference's sorted 64-bit keys (tile << 32 | depth bits) as an int64 tensor [R]."""
    lib = _abi.load()
    keys = torch.empty(max(int(R), 0), dtype=torch.int64, device=geom.device)
    if R > 0:
        with _on_device(geom.device):
            rc = lib.gsr_debug_sorted_keys(_stream(), geom.data_ptr(), _ptr(binning), point_list.data_ptr(), int(P),
                                           int(R), int(W), int(H), keys.data_ptr())
            _abi.check(rc, "gsr_debug_sorted_keys")
    return keys


def mark_visible(means3D, viewmatrix, projmatrix):
    """bool[P] = p_view.z > 0.2  (rasterize_points.cu:213-232)"""
    lib = _abi.load()
    P = int(means3D.size(0))
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    if P != 0:
        with _on_device(means3D.device):
            rc = lib.gsr_mark_visible(_stream(), P, _f32(means3D, "means3D").data_ptr(),
                                      _f32(viewmatrix, "viewmatrix").data_ptr(),
                                      _f32(projmatrix, "projmatrix").data_ptr(), present.data_ptr())
            _abi.check(rc, "gsr_mark_visible")
    return present


def rasterize_gaussians_backward_rows(background, means3D, radii, scales, rotations, scale_modifier, viewmatrix,
                                      projmatrix, projmatrix_raw, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth,
                                      geomBuffer, R, point_list, imageBuffer, rows, accumulate, chunks=None,
                                      chunk_done=None, want_means2D=False, debug=False, marks=None, rows_gate=None,
                                      l1_spec=None, image_hw=None, touched_only=False):
    """Backward of one keyframe of the sharded mapping window (gsr_backward_render + gsr_backward_rows): the
    Gaussian gradients are written / added into `rows` f32[P,16] (layout: include/gsr_b200.h) range by range;
    `chunk_done(p0, p1)` is called right after each range has been enqueued (the caller starts its allreduce of
    rows[p0:p1] there). `marks` = (touched_words_ptr, vis_u8_ptr): this rank's slot of the sparse window exchange
    (gsr_backward_rows_marked, monogs_b200.window.RowExchange.marks). `rows_gate`: a CUDA event the current stream
    waits for before the per-Gaussian half (keyframes of one window on several streams accumulate in order).
    `touched_only`: `rows` was zero-filled by the caller (or holds earlier keyframes' sums) and only the rows of
    Gaussians with a non-zero gradient are written / added (GSR_ROWS_TOUCHED_ONLY); column 15 is then not maintained.
    Returns (dL_dtau[6], dL_dmeans2D[P,3] or None)."""
    lib = _abi.load()
    P = int(means3D.size(0))
    dev = means3D.device
    if l1_spec is not None:   # fused L1 loss: dL/dpixel is formed inside the tile renderer (gsr_backward_render_l1)
        H, W = int(image_hw[0]), int(image_hw[1])
    else:
        H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    if rows.shape != (P, 16) or rows.dtype != torch.float32 or not rows.is_contiguous():
        raise RuntimeError("rows must be a contiguous float32 [P,16] tensor")
    tau = torch.zeros(6, dtype=torch.float32, device=dev)
    m2d = torch.empty(P, 3, dtype=torch.float32, device=dev) if want_means2D else None
    if P == 0:
        return tau, m2d
    chunks = chunks or [(0, P)]
    with _on_device(dev):
        scratch = torch.empty(lib.gsr_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
        gc = gd = None
        if l1_spec is None:
            gc, gd = _f32(dL_dout_color, "dL_dout_color"), _f32(dL_dout_depth, "dL_dout_depth")
        bg_c, m3, sc, ro = _f32(background, "bg"), _f32(means3D, "means3D"), _f32(scales, "scales"), _f32(rotations, "rotations")
        vm, pm, pr = _f32(viewmatrix, "viewmatrix"), _f32(projmatrix, "projmatrix"), _f32(projmatrix_raw, "projmatrix_raw")
        radii_c = radii.contiguous()
        st = _stream()
        if l1_spec is not None:
            rc = lib.gsr_backward_render_l1(st, P, W, H, _ptr(bg_c), geomBuffer.data_ptr(), imageBuffer.data_ptr(),
                                            _ptr(point_list), int(R), C.byref(l1_spec), scratch.data_ptr(),
                                            int(bool(debug)))
        else:
            rc = lib.gsr_backward_render(st, P, W, H, _ptr(bg_c), geomBuffer.data_ptr(), imageBuffer.data_ptr(),
                                         _ptr(point_list), int(R), _ptr(gc), _ptr(gd), scratch.data_ptr(),
                                         int(bool(debug)))
        _abi.check(rc, "gsr_backward_render")
        if rows_gate is not None:  # multi-stream window: the row accumulation is ordered across keyframes
            torch.cuda.current_stream(dev).wait_event(rows_gate)
        for (p0, p1) in chunks:
            rc = lib.gsr_backward_rows_marked(st, P, int(p0), int(p1), W, H, _ptr(m3), _ptr(sc),
                                              float(scale_modifier), _ptr(ro), _ptr(vm), _ptr(pm), _ptr(pr),
                                              float(tan_fovx), float(tan_fovy), radii_c.data_ptr(),
                                              geomBuffer.data_ptr(), scratch.data_ptr(), rows.data_ptr(),
                                              int(bool(accumulate)) | (2 if touched_only else 0), _ptr(m2d),
                                              tau.data_ptr(),
                                              marks[0] if marks else None, marks[1] if marks else None,
                                              int(bool(debug)))
            _abi.check(rc, "gsr_backward_rows_marked")
            if chunk_done is not None:
                chunk_done(p0, p1)
    return tau, m2d
