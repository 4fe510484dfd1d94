"""One tracking iteration as ONE CUDA graph (additive entry point; the reference API is untouched).

MonoGS's tracker runs, per frame, up to 100 iterations of
    render(viewpoint) -> get_loss_tracking -> loss.backward() -> pose optimiser step -> update_pose
(utils/slam_tracker.py:132-188) against a FIXED map.  At tracking sizes (config C2: 100 k Gaussians, 640x480) the
GPU work of an iteration is ~0.4 ms in 20 short kernels, so the loop is bound by host launch work, not by the
kernels.  `TrackingGraph` captures the whole device side of an iteration --

    gsr_camera_setup (R, T -> view / projection / camera centre)          utils/camera_utils.py:171-231
    gsr_forward_enqueue (the forward for a fixed list capacity, no host read)
    gsr_l1_loss_forward (get_loss_tracking, utils/slam_utils.py:58-98)
    gsr_backward_l1 (pose gradient dL/dtau, exposure gradients; Gaussian gradients only on request)

-- into one `torch.cuda.CUDAGraph` over static buffers and replays it per iteration: one launch instead of ~20 kernel
launches and ~15 framework ops.  The pose (R, T) and the exposure parameters are copied into the static inputs by
`step`; the map tensors are read in place (they do not change while tracking).  Results are the same kernels'
results as `fused_loss.render_l1_loss(kind="tracking")` + `backward()`.

The instance-list capacity is fixed at capture time (1.5 x the count of the first pose + slack).  A replay whose
`num_rendered` exceeds it writes nothing out of bounds, but its images are incomplete: the graph itself (device side,
no host read) then turns the loss `terms[0]` and the pose gradient `tau` of that replay into NaN, so a caller that never
asks `overflowed()` cannot silently optimise on a truncated render.  `overflowed()` (one host read of the pinned
count) says so explicitly; `recapture()` sizes the lists for the current pose and captures again.

The map tensors are captured BY ADDRESS: they must be contiguous fp32 CUDA tensors (a hidden converted copy would be
frozen into the graph and the caller's later in-place updates never seen) -- anything else is refused.  After an
in-place change of `scales` of an isotropic map ([P, 1], expanded here to [P, 3]) call `refresh_scales()`.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _abi
from .diff_gaussian_rasterization import _C
from .render import camera_constants


class TrackingGraph:
    def __init__(self, cam_intrinsics, means, rotations, scales, opacity, features, bg_color, gt_rgb, gt_depth=None,
                 mask=None, grad_mask=None, R0: Optional[torch.Tensor] = None, T0: Optional[torch.Tensor] = None,
                 want_gaussian_grads: bool = False, capacity_factor: float = 1.5, capacity_slack: int = 65536):
        lib = self.lib = _abi.load()
        if not means.is_cuda:
            raise RuntimeError("monogs_b200.tracking: CUDA tensors required (there is no CPU path)")
        dev = self.dev = means.device
        f = lambda t, n: _C._f32(t, n)
        self.P = P = int(means.shape[0])
        if P == 0:
            raise RuntimeError("empty map")
        self.W, self.H = int(cam_intrinsics.width), int(cam_intrinsics.height)
        W, H = self.W, self.H
        self.tanfovx, self.tanfovy, proj_T = camera_constants(cam_intrinsics)
        self.proj_T = proj_T.detach().contiguous()
        # the map: read in place (static storage for the lifetime of the graph)
        def static(t, n):
            t = t.detach()
            if t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
                raise RuntimeError(f"monogs_b200.tracking: `{n}` must be a contiguous float32 tensor on {dev} "
                                   "(the graph captures its address; a converted copy would go stale)")
            return t
        self.means, self.rots, self.opac, self.feat = (static(means, "means"), static(rotations, "rotations"),
                                                       static(opacity, "opacity"), static(features, "features"))
        self._scales_src = static(scales, "scales")
        self.scales = self._scales_src.expand(-1, 3).contiguous() if self._scales_src.shape[-1] == 1 else self._scales_src
        self.bg = f(bg_color, "bg")
        # per-frame targets (set_frame overwrites them in place)
        self.gt_rgb = f(gt_rgb, "gt_rgb").clone()
        self.gt_depth = f(gt_depth, "gt_depth").clone().view(1, H, W) if gt_depth is not None else None
        u8 = lambda m: None if m is None else (m.reshape(H, W) != 0).to(torch.uint8).contiguous()
        self.mask, self.grad_mask = u8(mask), u8(grad_mask)
        # static inputs updated by step()
        self.R = torch.eye(3, device=dev) if R0 is None else f(R0, "R").clone()
        self.T = torch.zeros(3, device=dev) if T0 is None else f(T0, "T").clone()
        self.exposure = torch.zeros(2, device=dev)
        # static outputs
        fo = dict(dtype=torch.float32, device=dev)
        self.cam = torch.empty(35, **fo)                   # view 16 | proj 16 | campos 3
        self.color, self.depth, self.opacity = torch.empty(3, H, W, **fo), torch.empty(1, H, W, **fo), torch.empty(1, H, W, **fo)
        self.radii = torch.empty(P, dtype=torch.int32, device=dev)
        self.n_touched = torch.zeros(P, dtype=torch.int32, device=dev)
        self.terms = torch.zeros(8, **fo)                  # loss, l1_rgb, l1_depth, k_rgb, k_depth, dL/da, dL/db, mean opacity
        self.tau = torch.zeros(6, **fo)                    # dL/d(rho | theta)
        self.grads = None
        if want_gaussian_grads:
            self.grads = {"means3D": torch.empty(P, 3, **fo), "means2D": torch.empty(P, 3, **fo),
                          "colors": torch.empty(P, 3, **fo), "opacity": torch.empty(P, 1, **fo),
                          "scales": torch.empty(P, 3, **fo), "rotations": torch.empty(P, 4, **fo)}
        ws = lambda n: torch.empty(max(int(n), 256), dtype=torch.uint8, device=dev)
        self.geom, self.img = ws(lib.gsr_geom_bytes(P, W, H)), ws(lib.gsr_image_bytes(W, H))
        self.scratch = ws(lib.gsr_backward_scratch_bytes(P))
        self.acc = torch.empty(8, dtype=torch.float64, device=dev)
        self.up = torch.ones(1, **fo)
        self.total = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.total_dev = torch.zeros(1, dtype=torch.int64, device=dev)   # the replay's num_rendered, device side
        self.over = torch.zeros(1, dtype=torch.bool, device=dev)
        self.capacity_factor, self.capacity_slack = float(capacity_factor), int(capacity_slack)
        self.graph = None
        self._size_and_capture()

    def _size_and_capture(self):
        """Capacity from the current pose (one synchronising plan call, outside the graph), buffers, capture."""
        lib, dev, P, W, H = self.lib, self.dev, self.P, self.W, self.H
        ws = lambda n: torch.empty(max(int(n), 256), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            self._camera_setup()
            R = C.c_int64(0)
            rc = lib.gsr_forward_plan(self._st(), P, 0, 0, W, H, self.means.data_ptr(), None, self.feat.data_ptr(),
                                      self.opac.data_ptr(), self.scales.data_ptr(), 1.0, self.rots.data_ptr(), None,
                                      self.cam.data_ptr(), self.cam.data_ptr() + 64, self.cam.data_ptr() + 128,
                                      self.tanfovx, self.tanfovy, 0, self.radii.data_ptr(), self.geom.data_ptr(),
                                      C.byref(R), 0)
            _abi.check(rc, "gsr_forward_plan")
        self.capacity = max(int(R.value * self.capacity_factor) + self.capacity_slack, 1)
        self.plist = ws(lib.gsr_point_list_bytes(self.capacity))
        self.binning = ws(lib.gsr_binning_bytes(self.capacity)) if lib.gsr_binning_needed(W, H) else None
        self.spec = _abi.L1LossSpec(_abi.LOSS_TRACKING, self.color.data_ptr(), self.depth.data_ptr(),
                                    self.opacity.data_ptr(), self.gt_rgb.data_ptr(), _C._ptr(self.gt_depth),
                                    _C._ptr(self.mask), _C._ptr(self.grad_mask), self.exposure.data_ptr(),
                                    self.exposure.data_ptr() + 4, self.terms.data_ptr(), self.up.data_ptr())
        self.graph = None
        self._capture()

    def refresh_scales(self):
        """Re-expand an isotropic map's [P, 1] scales into the captured [P, 3] buffer (after an in-place update)."""
        if self.scales is not self._scales_src:
            self.scales.copy_(self._scales_src.expand(-1, 3))

    def recapture(self):
        """Size the instance lists for the CURRENT pose (self.R, self.T) and capture the graph again."""
        torch.cuda.synchronize(self.dev)
        self.graph = None
        self._size_and_capture()

    # ------------------------------------------------------------------------------------------------------
    def _st(self):
        return C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)

    def _camera_setup(self):
        rc = self.lib.gsr_camera_setup(self._st(), self.R.data_ptr(), 3, 1, self.T.data_ptr(), 1,
                                       self.proj_T.data_ptr(), self.cam.data_ptr(), self.cam.data_ptr() + 64,
                                       self.cam.data_ptr() + 128)
        _abi.check(rc, "gsr_camera_setup")

    def _enqueue(self):
        """The device side of one iteration on the current stream (what the graph holds)."""
        lib, P, W, H, cam = self.lib, self.P, self.W, self.H, self.cam.data_ptr()
        self._camera_setup()
        self.n_touched.zero_()
        rc = lib.gsr_forward_enqueue(self._st(), P, 0, 0, W, H, self.bg.data_ptr(), self.means.data_ptr(), None,
                                     self.feat.data_ptr(), self.opac.data_ptr(), self.scales.data_ptr(), 1.0,
                                     self.rots.data_ptr(), None, cam, cam + 64, cam + 128, self.tanfovx, self.tanfovy,
                                     0, self.radii.data_ptr(), self.geom.data_ptr(), self.img.data_ptr(),
                                     _C._ptr(self.binning), self.plist.data_ptr(), self.capacity,
                                     self.color.data_ptr(), self.depth.data_ptr(), self.opacity.data_ptr(),
                                     self.n_touched.data_ptr(), self.total_dev.data_ptr())
        _abi.check(rc, "gsr_forward_enqueue")
        self.total.copy_(self.total_dev, non_blocking=True)
        rc = lib.gsr_l1_loss_forward(self._st(), _abi.LOSS_TRACKING, W, H, self.color.data_ptr(),
                                     self.depth.data_ptr(), self.opacity.data_ptr(), self.gt_rgb.data_ptr(),
                                     _C._ptr(self.gt_depth), _C._ptr(self.mask), _C._ptr(self.grad_mask),
                                     self.exposure.data_ptr(), self.exposure.data_ptr() + 4, 0.9,
                                     self.acc.data_ptr(), self.terms.data_ptr())
        _abi.check(rc, "gsr_l1_loss_forward")
        g = self.grads or {}
        p = lambda n: g[n].data_ptr() if n in g else None
        rc = lib.gsr_backward_l1(self._st(), P, 0, 0, W, H, self.bg.data_ptr(), self.means.data_ptr(), None,
                                 self.feat.data_ptr(), self.scales.data_ptr(), 1.0, self.rots.data_ptr(), None, cam,
                                 cam + 64, self.proj_T.data_ptr(), cam + 128, self.tanfovx, self.tanfovy,
                                 self.radii.data_ptr(), self.geom.data_ptr(), self.img.data_ptr(),
                                 self.plist.data_ptr(), self.capacity, C.byref(self.spec), self.scratch.data_ptr(),
                                 p("means2D"), p("colors"), p("opacity"), p("means3D"), None, None, p("scales"),
                                 p("rotations"), None, self.tau.data_ptr(), 0)
        _abi.check(rc, "gsr_backward_l1")
        # a replay beyond the captured capacity rendered a truncated list: poison its loss and pose gradient
        torch.gt(self.total_dev, self.capacity, out=self.over)
        self.terms[0:1].masked_fill_(self.over, float("nan"))
        self.tau.masked_fill_(self.over, float("nan"))

    def _capture(self):
        with torch.cuda.device(self.dev):
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(torch.cuda.current_stream(self.dev))
            with torch.cuda.stream(side):
                for _ in range(2):           # warm-up outside the capture (lazy function attributes, caches)
                    self._enqueue()
            torch.cuda.current_stream(self.dev).wait_stream(side)
            torch.cuda.synchronize(self.dev)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._enqueue()

    # ------------------------------------------------------------------------------------------------------
    def set_frame(self, gt_rgb, gt_depth=None, mask=None, grad_mask=None):
        """New target images for the same map / image size (copied into the static buffers)."""
        self.gt_rgb.copy_(gt_rgb)
        if self.gt_depth is not None and gt_depth is not None:
            self.gt_depth.copy_(gt_depth.reshape(1, self.H, self.W))
        if self.mask is not None and mask is not None:
            self.mask.copy_((mask.reshape(self.H, self.W) != 0).to(torch.uint8))
        if self.grad_mask is not None and grad_mask is not None:
            self.grad_mask.copy_((grad_mask.reshape(self.H, self.W) != 0).to(torch.uint8))

    def step(self, R: torch.Tensor, T: torch.Tensor, exposure_a=None, exposure_b=None):
        """Replay one iteration for the pose (R, T). Returns (loss, grad_rho[3], grad_theta[3], terms[8]) as views
        of the static outputs (device tensors, valid until the next step); dL/d exposure_a, b = terms[5], terms[6]."""
        self.R.copy_(R, non_blocking=True)
        self.T.copy_(T.reshape(3), non_blocking=True)
        if exposure_a is not None:
            self.exposure[0:1].copy_(exposure_a.detach().reshape(1), non_blocking=True)
        if exposure_b is not None:
            self.exposure[1:2].copy_(exposure_b.detach().reshape(1), non_blocking=True)
        self.graph.replay()
        return self.terms[0], self.tau[:3], self.tau[3:], self.terms

    def overflowed(self) -> bool:
        """True when the last completed replay had more instances than the captured capacity (synchronises)."""
        torch.cuda.current_stream(self.dev).synchronize()
        return int(self.total[0]) > self.capacity
