#!/usr/bin/env python
"""bench.py -- rasterize fwd+bwd iterations/s (BASELINE.json's metric) on N B200s of one node.

  python bench.py [--gpus N --steps K --warmup W] [--impl ours|reference] [--workload headline|c2|c3|c4|c5]

A "step" is one pass of the hot path over one keyframe: rasterizer forward (preprocess, depth sort, instance
emit, tile sort, ranges, render) + backward (render backward, fused preprocess backward incl. the SE(3) pose
gradient) of the synthetic 640x480 / 1 M-Gaussian scene (`headline`, the config the metric is quoted on).
At N > 1 the step is the sharded MAPPING WINDOW: N keyframes of the same Gaussian set, one per
GPU, with the Gaussian-gradient allreduce (NCCL) inside the timed step -- per-GPU work fixed, scaling "weak";
`value` = keyframe fwd+bwd iterations of all ranks / max-over-ranks device time.

  value        inputs resident in HBM, straight through the C ABI (include/gsr_b200.h), CUDA-event timed
  e2e          the same step through the public API with the Gaussian parameters and targets in PINNED HOST memory:
               H2D copies, L1 loss, backward and the D2H read of loss + pose gradient are inside the timed region.
               N = 1: GaussianRasterizer autograd (what MonoGS's render() calls). N > 1: the repo's own window API,
               monogs_b200.window.MappingWindow.step with the CUDA render backend and the RowExchange (sparse
               in-switch kernel / multimem / NCCL, whichever it selected) -- `e2e.resident` is the same leg with the
               host copies hoisted out, which separates H2D contention from the collective
  window_c4    BASELINE config 4: a FIXED window of 8 keyframes x 2 M Gaussians at 1200x680 through
               MappingWindow.step(params, 8), 8/N keyframes per rank (strong scaling), with the exchange mode
  roofline     the dominant kernel, timed live with CUDA events on the launching stream (gsr_profile_*),
               against MEASURED_PEAKS.json's HBM copy bandwidth, algorithmic bytes per DESIGN.md
  cpu_baseline the CPU oracle port (oracle/cpu_ref.cpp, OpenMP) on the box's host cores, one bounded sample
  --impl reference   the UNMODIFIED reference CUDA rasterizer (oracle/_ref, rebuilt for sm_100a) on the same
               workload and metric; falls back to the CPU oracle port when oracle/_ref did not travel
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

WORKLOADS = {
    # name: (P, camera kind, description)
    "headline": (1_000_000, "tum", "640x480, 1M Gaussians, rasterize fwd+bwd incl. pose grads (metric config)"),
    "c2": (100_000, "tum", "640x480, 100k Gaussians, tracking-step shape"),
    "c3": (500_000, "replica", "1200x680, 500k Gaussians, mapping keyframe"),
    "c4": (2_000_000, "replica", "1200x680, 2M Gaussians, mapping keyframe"),
    "c5": (4_000_000, "hd", "1920x1080, 4M Gaussians, stress"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="headline", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-window-c4", action="store_true")
    return ap.parse_args()


def dist_setup(n_gpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
    return rank, world, local


# ------------------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md "clocks line")
# ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.split(",") for r in Path(self.f.name).read_text().splitlines() if r.count(",") >= 8]
        os.unlink(self.f.name)
        sm, reasons = [], set()
        for r in rows:
            r = [x.strip() for x in r]
            try:
                sm.append(float(r[1]))
                out["sm_max_mhz"] = float(r[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]
        out["reasons"] = sorted(reasons)
        out["samples"] = len(sm)
        return out


# ------------------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------------------
class Workload:
    """Synthetic scene (monogs_b200.synth, seed 0) + one camera per rank on a 30 cm arc + fixed loss gradients."""

    def __init__(self, name, rank, world, dev):
        from monogs_b200 import synth
        self.name = name
        self.P, self.kind, self.desc = WORKLOADS[name]
        cams = synth.arc_cameras(self.kind, max(world, 1)) if world > 1 else [
            synth.make_camera(self.kind, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))]
        self.cam = cams[rank if world > 1 else 0]
        centre = synth.make_camera(self.kind)
        self.host = synth.make_scene(self.P, centre, seed=0, device="cpu")          # same Gaussians on every rank
        self.dev = dev
        self.scene = {k: v.to(dev) for k, v in self.host.items()}
        self.view, self.proj, self.proj_raw, self.campos = self.cam.matrices(dev)
        self.bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
        self.W, self.H = self.cam.W, self.cam.H
        g = torch.Generator().manual_seed(7 + rank)
        self.dcol = (torch.randn(3, self.H, self.W, generator=g) / (self.H * self.W)).to(dev)
        self.ddep = (torch.randn(1, self.H, self.W, generator=g) / (self.H * self.W)).to(dev)
        self.gt_image, self.gt_depth = synth.make_targets(self.cam, seed=rank, device="cpu")

    @property
    def N(self):
        return self.W * self.H


GRAD_COLS = (("means3D", 3), ("colors", 3), ("opacity", 1), ("scales", 3), ("rotations", 4))


class OursDeviceStep:
    """fwd+bwd straight through the C ABI with every buffer resident in HBM and reused across steps."""

    def __init__(self, wl: Workload, world):
        from monogs_b200 import _abi
        self.lib = _abi.load()
        self.wl, self.world = wl, world
        lib, P, W, H, dev = self.lib, wl.P, wl.W, wl.H, wl.dev
        u8 = lambda n: torch.empty(max(int(n), 256), dtype=torch.uint8, device=dev)
        self.radii = torch.empty(P, dtype=torch.int32, device=dev)
        self.n_touched = torch.zeros(P, dtype=torch.int32, device=dev)
        self.geom, self.img = u8(lib.gsr_geom_bytes(P, W, H)), u8(lib.gsr_image_bytes(W, H))
        self.scratch = u8(lib.gsr_backward_scratch_bytes(P))
        self.binning, self.plist, self.cap = None, None, -1
        self.color = torch.empty(3, H, W, device=dev)
        self.depth = torch.empty(1, H, W, device=dev)
        self.opacity = torch.empty(1, H, W, device=dev)
        # one flat gradient buffer (column blocks) = the allreduce payload of the mapping window
        self.flat = torch.empty(sum(n for _, n in GRAD_COLS) * P, device=dev)
        self.g, o = {}, 0
        for name, n in GRAD_COLS:
            self.g[name] = self.flat[o:o + n * P].view(P, n)
            o += n * P
        self.means2D = torch.empty(P, 3, device=dev)
        self.tau = torch.zeros(6, device=dev)
        self.R = 0
        self.rows, self.chunks = None, None
        self.zero_stream, self.zero_done = None, None

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def _grow(self, R):  # grow-only, like the reference's resizable byte tensors
        self.cap = int(R * 1.25) + 1024
        self.binning = torch.empty(self.lib.gsr_binning_bytes(self.cap), dtype=torch.uint8, device=self.wl.dev)
        self.plist = torch.empty(self.lib.gsr_point_list_bytes(self.cap), dtype=torch.uint8, device=self.wl.dev)

    def step(self):
        lib, wl, sc = self.lib, self.wl, self.wl.scene
        P, W, H = wl.P, wl.W, wl.H
        st = self._stream()
        R = C.c_int64(0)
        self.n_touched.zero_()
        self.zero_done = None
        ex = getattr(self, "exchange", None)
        if self.world > 1 and ex is not None and ex.decided and ex.mode == "sparse" and ex.marks is not None:
            if self.zero_stream is None:
                self.zero_stream = torch.cuda.Stream(device=wl.dev)
            self.zero_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.zero_stream):
                self.rows.zero_()
                self.zero_done = torch.cuda.Event()
                self.zero_done.record(self.zero_stream)
        if self.cap < 0:  # first call of this shape: two-phase forward sizes the instance buffers
            rc = lib.gsr_forward_plan(st, P, 0, 0, W, H, sc["means3D"].data_ptr(), None, sc["colors"].data_ptr(),
                                      sc["opacities"].data_ptr(), sc["scales"].data_ptr(), 1.0,
                                      sc["rotations"].data_ptr(), None, wl.view.data_ptr(), wl.proj.data_ptr(),
                                      wl.campos.data_ptr(), wl.cam.tanfovx, wl.cam.tanfovy, 0, self.radii.data_ptr(),
                                      self.geom.data_ptr(), C.byref(R), 0)
            assert rc == 0, rc
            self._grow(int(R.value))
            rc = lib.gsr_forward_render(st, P, W, H, wl.bg.data_ptr(), self.radii.data_ptr(), self.geom.data_ptr(),
                                        self.img.data_ptr(), self.binning.data_ptr(), self.plist.data_ptr(),
                                        int(R.value), self.color.data_ptr(), self.depth.data_ptr(),
                                        self.opacity.data_ptr(), self.n_touched.data_ptr(), 0)
            assert rc == 0, rc
        else:             # steady state: one bubble-free call for the remembered capacity
            rc = lib.gsr_forward(st, P, 0, 0, W, H, wl.bg.data_ptr(), sc["means3D"].data_ptr(), None,
                                 sc["colors"].data_ptr(), sc["opacities"].data_ptr(), sc["scales"].data_ptr(), 1.0,
                                 sc["rotations"].data_ptr(), None, wl.view.data_ptr(), wl.proj.data_ptr(),
                                 wl.campos.data_ptr(), wl.cam.tanfovx, wl.cam.tanfovy, 0, self.radii.data_ptr(),
                                 self.geom.data_ptr(), self.img.data_ptr(), self.binning.data_ptr(),
                                 self.plist.data_ptr(), self.cap, self.color.data_ptr(), self.depth.data_ptr(),
                                 self.opacity.data_ptr(), self.n_touched.data_ptr(), C.byref(R), 0)
            assert rc == 0, rc  # (a GSR_ECAPACITY retry is not needed here: the scene is fixed)
        R = int(R.value)
        if self.world > 1:
            self._backward_window(st, R)
        else:
            g = self.g
            rc = lib.gsr_backward(st, P, 0, 0, W, H, wl.bg.data_ptr(), sc["means3D"].data_ptr(), None,
                                  sc["colors"].data_ptr(), sc["scales"].data_ptr(), 1.0, sc["rotations"].data_ptr(),
                                  None, wl.view.data_ptr(), wl.proj.data_ptr(), wl.proj_raw.data_ptr(),
                                  wl.campos.data_ptr(), wl.cam.tanfovx, wl.cam.tanfovy, self.radii.data_ptr(),
                                  self.geom.data_ptr(), self.img.data_ptr(), self.plist.data_ptr(), R,
                                  wl.dcol.data_ptr(), wl.ddep.data_ptr(), self.scratch.data_ptr(),
                                  self.means2D.data_ptr(), g["colors"].data_ptr(), g["opacity"].data_ptr(),
                                  g["means3D"].data_ptr(), None, None, g["scales"].data_ptr(),
                                  g["rotations"].data_ptr(), None, self.tau.data_ptr(), 0)
            assert rc == 0, rc
        self.R = R

    def _backward_window(self, st, R):
        """Sharded mapping window (monogs_b200/window.py): per-pixel half, then the per-Gaussian half range by
        range into 64-byte window rows; each finished range goes to an async NCCL allreduce (the one exchange step
        of the path) while the next range is computed; the step ends when the last allreduce has completed."""
        from monogs_b200.window import RowExchange, window_chunks
        lib, wl, sc = self.lib, self.wl, self.wl.scene
        P, W, H = wl.P, wl.W, wl.H
        if self.rows is None:
            self.exchange = RowExchange()
            self.rows = self.exchange.allocate(P, wl.dev)   # symmetric memory when NVLS multicast is available
            self.chunks = window_chunks(P, 1)
        rc = lib.gsr_backward_render(st, P, W, H, wl.bg.data_ptr(), self.geom.data_ptr(), self.img.data_ptr(),
                                     self.plist.data_ptr(), R, wl.dcol.data_ptr(), wl.ddep.data_ptr(),
                                     self.scratch.data_ptr(), 0)
        assert rc == 0, rc
        # sparse exchange decided: rows zero-filled on a side stream during the forward (step()), only touched rows
        # written by the per-Gaussian backward (GSR_ROWS_TOUCHED_ONLY) -- what MappingWindow.step does
        flags = 0
        if self.zero_done is not None:
            torch.cuda.current_stream().wait_event(self.zero_done)
            flags = 2
        works = []
        for (p0, p1) in self.chunks:
            mk = self.exchange.marks or (None, None)
            rc = lib.gsr_backward_rows_marked(st, P, p0, p1, W, H, sc["means3D"].data_ptr(),
                                              sc["scales"].data_ptr(), 1.0, sc["rotations"].data_ptr(),
                                              wl.view.data_ptr(), wl.proj.data_ptr(), wl.proj_raw.data_ptr(),
                                              wl.cam.tanfovx, wl.cam.tanfovy, self.radii.data_ptr(),
                                              self.geom.data_ptr(), self.scratch.data_ptr(), self.rows.data_ptr(),
                                              flags, self.means2D.data_ptr(), self.tau.data_ptr(), mk[0], mk[1], 0)
            assert rc == 0, rc
            w = self.exchange.reduce(self.rows, p0, p1)
            if w is not None:
                works.append(w)
        for w in works:
            w.wait()  # the current stream waits for the exchange: it is part of the timed step

    def units(self):
        """P, V, R, consumed instances (sum over tiles of the largest n_contrib) and N of the last step."""
        wl = self.wl
        nc_p, ft_p, rg_p = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self.lib.gsr_debug_image_views(self.img.data_ptr(), wl.W, wl.H, C.byref(ft_p), C.byref(nc_p), C.byref(rg_p))
        off = nc_p.value - self.img.data_ptr()
        nc = self.img[off:off + 4 * wl.N].view(torch.int32).view(wl.H, wl.W).float()
        tiles = torch.nn.functional.max_pool2d(nc[None, None], 16, ceil_mode=True)
        return dict(P=wl.P, V=int((self.radii > 0).sum()), R=int(self.R), R_consumed=int(tiles.sum()), N=wl.N)


class RefDeviceStep:
    """The unmodified reference CUDA rasterizer (oracle/_ref) on the same inputs, same outputs."""

    def __init__(self, wl: Workload, world):
        from oracle import ref_cuda
        self.ref, self.wl, self.world = ref_cuda, wl, world
        ref_cuda.load()
        self.R = 0
        self.radii = None

    def step(self):
        wl, sc, ref = self.wl, self.wl.scene, self.ref
        e = torch.empty(0, device=wl.dev)
        R, col, radii, _, _, _, dep, opa, nt = ref.rasterize_gaussians(
            wl.bg, sc["means3D"], sc["colors"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, wl.view,
            wl.proj, wl.proj_raw, wl.cam.tanfovx, wl.cam.tanfovy, wl.H, wl.W, e, 0, wl.campos, False, False)
        out = ref.rasterize_gaussians_backward(
            wl.bg, sc["means3D"], radii, sc["colors"], sc["scales"], sc["rotations"], 1.0, e, wl.view, wl.proj,
            wl.proj_raw, wl.cam.tanfovx, wl.cam.tanfovy, wl.dcol, wl.ddep, e, 0, wl.campos, None, R, None, None, False)
        tau = out[8].sum(0)  # diff_gaussian_rasterization/__init__.py:152-154
        if self.world > 1:
            flat = torch.cat([out[3].reshape(-1), out[1].reshape(-1), out[2].reshape(-1), out[6].reshape(-1),
                              out[7].reshape(-1)])
            dist.all_reduce(flat)
        self.R, self.radii, self.tau = R, radii, tau

    def units(self):
        wl = self.wl
        return dict(P=wl.P, V=int((self.radii > 0).sum()), R=int(self.R), R_consumed=None, N=wl.N)


# ------------------------------------------------------------------------------------------------------------
# end to end through the public API, inputs in pinned host memory
# ------------------------------------------------------------------------------------------------------------
class _RefFunction(torch.autograd.Function):
    """Autograd node over oracle/_ref with the data flow of the reference's own Python wrapper
    (diff_gaussian_rasterization/__init__.py:48-171): forward, backward, torch.sum of dL_dtau."""

    @staticmethod
    def forward(ctx, ref, wl, means3D, means2D, colors, opacities, scales, rotations, theta, rho):
        e = torch.empty(0, device=means3D.device)
        R, col, radii, _, _, _, dep, opa, nt = ref.rasterize_gaussians(
            wl.bg, means3D, colors, opacities, scales, rotations, 1.0, e, wl.view, wl.proj, wl.proj_raw,
            wl.cam.tanfovx, wl.cam.tanfovy, wl.H, wl.W, e, 0, wl.campos, False, False)
        ctx.ref, ctx.wl, ctx.R = ref, wl, R
        ctx.save_for_backward(means3D, colors, scales, rotations, radii)
        ctx.mark_non_differentiable(radii, nt)
        return col, radii, dep, opa, nt

    @staticmethod
    def backward(ctx, g_col, g_radii, g_dep, g_opa, g_nt):
        means3D, colors, scales, rotations, radii = ctx.saved_tensors
        wl = ctx.wl
        e = torch.empty(0, device=means3D.device)
        if g_dep is None:
            g_dep = torch.zeros(1, wl.H, wl.W, device=means3D.device)
        o = ctx.ref.rasterize_gaussians_backward(
            wl.bg, means3D, radii, colors, scales, rotations, 1.0, e, wl.view, wl.proj, wl.proj_raw, wl.cam.tanfovx,
            wl.cam.tanfovy, g_col.contiguous(), g_dep.contiguous(), e, 0, wl.campos, None, ctx.R, None, None, False)
        tau = torch.sum(o[8].view(-1, 6), dim=0)
        return None, None, o[3], o[0], o[1], o[2], o[6], o[7], tau[3:].view(1, -1), tau[:3].view(1, -1)


class E2EStep:
    """One mapping-shaped iteration the way a MonoGS caller does it: parameters arrive from PINNED HOST memory,
    render through the public API, the mapping loss 0.9 L1(rgb) + 0.1 L1(depth over gt_depth > 0) (utils/slam_utils.py:101-146),
    backward, and the loss + pose gradient are read back to the host."""

    NAMES = ("means3D", "colors", "opacities", "scales", "rotations")
    SLOTS = 3

    def __init__(self, wl: Workload, world, impl):
        self.wl, self.world, self.impl = wl, world, impl
        self.pinned = {k: wl.host[k].contiguous().pin_memory() for k in self.NAMES}
        self.pinned_img = wl.gt_image.contiguous().pin_memory()
        self.pinned_dep = wl.gt_depth.contiguous().pin_memory()
        # triple-buffered device staging: the H2D copies of the next two steps' inputs run on a copy stream while
        # step i computes, so the copy engine never waits for the host (every step still copies its own 61 MB from
        # pinned host memory inside the timed region)
        S = self.SLOTS
        self.dev_buf = [{k: torch.empty_like(wl.host[k], device=wl.dev) for k in self.NAMES} for _ in range(S)]
        self.dev_img = [torch.empty_like(wl.gt_image, device=wl.dev) for _ in range(S)]
        self.dev_dep = [torch.empty_like(wl.gt_depth, device=wl.dev) for _ in range(S)]
        self.copy_stream = torch.cuda.Stream(device=wl.dev)
        self.ready = [torch.cuda.Event() for _ in range(S)]
        self.consumed = [torch.cuda.Event() for _ in range(S)]
        self.i = 0
        self._prefetch(0)
        self._prefetch(1)
        self.host_out = torch.empty(7, dtype=torch.float32).pin_memory()
        self.h2d = sum(t.numel() * 4 for t in self.pinned.values()) + (self.pinned_img.numel() + self.pinned_dep.numel()) * 4
        self.d2h = 7 * 4
        if impl in ("ours", "ours-fused"):
            from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
            s = GaussianRasterizationSettings(wl.H, wl.W, wl.cam.tanfovx, wl.cam.tanfovy, wl.bg, 1.0, wl.view, wl.proj,
                                              wl.proj_raw, 0, wl.campos, False, False)
            self.rasterizer = GaussianRasterizer(s)
        else:
            from oracle import ref_cuda
            self.ref = ref_cuda

    def _prefetch(self, slot):
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.consumed[slot])  # do not overwrite inputs still being read
            for k in self.NAMES:
                self.dev_buf[slot][k].copy_(self.pinned[k], non_blocking=True)
            self.dev_img[slot].copy_(self.pinned_img, non_blocking=True)
            self.dev_dep[slot].copy_(self.pinned_dep, non_blocking=True)
            self.ready[slot].record(self.copy_stream)

    def step(self):
        wl = self.wl
        slot = self.i % self.SLOTS
        self.i += 1
        torch.cuda.current_stream().wait_event(self.ready[slot])
        self._prefetch((slot + 2) % self.SLOTS)               # inputs of step i+2, overlapped with steps i, i+1
        leaves = {k: self.dev_buf[slot][k].detach().requires_grad_(True) for k in self.NAMES}
        dev_img, dev_dep = self.dev_img[slot], self.dev_dep[slot]
        means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
        theta = torch.zeros(3, device=wl.dev, requires_grad=True)
        rho = torch.zeros(3, device=wl.dev, requires_grad=True)
        if self.impl == "ours-fused":
            from monogs_b200.fused_loss import render_l1_loss
            loss = render_l1_loss(self.rasterizer.raster_settings, leaves["means3D"], means2D, leaves["opacities"],
                                  colors_precomp=leaves["colors"], scales=leaves["scales"],
                                  rotations=leaves["rotations"], theta=theta, rho=rho, gt_rgb=dev_img,
                                  gt_depth=dev_dep, kind="mapping")[0]
        elif self.impl == "ours":
            color, radii, depth, opacity, n_touched = self.rasterizer(
                means3D=leaves["means3D"], means2D=means2D, shs=None, colors_precomp=leaves["colors"],
                opacities=leaves["opacities"], scales=leaves["scales"], rotations=leaves["rotations"],
                cov3D_precomp=None, theta=theta, rho=rho)
        else:
            color, radii, depth, opacity, n_touched = _RefFunction.apply(
                self.ref, wl, leaves["means3D"], means2D, leaves["colors"], leaves["opacities"], leaves["scales"],
                leaves["rotations"], theta, rho)
        if self.impl != "ours-fused":
            # get_loss_mapping (utils/slam_utils.py:101-146): 0.9 L1(rgb) + 0.1 L1(depth over gt_depth > 0)
            valid = (dev_dep > 0).float()
            loss = 0.9 * (color - dev_img).abs().mean() + \
                0.1 * ((depth - dev_dep).abs() * valid).sum() / valid.sum().clamp_min(1.0)
        loss.backward()
        if self.world > 1:
            flat = torch.cat([leaves[k].grad.reshape(-1) for k in self.NAMES])
            dist.all_reduce(flat)
        out = torch.cat([loss.detach().view(1), rho.grad.view(-1), theta.grad.view(-1)])
        self.host_out.copy_(out, non_blocking=True)
        self.consumed[slot].record()
        torch.cuda.current_stream().synchronize()  # the caller reads the loss / pose gradient on the host
        return float(self.host_out[0])


class _SlotTargets:
    """targets[k] of make_cuda_render_fn -> the (image, depth) of the staging slot currently being consumed."""

    def __init__(self, owner):
        self.owner = owner

    def __getitem__(self, k):
        o = self.owner
        return o.dev_img[o.cur_slot], o.dev_dep[o.cur_slot]


class E2EWindowStep(E2EStep):
    """N > 1: one window iteration through the repo's own multi-GPU API -- MappingWindow.step with the CUDA render
    backend (fused mapping loss) and the RowExchange -- one keyframe per rank, parameters and targets arriving from
    pinned host memory every step (`copies=False`: the same with the inputs already resident).

    The Gaussian parameters are REPLICATED over the ranks, so they cross PCIe once, not N times: every rank copies
    its 1/N row slice of each parameter tensor from pinned host memory and the slices are all-gathered over NVLink
    (torch.distributed all_gather_into_tensor on the copy stream: plumbing); the per-rank targets are copied whole.
    Per step the box still moves every input byte from host memory inside the timed region -- 61 MB in total instead
    of N x 61 MB, which at 8 GPUs is what the host memory system can feed (8 x 61 MB per 0.6 ms step = 800 GB/s
    would be needed; measured ceiling of the dense scheme: 189 GB/s aggregate = 3100 it/s)."""

    def __init__(self, wl: Workload, world, copies=True):
        self.rank = dist.get_rank()
        self.world_n = world
        self.copy_group = dist.new_group(backend="nccl")   # own communicator: the input all-gathers do not queue
        super().__init__(wl, world, "ours")                 # behind / in front of the window's collectives
        from monogs_b200 import synth, window
        self.copies = copies
        self.cur_slot = 0
        cams = synth.arc_cameras(wl.kind, world)
        self.window = window.MappingWindow(window.make_cuda_render_fn(cams, _SlotTargets(self), wl.bg))
        self.host_out = torch.empty(7, dtype=torch.float32).pin_memory()
        P = wl.P
        lo, hi = self._slice(P)
        row_bytes = sum(wl.host[k][0].numel() * 4 for k in self.NAMES)
        self.h2d = (hi - lo) * row_bytes + (self.pinned_img.numel() + self.pinned_dep.numel()) * 4

    def _slice(self, P):
        per = (P + self.world_n - 1) // self.world_n
        return min(P, self.rank * per), min(P, (self.rank + 1) * per)

    def _prefetch(self, slot):
        if getattr(self, "world_n", 1) <= 1 or self.wl.P % self.world_n != 0:
            return super()._prefetch(slot)
        P = self.wl.P
        lo, hi = self._slice(P)
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.consumed[slot])  # do not overwrite inputs still being read
            for k in self.NAMES:
                full = self.dev_buf[slot][k]
                full[lo:hi].copy_(self.pinned[k][lo:hi], non_blocking=True)      # my rows: host -> device
                dist.all_gather_into_tensor(full, full[lo:hi], group=self.copy_group)   # everyone's rows: NVLink
            self.dev_img[slot].copy_(self.pinned_img, non_blocking=True)
            self.dev_dep[slot].copy_(self.pinned_dep, non_blocking=True)
            self.ready[slot].record(self.copy_stream)

    def step(self):
        wl = self.wl
        if self.copies:
            slot = self.i % self.SLOTS
            self.i += 1
            torch.cuda.current_stream().wait_event(self.ready[slot])
            self._prefetch((slot + 2) % self.SLOTS)
        else:
            slot = 0
            torch.cuda.current_stream().wait_event(self.ready[0])   # the initial prefetch (complete after warm-up)
        self.cur_slot = slot
        b = self.dev_buf[slot]
        params = {"means3D": b["means3D"], "colors": b["colors"], "opacities": b["opacities"], "scales": b["scales"],
                  "rotations": b["rotations"]}
        res = self.window.step(params, self.world)
        kf = res.local[0]
        out = torch.cat([kf.loss.detach().view(1), kf.grad_tau.view(-1)])
        self.host_out.copy_(out, non_blocking=True)
        if self.copies:
            self.consumed[slot].record()
        torch.cuda.current_stream().synchronize()  # the caller reads the loss / pose gradient on the host
        return float(self.host_out[0])


def window_c4_leg(rank, world, dev, iters=6):
    """BASELINE config 4 as a FIXED window: 8 keyframes x 2 M Gaussians at 1200x680 through MappingWindow.step(params, 8),
    keyframes k = rank, rank + N, ... on this rank, one exchange per window iteration. Device-timed, max over ranks."""
    from monogs_b200 import synth, window
    K, P, kind = 8, 2_000_000, "replica"
    cams = synth.arc_cameras(kind, K)
    sc = {k: v.to(dev) for k, v in synth.make_scene(P, synth.make_camera(kind), seed=0, device="cpu").items()}
    targets = [tuple(t.to(dev) for t in synth.make_targets(c, seed=i)) for i, c in enumerate(cams)]
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    w = window.MappingWindow(window.make_cuda_render_fn(cams, targets, bg))
    for _ in range(3):
        w.step(sc, K)
    barrier(world)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        w.step(sc, K)
    b.record()
    barrier(world)
    secs = max_over_ranks(a.elapsed_time(b) / 1e3, world, dev)
    rep = dict(w.exchange.report)
    return {"metric": "mapping-window it/s (keyframe fwd+bwd iterations of a fixed 8-keyframe window)",
            "value": K * iters / secs, "unit": "it/s", "windows_per_s": iters / secs, "ms_per_window": secs / iters * 1e3,
            "keyframes": K, "keyframes_per_rank": len(window.shard_keyframes(K, world, rank)), "gaussians": P,
            "image": [cams[0].W, cams[0].H], "scaling": "strong", "api": "MappingWindow.step(params, 8)",
            "exchange": w.exchange.mode if world > 1 else None, "exchange_times_ms": rep.get("times_ms"),
            "touched_rows_local": rep.get("touched_rows_local")}


# ------------------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port on the host cores (bounded sample)
# ------------------------------------------------------------------------------------------------------------
def cpu_baseline(wl: Workload, budget_s=25.0):
    from oracle import cpu_oracle as co
    n = lambda t: t.detach().cpu().numpy()
    h = wl.host
    view, proj, proj_raw, campos = [n(t) for t in (wl.view, wl.proj, wl.proj_raw, wl.campos)]
    dcol, ddep, bg = n(wl.dcol), n(wl.ddep), n(wl.bg)
    co.load()
    times = []
    t_begin = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        f = co.forward(n(h["means3D"]), n(h["scales"]), n(h["rotations"]), n(h["opacities"]), n(h["colors"]), view,
                       proj, campos, wl.W, wl.H, wl.cam.tanfovx, wl.cam.tanfovy, bg)
        co.backward(f, n(h["means3D"]), n(h["scales"]), n(h["rotations"]), view, proj, proj_raw, wl.W, wl.H,
                    wl.cam.tanfovx, wl.cam.tanfovy, bg, dcol, ddep)
        times.append(time.perf_counter() - t0)
        if len(times) >= 3 or time.perf_counter() - t_begin > budget_s * 0.6:
            break
    best = min(times)
    cores = int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1))
    return {"value": 1.0 / best, "unit": "it/s", "cores": cores, "kind": "port",
            "sample": f"{len(times)} full fwd+bwd iteration(s) of the same {wl.name} workload "
                      f"({wl.W}x{wl.H}, P={wl.P}) with oracle/cpu_ref.cpp (OpenMP), best of {len(times)}"}


def cpu_torch_c1():
    """BASELINE.json's reported baseline "a host-core PyTorch reference of the same render() math, with core count
    stated", on its config C1 (10 k Gaussians, 320x240, one tracking-shaped step fwd + autograd bwd): oracle/torch_ref.py
    on all host cores, best of 3."""
    from monogs_b200 import synth
    from oracle import torch_ref as tr
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cam = synth.make_camera("tum_half", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(10_000, cam, seed=0, device="cpu")
    view, proj, _, _ = cam.matrices("cpu")
    g = torch.Generator().manual_seed(7)
    dcol = torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)
    ddep = torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)
    bg = torch.tensor([0.1, 0.2, 0.3])
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        tr.tracking_step(sc, view, proj, cam.W, cam.H, cam.tanfovx, cam.tanfovy, bg, dcol, ddep)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": 1.0 / best, "unit": "it/s", "cores": cores, "kind": "port",
            "sample": "config C1 (10 k Gaussians, 320x240), one fwd + autograd bwd with oracle/torch_ref.py "
                      "(vectorised PyTorch on the host cores), best of 3"}


# ------------------------------------------------------------------------------------------------------------
# timing helpers
# ------------------------------------------------------------------------------------------------------------
def barrier(world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world, dev):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed_steps(step_fn, steps, warmup, world, dev, flush):
    """W untimed steps, then K steps each bracketed by its own CUDA-event pair on the launching stream, the L2
    flushed (256 MiB write, outside the event pairs) before every step. Returns device seconds for K steps
    (max over ranks), bracketed by barrier + synchronize on both sides."""
    for _ in range(warmup):
        flush.zero_()
        step_fn()
    barrier(world)
    pairs = []
    for _ in range(steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step_fn()
        b.record()
        pairs.append((a, b))
    barrier(world)
    ms = sum(a.elapsed_time(b) for a, b in pairs)
    return max_over_ranks(ms / 1e3, world, dev)


def stage_bytes(u):
    """Algorithmic bytes per launch of each stage (SURVEY 8d terms; DESIGN.md 'Kernels and rooflines').
    Render stages use the instances actually consumed (early termination skips the rest of a tile's list)."""
    P, V, R, N = u["P"], u["V"], u["R"], u["N"]
    Rc = u["R_consumed"] if u.get("R_consumed") is not None else R
    T = (N + 255) // 256  # tiles (exact for the bench's 16-divisible image sizes)
    if T <= 4096:
        # sort-free binning (tile_scatter.cu): count matrix of nb CTAs x T tiles instead of instance keys
        nb = 148 * max(1, min(4, 232448 // (16512 + 8448 + 24 * T + 1024)))
        emit, tile_sort, ranges = 8 * P + 4 * nb * T, 12 * nb * T + 12 * P + 4 * R + 16 * T, 0
    else:
        emit, tile_sort, ranges = 12 * R, 24 * R, 8 * R
    return {
        "preprocess": 44 * P + 8 * P + 52 * V,
        "depth_sort": 16 * P,
        "scan": 8 * P,
        "emit": emit,
        "tile_sort": tile_sort,
        "ranges": ranges,
        "render_fwd": 44 * Rc + 28 * N + 4 * P,
        "render_bwd": 84 * Rc + 24 * N,
        "preprocess_bwd": 164 * V + 48 * P,
    }


def step_bytes(u):
    """Whole fwd+bwd iteration, the single formula of SURVEY 8d / BASELINE.md section 4 (with the full R)."""
    P, V, R, N = u["P"], u["V"], u["R"], u["N"]
    return (64 * P + 52 * V + 88 * R + 28 * N) + (84 * R + 24 * N + 164 * V + 48 * P)


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full summary, if present."""
    p = ROOT / "profiles" / "traffic.json"
    if p.exists():
        try:
            return json.loads(p.read_text()).get(kernel)
        except Exception:
            return None
    return None


def main():
    a = parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path (use gpurun)")
    rank, world, local = dist_setup(a.gpus)
    dev = torch.device("cuda", local if world > 1 else 0)
    wl_name = a.workload

    # ---- which implementation is measured --------------------------------------------------------------------
    impl, ref_kind = a.impl, None
    if impl == "reference":
        from oracle import ref_cuda
        if ref_cuda.available():
            ref_kind = "reference"  # the unmodified reference CUDA kernels, rebuilt for sm_100a
        else:
            ref_kind = "port"       # oracle/_ref did not travel: the CPU oracle port on the host cores
    if impl == "reference" and ref_kind == "port":
        if rank == 0:
            wl = Workload(wl_name, 0, 1, dev)
            cb = cpu_baseline(wl, budget_s=60.0)
            line = {"impl": "reference", "metric": "rasterize fwd+bwd it/s", "value": cb["value"], "unit": "it/s",
                    "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 / cb["value"],
                    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                    "data": "synthetic", "config": {"workload": wl_name, "detail": wl.desc},
                    "cpu_baseline": cb,
                    "e2e": {"value": cb["value"], "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    wl = Workload(wl_name, rank, world, dev)
    stepper = OursDeviceStep(wl, world) if impl == "ours" else RefDeviceStep(wl, world)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    lib = None
    if impl == "ours":
        from monogs_b200 import _abi
        lib = _abi.load()

    # ---- device-resident value --------------------------------------------------------------------------------
    sampler = ClockSampler(local if world > 1 else 0) if rank == 0 else None
    launches0 = lib.gsr_profile_kernel_launches() if lib else 0
    for _ in range(a.warmup):
        flush.zero_()
        stepper.step()
    barrier(world)
    # Stage times: (1) a short untimed pass with all nine stages bracketed by events gives the stage table and names
    # the dominant stage; (2) inside the TIMED region only that stage is bracketed (its live launch duration is what
    # `roofline.achieved` uses) -- an event record between two kernels removes their programmatic-launch overlap, so
    # bracketing all nine inside the timed region would tax `value` itself.
    stage_ms, stage_n, dom_ms, dom_n, dom_idx = None, None, None, None, None
    if lib:
        lib.gsr_profile_collect((C.c_double * 9)(), (C.c_int * 9)())  # drop anything recorded so far
        lib.gsr_profile_enable(1)
        for _ in range(min(10, max(3, a.steps))):
            if flush is not None:
                flush.zero_()
            stepper.step()
        torch.cuda.synchronize(dev)
        lib.gsr_profile_enable(0)
        ms_arr, n_arr = (C.c_double * 9)(), (C.c_int * 9)()
        rc = lib.gsr_profile_collect(ms_arr, n_arr)
        assert rc == 0, rc
        stage_ms, stage_n = list(ms_arr), list(n_arr)
        per = [stage_ms[i] / stage_n[i] if stage_n[i] else 0.0 for i in range(9)]
        dom_idx = max(range(9), key=lambda i: per[i])
        barrier(world)
        lib.gsr_profile_enable(1 | (1 << (8 + dom_idx)))
        launches0 = lib.gsr_profile_kernel_launches()
    secs = timed_steps(stepper.step, a.steps, 0, world, dev, flush)
    gpu_launches = None
    if lib:
        gpu_launches = int(lib.gsr_profile_kernel_launches() - launches0)
        ms_arr, n_arr = (C.c_double * 9)(), (C.c_int * 9)()
        lib.gsr_profile_enable(0)
        rc = lib.gsr_profile_collect(ms_arr, n_arr)
        assert rc == 0, rc
        dom_ms, dom_n = ms_arr[dom_idx], n_arr[dom_idx]
    clocks = sampler.stop() if sampler else None
    units = stepper.units()
    value = world * a.steps / secs

    # ---- end to end ---------------------------------------------------------------------------------------------
    e2e = None
    def e2e_leg(stepper_, k):
        """k steps through the public API with the host copies inside, device-timed, max over ranks; two passes, the
        better one is kept (a 70 ms leg is easily disturbed by host jitter on a fresh box)."""
        for _ in range(5):
            stepper_.step()
        best, wall = None, None
        for _ in range(2):
            barrier(world)
            t0 = torch.cuda.Event(enable_timing=True)
            t1 = torch.cuda.Event(enable_timing=True)
            w0 = time.perf_counter()
            t0.record()
            for _ in range(k):
                stepper_.step()
            t1.record()
            barrier(world)
            w = time.perf_counter() - w0
            s_ = max_over_ranks(max(t0.elapsed_time(t1) / 1e3, 0.0), world, dev)
            if best is None or s_ < best:
                best, wall = s_, w
        return best, wall

    if not a.no_e2e:
        windowed = world > 1 and impl == "ours"
        es = E2EWindowStep(wl, world) if windowed else E2EStep(wl, world, impl)
        k_e2e = max(3, min(2 * a.steps, 60))  # longer than the device leg: host-side jitter averages out
        e2e_s, wall = e2e_leg(es, k_e2e)
        e2e = {"value": world * k_e2e / e2e_s, "unit": "it/s", "h2d_bytes_per_step": es.h2d,
               "d2h_bytes_per_step": es.d2h, "steps": k_e2e, "passes": 2, "wall_s": round(wall, 4),
               "api": ("MappingWindow.step (CUDA render backend, fused mapping loss, RowExchange mode "
                       + es.window.exchange.mode + "); replicated parameters: 1/N row slice per rank from pinned host "
                       "memory + NVLink all-gather, per-rank targets whole") if windowed else
                      ("GaussianRasterizer.forward + loss.backward()" if impl == "ours" else
                       "reference CUDA via oracle/_ref + loss.backward()" + (" + ncclAllReduce" if world > 1 else "")),
               "h2d_gb_per_s_per_gpu": round(es.h2d * k_e2e / e2e_s / 1e9, 1)}
        if e2e["h2d_gb_per_s_per_gpu"] > 40.0:
            e2e["note"] = ("H2D-bound: %.1f MB of pinned-host inputs per step and GPU at %.1f GB/s (PCIe gen5 x16 tops "
                           "out near 55 GB/s); kernel speed shows in `value`, not here"
                           % (es.h2d / 1e6, e2e["h2d_gb_per_s_per_gpu"]))
        if windowed:
            er = E2EWindowStep(wl, world, copies=False)
            er_s, _ = e2e_leg(er, k_e2e)
            e2e["resident"] = {"value": world * k_e2e / er_s, "unit": "it/s",
                               "what": "the same MappingWindow.step leg with the host->device copies hoisted out"}
            del er
        del es

    e2e_fused = None
    if not a.no_e2e and impl == "ours" and world == 1:
        # the same iteration through the ADDITIONAL fused entry point (monogs_b200.fused_loss.render_l1_loss)
        ef = E2EStep(wl, world, "ours-fused")
        ef_s, _ = e2e_leg(ef, k_e2e)
        e2e_fused = {"value": world * k_e2e / ef_s, "unit": "it/s", "h2d_bytes_per_step": ef.h2d,
                     "d2h_bytes_per_step": ef.d2h, "api": "fused_loss.render_l1_loss + loss.backward()"}

    window_c4 = None
    if impl == "ours" and wl_name == "headline" and not a.no_window_c4:
        del stepper.scratch, stepper.geom                      # free the headline buffers first
        torch.cuda.empty_cache()
        window_c4 = window_c4_leg(rank, world, dev)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    line = {
        "metric": "rasterize fwd+bwd it/s", "value": value, "unit": "it/s", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": secs / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl_name, "detail": wl.desc, "image": [wl.W, wl.H], "gaussians": wl.P,
                   "l2": "256 MiB flush write before every timed step (outside the event pairs)",
                   "schedule": "render CTAs start heaviest tiles first: backward from this step's forward, forward from "
                               "the previous step's per-tile work left in the image workspace (a hint only; "
                               "GSR_FWD_TILE_ORDER=0 / GSR_TILE_ORDER=0 switch it off)",
                   "multi_gpu": ("mapping window: one keyframe per GPU; sum of the [P,16] window rows (14 gradient + 2 "
                                 "densification columns) over the ranks inside the step, exchange = "
                                 + getattr(getattr(stepper, "exchange", None), "mode", "all_reduce"))
                                if world > 1 else "single keyframe",
                   "units": units},
        "clocks": clocks, "e2e": e2e, "gpu_launches": gpu_launches,
    }
    if e2e_fused:
        line["e2e_fused_loss"] = e2e_fused
    if window_c4:
        line["window_c4"] = window_c4
    if world > 1 and getattr(stepper, "exchange", None) is not None:
        line["exchange"] = dict(getattr(stepper.exchange, "report", {}), mode=stepper.exchange.mode)
    if impl == "reference":
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": value, "unit": "it/s", "cores": 0, "kind": "reference",
                                "sample": "the reference has no CPU implementation of this path: these are its own "
                                          "CUDA kernels (oracle/_ref, unmodified sources rebuilt for sm_100a) on the GPU"}
        line["gpu_launches"] = None
    else:
        sb = stage_bytes(units)
        from monogs_b200._abi import STAGES
        per_launch_ms = {s: (stage_ms[i] / stage_n[i] if stage_n[i] else 0.0) for i, s in enumerate(STAGES)}
        dom = STAGES[dom_idx]
        if dom_n:   # the dominant stage's launches inside the timed region (the table's other stages: the pass before)
            per_launch_ms[dom] = dom_ms / dom_n
        achieved = sb[dom] / (per_launch_ms[dom] * 1e-3) / 1e9 if per_launch_ms[dom] > 0 else 0.0
        line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                            "frac": achieved / peak, "traffic": ncu_traffic(dom), "peak_source": peak_src,
                            "algorithmic_bytes_per_launch": sb[dom], "ms_per_launch": per_launch_ms[dom]}
        # the render kernels are instruction-issue bound, not HBM bound: the same launch against the
        # issue-slot ceiling = 148 SMs x 4 schedulers x SM clock, with the warp-instruction count of the committed
        # ncu capture (valid for the headline workload it was taken on)
        try:
            wi = json.loads((ROOT / "profiles" / "traffic.json").read_text()).get("_warp_inst", {}).get(dom)
        except Exception:
            wi = None
        if wi and wl_name == "headline" and world == 1 and per_launch_ms[dom] > 0:
            mhz = (clocks or {}).get("sm_mhz") or 1965.0
            peak_issue = 148 * 4 * mhz * 1e6
            ach = wi / (per_launch_ms[dom] * 1e-3)
            line["roofline"]["issue_slots"] = {"warp_inst_per_launch": wi, "achieved_ginst_s": ach / 1e9,
                                               "peak_ginst_s": peak_issue / 1e9, "frac": ach / peak_issue}
        # the same kernels against SURVEY 8d's per-instance terms with the FULL instance count R (the formula
        # assumes every instance of a tile list is read; early termination makes the kernels skip most of them)
        full = dict(units, R_consumed=None)
        sbf = stage_bytes(full)
        line["roofline"]["survey_formula_full_R"] = {
            s: {"bytes": sbf[s], "frac": round(sbf[s] / (per_launch_ms[s] * 1e-3) / 1e9 / peak, 4)}
            for s in ("render_fwd", "render_bwd") if per_launch_ms[s] > 0}
        line["stages_ms_per_step"] = {s: round(per_launch_ms[s], 5) for s in STAGES}
        line["stages_note"] = ("the dominant stage (" + dom + ") is timed with CUDA events inside the timed region; the "
                               "other stages in a 10-step pass right before it, every stage bracketed by events (which "
                               "removes the programmatic-launch overlap between kernels: their sum exceeds ms_per_step's share)")
        line["stages_frac_of_hbm_peak"] = {s: round(sb[s] / (per_launch_ms[s] * 1e-3) / 1e9 / peak, 4)
                                           if per_launch_ms[s] > 0 else None for s in STAGES}
        tb = step_bytes(units)
        line["step_roofline"] = {"algorithmic_bytes": tb, "achieved": tb / (secs / a.steps) / 1e9, "unit": "GB/s",
                                 "frac": tb / (secs / a.steps) / 1e9 / peak}
        if world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl)
            try:
                line["cpu_baseline_torch_c1"] = cpu_torch_c1()
            except Exception as e:  # noqa: BLE001  (reported extra; never fails the bench line)
                line["cpu_baseline_torch_c1"] = {"error": repr(e)[:200]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
