"""Mapping-window sharding across the GPUs of one box (the only stage of the path that shards).

The reference's `Mapper.optimize_map` (utils/slam_mapper.py:244-500) renders every keyframe of the window,
sums the mapping losses and calls ONE backward, so the Gaussian-parameter gradients are sums over keyframes
(:273-322, :394).  Keyframes are independent rasterizer problems over the same Gaussian set, so here keyframe k
is rendered on rank k mod G and the per-rank partial sums meet in ONE exchange step:

  * allreduce(sum) of a flat [P, 14+2] fp32 buffer: d/d(xyz 3, colour 3, opacity 1, scale 3, rotation 4) plus the
    two densification accumulators (sum of ||d mean2D[:, :2]|| and the visibility count,
    gaussian_model.py:888-892) -- NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests;
  * allreduce(max) of the screen radii (max_radii_2d, slam_mapper.py:453-457);
  * pose gradients (theta/rho) stay with the rank that owns the keyframe (they belong to one keyframe).

One process per GPU (torchrun); the Gaussian parameters are replicated and every rank applies the identical
optimizer step afterwards, so no broadcast is needed.  The render backend is injected (`render_fn`) so the
host logic is testable on CPU with world_size 2 / gloo; the product default is the sm_100a rasterizer.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

PARAM_COLUMNS = (("means3D", 3), ("colors", 3), ("opacities", 1), ("scales", 3), ("rotations", 4))
AUX_COLUMNS = (("grad2d_norm", 1), ("vis_count", 1))
N_COLS = sum(n for _, n in PARAM_COLUMNS) + sum(n for _, n in AUX_COLUMNS)  # 16


def shard_keyframes(n_keyframes: int, world_size: int, rank: int) -> List[int]:
    """Indices of the window's keyframes rendered by `rank` (round robin: k -> k mod G)."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    return list(range(rank, n_keyframes, world_size))


def column_slices() -> Dict[str, slice]:
    out, c = {}, 0
    for name, n in PARAM_COLUMNS + AUX_COLUMNS:
        out[name] = slice(c, c + n)
        c += n
    return out


def window_chunks(P: int, n_chunks: int = 4) -> List[tuple]:
    """Gaussian ranges [p0, p1) of the window backward: 256-aligned, shrinking (40/30/20/10 %) so that the
    allreduce of the LAST range -- the only one nothing is left to overlap with -- is the smallest. A pure
    function of (P, n_chunks): every rank derives the same ranges, so the collectives line up."""
    if P <= 0:
        return []
    if n_chunks <= 1 or P < 4096:
        return [(0, P)]
    weights = list(range(n_chunks, 0, -1))
    total, bounds, acc = float(sum(weights)), [0], 0
    for w in weights[:-1]:
        acc += w
        bounds.append(min(P, int(P * acc / total) // 256 * 256))
    bounds.append(P)
    return [(a, b) for a, b in zip(bounds[:-1], bounds[1:]) if b > a]


@dataclass
class KeyframeResult:
    """What one keyframe contributes (the per-keyframe items of slam_mapper.py:319-322)."""
    index: int
    radii: torch.Tensor                 # i32[P]
    n_touched: torch.Tensor             # i32[P]
    grad_tau: Optional[torch.Tensor]    # f32[6] = [rho | theta], stays local
    loss: Optional[torch.Tensor] = None


@dataclass
class WindowStepResult:
    grads: Dict[str, torch.Tensor]            # summed over ALL keyframes of the window (every rank identical)
    grad2d_norm: torch.Tensor                 # f32[P]  sum over keyframes of ||dL/dmean2D[:, :2]||
    vis_count: torch.Tensor                   # f32[P]  number of keyframes in which the Gaussian is visible
    max_radii: torch.Tensor                   # i32[P]
    local: List[KeyframeResult] = field(default_factory=list)


# render_fn(keyframe_index, params, flat, accumulate, chunks, chunk_done) -> KeyframeResult.
# It must WRITE (accumulate False) or ADD (accumulate True) this keyframe's Gaussian-parameter gradients and aux
# columns into flat[P, N_COLS] (fp32, on the parameters' device), and -- when chunk_done is given -- call
# chunk_done(p0, p1) for every range of `chunks`, in order, as soon as rows [p0, p1) are final on the current stream.
RenderFn = Callable[..., KeyframeResult]


class RowExchange:
    """The one exchange step: sum of the [P,16] window rows over the ranks.

    On NVSwitch boxes the rows live in symmetric memory and are reduced IN THE SWITCH (NVLS multimem
    ld_reduce / st through torch.distributed._symmetric_memory): measured on 8 x B200 at 64 MB: 0.172 ms vs
    0.278 ms for ncclAllReduce (gpurun_out/nccl_probe2_n8.json, profiles/r01_nccl_probe_n8.json). Elsewhere
    (gloo on CPU, no multicast support) it is a plain all_reduce. Splitting the rows into ranges to overlap the
    allreduce with the per-Gaussian backward was measured and rejected: 4 pieces cost 0.47 ms (latency-bound
    collectives) against 0.064 ms of overlappable compute -- so the default is ONE range."""

    def __init__(self, group=None):
        self.group = group
        self.mode = "all_reduce"
        self._name = None

    def allocate(self, P: int, device) -> torch.Tensor:
        device = torch.device(device)
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        if device.type == "cuda" and world > 1:
            try:
                import torch.distributed._symmetric_memory as symm_mem
                grp = self.group if self.group is not None else dist.group.WORLD
                t = symm_mem.empty(P * N_COLS, dtype=torch.float32, device=device)
                hdl = symm_mem.rendezvous(t, grp.group_name)
                if getattr(hdl, "multicast_ptr", 0) and hasattr(torch.ops.symm_mem, "multimem_all_reduce_"):
                    self._name = grp.group_name
                    flat = t.view(P, N_COLS)
                    self.mode = self._pick(flat)
                    return flat
            except Exception:  # no symmetric memory / multicast on this system: NCCL ring/tree instead
                pass
        return torch.empty(P, N_COLS, dtype=torch.float32, device=device)

    def _pick(self, flat: torch.Tensor) -> str:
        """Time both exchanges once on the real buffer and keep the faster (the in-switch reduction wins at 8 GPUs,
        ncclAllReduce at 2); the decision is made identical on every rank with a MAX-allreduce of the timings."""
        flat.zero_()
        times = []
        for mode in ("multimem", "all_reduce"):
            self.mode = mode
            for _ in range(2):
                w = self.reduce(flat, 0, flat.shape[0])
                if w is not None:
                    w.wait()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(flat.device)
            a.record()
            for _ in range(3):
                w = self.reduce(flat, 0, flat.shape[0])
                if w is not None:
                    w.wait()
            b.record()
            torch.cuda.synchronize(flat.device)
            times.append(a.elapsed_time(b))
        t = torch.tensor(times, dtype=torch.float64, device=flat.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return "multimem" if float(t[0]) < float(t[1]) else "all_reduce"

    def reduce(self, flat: torch.Tensor, p0: int, p1: int):
        """Start the sum of rows [p0, p1); returns a handle with .wait() or None when already stream-ordered."""
        if self.mode == "multimem":
            torch.ops.symm_mem.multimem_all_reduce_(flat.view(-1)[p0 * N_COLS:p1 * N_COLS], "sum", self._name)
            return None
        return dist.all_reduce(flat[p0:p1], op=dist.ReduceOp.SUM, group=self.group, async_op=True)


class MappingWindow:
    def __init__(self, render_fn: RenderFn, group: Optional[dist.ProcessGroup] = None, n_chunks: int = 1):
        self.render_fn = render_fn
        self.group = group
        self.n_chunks = n_chunks
        self.exchange = RowExchange(group)
        self._flat: Optional[torch.Tensor] = None
        self._radii: Optional[torch.Tensor] = None

    @property
    def world_size(self) -> int:
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    @property
    def rank(self) -> int:
        return dist.get_rank(self.group) if dist.is_available() and dist.is_initialized() else 0

    def _buffers(self, P: int, device) -> None:
        if self._flat is None or self._flat.shape[0] != P or self._flat.device != torch.device(device):
            self._flat = self.exchange.allocate(P, device)
            self._radii = torch.empty(P, dtype=torch.int32, device=device)

    def step(self, params: Dict[str, torch.Tensor], n_keyframes: int) -> WindowStepResult:
        """One window iteration: local keyframes -> partial sums -> exchange. `params` maps the names of
        PARAM_COLUMNS to replicated tensors. The exchange is started range by range while the last local
        keyframe's per-Gaussian backward is still running (async allreduce per finished range of rows)."""
        P = int(params["means3D"].shape[0])
        dev = params["means3D"].device
        self._buffers(P, dev)
        flat, radii = self._flat, self._radii
        world = self.world_size
        chunks = window_chunks(P, self.n_chunks if world > 1 else 1)
        works = []

        def chunk_done(p0, p1):
            if world > 1:
                w = self.exchange.reduce(flat, p0, p1)
                if w is not None:
                    works.append(w)

        mine = shard_keyframes(n_keyframes, world, self.rank)
        local = []
        radii.zero_()
        if not mine:  # more GPUs than keyframes: contribute zeros, but take part in every collective
            flat.zero_()
            for c in chunks:
                chunk_done(*c)
        for i, k in enumerate(mine):
            last = i == len(mine) - 1
            r = self.render_fn(k, params, flat, i > 0, chunks, chunk_done if last else None)
            torch.maximum(radii, r.radii.to(torch.int32), out=radii)
            local.append(r)
        if world > 1:
            works.append(dist.all_reduce(radii, op=dist.ReduceOp.MAX, group=self.group, async_op=True))
            for w in works:
                w.wait()
        sl = column_slices()
        grads = {name: flat[:, sl[name]] for name, _ in PARAM_COLUMNS}
        return WindowStepResult(grads=grads, grad2d_norm=flat[:, sl["grad2d_norm"]].squeeze(1),
                                vis_count=flat[:, sl["vis_count"]].squeeze(1), max_radii=radii, local=local)


def mapping_loss_grad(color, depth, gt_image, gt_depth):
    """0.9 L1(rgb) + 0.1 L1(depth over gt_depth > 0) and its gradient w.r.t. the rendered images
    (get_loss_mapping, utils/slam_utils.py:101-146, RGB-D branch without exposure / rgb-boundary masks)."""
    diff = color - gt_image
    n_rgb = diff.numel()
    valid = gt_depth > 0
    n_d = valid.sum().clamp_min(1)
    ddiff = (depth - gt_depth) * valid
    loss = 0.9 * diff.abs().sum() / n_rgb + 0.1 * ddiff.abs().sum() / n_d
    return loss, 0.9 * torch.sign(diff) / n_rgb, 0.1 * torch.sign(ddiff) / n_d


def make_cuda_render_fn(cameras: Sequence, targets: Sequence, bg: torch.Tensor,
                        loss_grad: Optional[Callable] = None) -> RenderFn:
    """Product backend: the sm_100a rasterizer through the `_C` mirror (C ABI). The per-Gaussian half of the
    backward writes 64-byte window rows straight into `flat` (gsr_backward_rows), range by range.

    cameras[k] is a synth.Camera-like object with .matrices(device), .W, .H, .tanfovx, .tanfovy; targets[k] =
    (gt_image[3,H,W], gt_depth[1,H,W]).  loss_grad(color, depth, gt_image, gt_depth) -> (loss, dL_dcolor,
    dL_ddepth); default = mapping_loss_grad.
    """
    from .diff_gaussian_rasterization import _C

    cached = {}
    lg = loss_grad or mapping_loss_grad

    def render(k, params, flat, accumulate, chunks, chunk_done) -> KeyframeResult:
        cam = cameras[k]
        dev = flat.device
        if k not in cached:
            cached[k] = cam.matrices(dev)
        view, proj, proj_raw, campos = cached[k]
        e = torch.empty(0, device=dev)
        R, color, radii, geom, plist, img, depth, opacity, n_touched = _C.rasterize_gaussians(
            bg, params["means3D"], params["colors"], params["opacities"], params["scales"], params["rotations"], 1.0,
            e, view, proj, proj_raw, cam.tanfovx, cam.tanfovy, cam.H, cam.W, e, 0, campos, False, False)
        gt_image, gt_depth = targets[k]
        loss, dcol, ddep = lg(color, depth, gt_image, gt_depth)
        tau, _ = _C.rasterize_gaussians_backward_rows(
            bg, params["means3D"], radii, params["scales"], params["rotations"], 1.0, view, proj, proj_raw,
            cam.tanfovx, cam.tanfovy, dcol.contiguous(), ddep.contiguous(), geom, R, plist, img, flat, accumulate,
            chunks=chunks, chunk_done=chunk_done)
        return KeyframeResult(index=k, radii=radii, n_touched=n_touched, grad_tau=tau, loss=loss)

    return render
