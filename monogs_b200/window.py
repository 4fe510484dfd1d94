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

import os
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

    On NVSwitch boxes the rows live in symmetric memory (torch.distributed._symmetric_memory supplies the peer
    mappings and the NVLS multicast address; nothing else) and three exchanges are available:
      * "sparse"     OUR kernel (csrc/rows_exchange.cu, gsr_rows_exchange): only rows that are non-zero on some rank
                     are summed in the switch (multimem.ld_reduce) and multicast back, found through per-rank bitmaps
                     the per-Gaussian backward writes (gsr_backward_rows_marked); visibility counts travel as bytes;
      * "multimem"   torch's dense in-switch allreduce of the whole buffer (0.172 ms for 64 MB at 8 GPUs);
      * "all_reduce" ncclAllReduce (0.278 ms; the only mode on CPU / gloo and without multicast support).
    The first real exchange times all available modes ON THE REAL ROWS (rows and marks are restored between trials),
    checks the sparse result against the dense one, and keeps the fastest; the decision is made identical on every
    rank with a MAX-allreduce of the timings. Splitting the rows into ranges to overlap the allreduce with the
    per-Gaussian backward was measured and rejected (4 pieces 0.47 ms, latency-bound)."""

    def __init__(self, group=None):
        self.group = group
        self.mode = "all_reduce"
        self.candidates = ["all_reduce"]
        self.decided = True
        self.report = {}
        self.marks = None          # (touched_words_ptr, vis_u8_ptr) of this rank's slot, or None
        self.vis = None            # f32[P] visibility sum of the sparse mode
        self._name = None
        self._buf = None           # symmetric block (f32), sized for self._cap rows
        self._plain = None         # ordinary rows buffer of the NCCL / gloo mode, f32[cap, 16]
        self._hdl = self._peers = None
        self._P = 0
        self._cap = 0              # grow-only row capacity of the current buffer
        self._decided_mode = None  # survives re-allocations

    # -- allocation ------------------------------------------------------------------------------------------
    @staticmethod
    def _capacity_for(P: int, old: int) -> int:
        """Grow-only row capacity: at least 1.5x the previous one, a multiple of 4096 rows."""
        want = max(P, old + old // 2, 4096)
        return (want + 4095) // 4096 * 4096

    def allocate(self, P: int, device) -> torch.Tensor:
        """-> flat f32[P, 16] rows for the live P. The buffer behind it is sized for a grow-only CAPACITY, so a map
        that grows by keyframe insertion / densification (every few iterations in MonoGS) re-uses the symmetric
        block, its peer mappings and the mode decided on the first exchange; only outgrowing the capacity allocates
        (and rendezvouses) again, and then the previous block and handle are dropped first."""
        device = torch.device(device)
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self._P = P
        if self._buf is not None and P <= self._cap and self._buf.device == device:
            return self._buf[:P * N_COLS].view(P, N_COLS)          # fits: same block, same marks, same decision
        if self._plain is not None and P <= self._cap and self._plain.device == device:
            return self._plain[:P]
        cap = self._capacity_for(P, self._cap)
        if device.type == "cuda" and world > 1 and os.environ.get("GSR_XCHG", "auto") != "nccl":
            self._release()
            flat, err = None, None
            try:
                flat = self._allocate_symmetric(P, cap, device, world)
            except Exception as e:  # no symmetric memory / multicast on this system
                err = repr(e)[:200]
            # every rank must take the same branch (a rank-local failure would pair our spin-barrier kernel on one
            # side with ncclAllReduce on the other and hang): agree with one MIN-allreduce of the outcome
            ok = torch.tensor([1 if flat is not None else 0], dtype=torch.int32, device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
            if int(ok.item()) == 1:
                return flat
            if err is not None:
                self.report["symmetric_memory_error"] = err
            self._release()
            self.candidates, self.mode, self.decided = ["all_reduce"], "all_reduce", True
        self._cap = cap
        self._plain = torch.empty(cap, N_COLS, dtype=torch.float32, device=device)
        return self._plain[:P]

    def _release(self):
        """Drop the symmetric block, its rendezvous handle and the peer-pointer tensor before replacing them."""
        self._buf = self._hdl = self._peers = self._plain = None
        self.marks = self.vis = None

    def _allocate_symmetric(self, P, cap, device, world):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm_mem
        from . import _abi
        lib = _abi.load()
        grp = self.group if self.group is not None else dist.group.WORLD
        total = int(lib.gsr_rows_exchange_bytes(cap, world))
        if total == 0:
            return None
        t = symm_mem.empty(total // 4, dtype=torch.float32, device=device)
        hdl = symm_mem.rendezvous(t, grp.group_name)
        if not getattr(hdl, "multicast_ptr", 0):
            return None
        t.zero_()
        torch.cuda.synchronize(device)
        hdl.barrier()
        torch.cuda.synchronize(device)
        offs = (C.c_size_t * 4)()
        wpr, vpr = C.c_int(0), C.c_int(0)
        _abi.check(lib.gsr_rows_exchange_layout(cap, world, offs, C.byref(wpr), C.byref(vpr)), "gsr_rows_exchange_layout")
        rank = dist.get_rank(self.group)
        self._lib, self._hdl, self._buf, self._name = lib, hdl, t, grp.group_name
        self._cap = cap
        self._rank, self._world = rank, world
        self._offs, self._wpr, self._vpr = [int(o) for o in offs], int(wpr.value), int(vpr.value)
        base_off = int(getattr(hdl, "offset", 0) or 0)
        self._local_ptr = t.data_ptr()
        self._mc_ptr = int(hdl.multicast_ptr) + base_off
        # device array of the peers' buffer addresses (+ this allocation's offset inside the symmetric block)
        self._peers = torch.tensor([int(p) + base_off for p in hdl.buffer_ptrs], dtype=torch.int64, device=device)
        self.marks = (self._local_ptr + self._offs[1] + 4 * rank * self._wpr,
                      self._local_ptr + self._offs[2] + rank * self._vpr)
        self.vis = torch.zeros(cap, dtype=torch.float32, device=device)
        self._status = torch.zeros(1, dtype=torch.int32, device=device)
        self._epoch = 1   # a fresh (zero-filled) block restarts the barrier epochs on every rank alike
        sms = torch.cuda.get_device_properties(device).multi_processor_count
        self._ctas = max(1, min(int(os.environ.get("GSR_XCHG_CTAS", "148")), sms, 256))
        forced = os.environ.get("GSR_XCHG", "auto")
        if self._decided_mode is not None and self._decided_mode != "all_reduce":
            # a mode was already chosen (and the sparse kernel verified) on a smaller block of this window
            self.candidates, self.mode, self.decided = [self._decided_mode], self._decided_mode, True
        else:
            self.candidates = []
            if forced in ("auto", "sparse"):
                self.candidates.append("sparse")
            if forced in ("auto", "multimem") and hasattr(torch.ops.symm_mem, "multimem_all_reduce_"):
                self.candidates.append("multimem")
            if forced in ("auto", "all_reduce") or not self.candidates:
                self.candidates.append("all_reduce")
            # never "decided" up front while the sparse kernel is a candidate: its result is always checked once
            # against a dense sum of the same rows (also when GSR_XCHG=sparse forces it)
            self.mode = self.candidates[0]
            self.decided = len(self.candidates) == 1 and self.candidates[0] != "sparse"
        return t[:P * N_COLS].view(P, N_COLS)

    def zero_marks(self):
        """A rank without keyframes contributes no rows: clear its bitmap / visibility slot."""
        if self._buf is not None:
            b = self._buf.view(torch.uint8)
            o1 = self._offs[1] + 4 * self._rank * self._wpr
            o2 = self._offs[2] + self._rank * self._vpr
            b[o1:o1 + 4 * self._wpr].zero_()
            b[o2:o2 + self._vpr].zero_()

    def vis_count(self, flat: torch.Tensor) -> torch.Tensor:
        """Number of keyframes (over all ranks) in which each Gaussian is visible, after the exchange."""
        return self.vis[:flat.shape[0]] if self.mode == "sparse" else flat[:, N_COLS - 1]

    # -- the exchange ----------------------------------------------------------------------------------------
    def _run(self, mode, flat, p0, p1):
        if mode == "sparse":
            if p1 < self._P:
                return None  # ranges: the sparse kernel runs once, when the last range is in place
            rc = self._lib.gsr_rows_exchange(torch.cuda.current_stream(flat.device).cuda_stream, self._P, self._cap,
                                             self._rank, self._world, self._ctas, self._local_ptr, self._mc_ptr,
                                             self._peers.data_ptr(), self._epoch, self.vis.data_ptr(),
                                             self._status.data_ptr())
            self._epoch = (self._epoch + 2) & 0xFFFFFFFF
            if rc != 0:
                raise RuntimeError(f"gsr_rows_exchange failed: {rc}")
            return None
        if mode == "multimem":
            torch.ops.symm_mem.multimem_all_reduce_(flat.view(-1)[p0 * N_COLS:p1 * N_COLS], "sum", self._name)
            return None
        return dist.all_reduce(flat[p0:p1], op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def _decide(self, flat: torch.Tensor):
        """First real exchange: time every candidate on the real rows, verify the sparse result, keep the fastest."""
        P, dev = self._P, flat.device
        b = self._buf.view(torch.uint8)
        regions = [(0, 4 * P * N_COLS), (self._offs[1] + 4 * self._rank * self._wpr, 4 * self._wpr),
                   (self._offs[2] + self._rank * self._vpr, self._vpr)]
        saved = [b[o:o + n].clone() for o, n in regions]

        def restore():
            for (o, n), s in zip(regions, saved):
                b[o:o + n].copy_(s)

        times, results = {}, {}
        trial = list(self.candidates)
        if trial == ["sparse"]:            # forced sparse: still verified once, against ncclAllReduce of the same rows
            trial.append("all_reduce")
        for mode in trial:
            ts = []
            for rep in range(4):
                restore()
                a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                w = self._run(mode, flat, 0, P)
                if w is not None:
                    w.wait()
                e.record()
                torch.cuda.synchronize(dev)
                if rep > 0:
                    ts.append(a.elapsed_time(e))
            times[mode] = sorted(ts)[len(ts) // 2]
            results[mode] = (flat.clone(), self.vis[:P].clone())
        ok = 1.0
        dense = next((m for m in ("multimem", "all_reduce") if m in results), None)
        if "sparse" in results and dense is not None:
            rs, vs = results["sparse"]
            rd, _ = results[dense]
            scale = float(rd[:, :N_COLS - 1].abs().max().clamp_min(1e-30))
            err = float((rs[:, :N_COLS - 1] - rd[:, :N_COLS - 1]).abs().max()) / scale
            verr = float((vs - rd[:, N_COLS - 1]).abs().max())
            self.report.update(sparse_vs_dense_rel_err=err, sparse_vis_err=verr)
            if not (err <= 1e-5 and verr == 0.0) or int(self._status.item()) != 0:
                ok = 0.0
        order = trial
        t = torch.tensor([times[m] for m in order] + [-ok], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        tl = [float(x) for x in t.tolist()]
        agreed = dict(zip(order, tl[:-1]))
        if -tl[-1] < 0.5:  # some rank saw a wrong sparse result: never use it
            agreed.pop("sparse", None)
            self.report["sparse_rejected"] = True
        allowed = {m: v for m, v in agreed.items() if m in self.candidates} or agreed
        self.mode = min(allowed, key=allowed.get)
        self.decided = True
        self._decided_mode = self.mode
        import numpy as np
        o = self._offs[1] // 4 + self._rank * self._wpr
        words = self._buf.view(torch.int32)[o:o + self._wpr].cpu().numpy()
        self.report.update(times_ms=agreed, mode=self.mode, rows=P,
                           touched_rows_local=int(np.unpackbits(words.view(np.uint8)).sum()))
        restore()

    def reduce(self, flat: torch.Tensor, p0: int, p1: int, marks_written: bool = True):
        """Start the sum of rows [p0, p1); returns a handle with .wait() or None when already stream-ordered.
        marks_written=False (a render backend that does not fill the touched-row bitmaps) rules the sparse kernel
        out: it would exchange nothing."""
        if not marks_written and "sparse" in self.candidates:
            self.candidates = [m for m in self.candidates if m != "sparse"] or ["all_reduce"]
            if self.mode == "sparse":
                self.mode, self.decided = self.candidates[0], len(self.candidates) == 1
        if not self.decided:
            if (p0, p1) != (0, self._P):   # ranges are not supported by the sparse kernel's trial: drop it
                self.candidates = [m for m in self.candidates if m != "sparse"] or ["all_reduce"]
                self.mode, self.decided = self.candidates[0], len(self.candidates) == 1
            if not self.decided:
                self._decide(flat)
        return self._run(self.mode, flat, p0, p1)


class MappingWindow:
    def __init__(self, render_fn: RenderFn, group: Optional[dist.ProcessGroup] = None, n_chunks: int = 1,
                 n_streams: Optional[int] = None):
        self.render_fn = render_fn
        self.group = group
        self.n_chunks = n_chunks
        # several local keyframes are issued round-robin on `n_streams` CUDA streams: one keyframe's pipeline is a
        # chain of short, mostly latency-bound kernels that leaves SMs idle, a second keyframe fills them (measured
        # on one B200, 8 keyframes: -7 % at the headline scene, -9 % at C3, -19 % at C2; 3+ streams add nothing).
        # Only the row accumulation is ordered across keyframes (render backends with `accepts_rows_gate`).
        self.n_streams = int(os.environ.get("GSR_WINDOW_STREAMS", "2")) if n_streams is None else int(n_streams)
        self._streams: List = []
        self.exchange = RowExchange(group)
        self._flat: Optional[torch.Tensor] = None
        self._radii: Optional[torch.Tensor] = None
        self._radii_buf: Optional[torch.Tensor] = None
        self._zero_stream = None

    @property
    def world_size(self) -> int:
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    @property
    def rank(self) -> int:
        return dist.get_rank(self.group) if dist.is_available() and dist.is_initialized() else 0

    def _buffers(self, P: int, device) -> None:
        if self._flat is None or self._flat.shape[0] != P or self._flat.device != torch.device(device):
            self._flat = self.exchange.allocate(P, device)   # same block while P stays inside its capacity
            if self._radii_buf is None or self._radii_buf.numel() < P or self._radii_buf.device != torch.device(device):
                self._radii_buf = torch.empty(self.exchange._cap or P, dtype=torch.int32, device=device)
            self._radii = self._radii_buf[:P]

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

        marks_written = bool(getattr(self.render_fn, "accepts_marks", False))

        def chunk_done(p0, p1):
            if world > 1:
                w = self.exchange.reduce(flat, p0, p1, marks_written=marks_written)
                if w is not None:
                    works.append(w)

        mine = shard_keyframes(n_keyframes, world, self.rank)
        local = []
        radii.zero_()
        if not mine:  # more GPUs than keyframes: contribute zeros, but take part in every collective
            flat.zero_()
            self.exchange.zero_marks()
            for c in chunks:
                chunk_done(*c)
        kwargs = {}
        if self.exchange.marks is not None and getattr(self.render_fn, "accepts_marks", False):
            kwargs["marks"] = self.exchange.marks
        # Sparse exchange decided: the per-Gaussian backward only writes the rows it has a gradient for (a few % of the
        # map per keyframe) into rows zero-filled HERE, on a side stream, while the forward renders -- instead of
        # streaming 64 bytes per Gaussian through the rows kernel in front of the exchange.
        zero_done = None
        if (world > 1 and mine and dev.type == "cuda" and self.exchange.decided and self.exchange.mode == "sparse"
                and "marks" in kwargs and getattr(self.render_fn, "accepts_touched_only", False) and len(chunks) == 1):
            if self._zero_stream is None:
                self._zero_stream = torch.cuda.Stream(device=dev)
            main0 = torch.cuda.current_stream(dev)
            self._zero_stream.wait_stream(main0)
            with torch.cuda.stream(self._zero_stream):
                flat.zero_()
                zero_done = torch.cuda.Event()
                zero_done.record(self._zero_stream)
            kwargs["touched_only"] = True
        n_streams = min(self.n_streams, len(mine))
        if (n_streams > 1 and dev.type == "cuda" and len(chunks) == 1
                and getattr(self.render_fn, "accepts_rows_gate", False)):
            # keyframes round-robin on side streams; the rows kernel of keyframe i+1 waits for keyframe i's
            while len(self._streams) < n_streams:
                self._streams.append(torch.cuda.Stream(device=dev))
            main = torch.cuda.current_stream(dev)
            side = self._streams[:n_streams]
            for st in side:
                st.wait_stream(main)
            gate = zero_done
            for i, k in enumerate(mine):
                st = side[i % n_streams]
                with torch.cuda.stream(st):
                    r = self.render_fn(k, params, flat, i > 0, chunks, None, rows_gate=gate, **kwargs)
                    gate = torch.cuda.Event()
                    gate.record(st)
                local.append(r)
            for st in side:
                main.wait_stream(st)
            for r in local:
                torch.maximum(radii, r.radii.to(torch.int32), out=radii)
            for c in chunks:
                chunk_done(*c)
        else:
            for i, k in enumerate(mine):
                last = i == len(mine) - 1
                if i == 0 and zero_done is not None:
                    kwargs["rows_gate"] = zero_done
                else:
                    kwargs.pop("rows_gate", None)
                r = self.render_fn(k, params, flat, i > 0, chunks, chunk_done if last else None, **kwargs)
                torch.maximum(radii, r.radii.to(torch.int32), out=radii)
                local.append(r)
        if world > 1:
            works.append(dist.all_reduce(radii, op=dist.ReduceOp.MAX, group=self.group, async_op=True))
            for w in works:
                w.wait()
        sl = column_slices()
        grads = {name: flat[:, sl[name]] for name, _ in PARAM_COLUMNS}
        return WindowStepResult(grads=grads, grad2d_norm=flat[:, sl["grad2d_norm"]].squeeze(1),
                                vis_count=self.exchange.vis_count(flat) if world > 1 else
                                flat[:, sl["vis_count"]].squeeze(1), max_radii=radii, local=local)


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
                        loss_grad: Optional[Callable] = None, fused_loss: Optional[bool] = None) -> RenderFn:
    """Product backend: the sm_100a rasterizer through the `_C` mirror (C ABI). The per-Gaussian half of the
    backward writes 64-byte window rows straight into `flat` (gsr_backward_rows), range by range.

    cameras[k] is a synth.Camera-like object with .matrices(device), .W, .H, .tanfovx, .tanfovy; targets[k] =
    (gt_image[3,H,W], gt_depth[1,H,W]).  loss_grad(color, depth, gt_image, gt_depth) -> (loss, dL_dcolor,
    dL_ddepth) is a custom loss evaluated with framework ops.  Without one (fused_loss defaults to True then) the
    mapping loss 0.9 L1(rgb) + 0.1 L1(depth over gt_depth > 0) (get_loss_mapping, utils/slam_utils.py:101-146) is
    reduced by ONE kernel (gsr_l1_loss_forward) and its gradient is formed inside the backward tile renderer
    (gsr_backward_render_l1): no gradient images, ~12 framework ops less per keyframe -- the Python window loop is
    host-bound at tracking-sized maps.
    """
    import ctypes as C

    from . import _abi
    from .diff_gaussian_rasterization import _C

    cached = {}
    lg = loss_grad or mapping_loss_grad
    fused = (loss_grad is None) if fused_loss is None else bool(fused_loss)
    if fused and loss_grad is not None:
        raise ValueError("fused_loss=True evaluates the built-in mapping loss; it cannot be combined with loss_grad")

    def render(k, params, flat, accumulate, chunks, chunk_done, marks=None, rows_gate=None,
               touched_only=False) -> KeyframeResult:
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
        kw = dict(chunks=chunks, chunk_done=chunk_done, marks=marks, rows_gate=rows_gate, touched_only=touched_only)
        if fused:
            lib = _abi.load()
            terms = torch.empty(8, dtype=torch.float32, device=dev)
            acc = torch.empty(8, dtype=torch.float64, device=dev)
            gi, gd = _C._f32(gt_image, "gt_image"), _C._f32(gt_depth, "gt_depth")
            with _C._on_device(dev):
                rc = lib.gsr_l1_loss_forward(_C._stream(), _abi.LOSS_MAPPING, cam.W, cam.H, color.data_ptr(),
                                             depth.data_ptr(), opacity.data_ptr(), gi.data_ptr(), gd.data_ptr(),
                                             None, None, None, None, 0.9, acc.data_ptr(), terms.data_ptr())
            _abi.check(rc, "gsr_l1_loss_forward")
            spec = _abi.L1LossSpec(_abi.LOSS_MAPPING, color.data_ptr(), depth.data_ptr(), opacity.data_ptr(),
                                   gi.data_ptr(), gd.data_ptr(), None, None, None, None, terms.data_ptr(), None)
            tau, _ = _C.rasterize_gaussians_backward_rows(
                bg, params["means3D"], radii, params["scales"], params["rotations"], 1.0, view, proj, proj_raw,
                cam.tanfovx, cam.tanfovy, None, None, geom, R, plist, img, flat, accumulate, l1_spec=spec,
                image_hw=(cam.H, cam.W), **kw)
            loss = terms[0]
        else:
            loss, dcol, ddep = lg(color, depth, gt_image, gt_depth)
            tau, _ = _C.rasterize_gaussians_backward_rows(
                bg, params["means3D"], radii, params["scales"], params["rotations"], 1.0, view, proj, proj_raw,
                cam.tanfovx, cam.tanfovy, dcol.contiguous(), ddep.contiguous(), geom, R, plist, img, flat, accumulate,
                **kw)
        return KeyframeResult(index=k, radii=radii, n_touched=n_touched, grad_tau=tau, loss=loss)

    render.accepts_marks = True
    render.accepts_rows_gate = True
    render.accepts_touched_only = True
    return render
