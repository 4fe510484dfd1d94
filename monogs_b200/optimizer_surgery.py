"""Map growth and pruning next to the rasterizer (second half): drop-ins for the tensor / Adam-state
surgery of the reference's GaussianModel (gaussian_splatting/scene/gaussian_model.py):

    cat_tensors_to_optimizer(model, tensors_dict)      :697-728
    densification_postfix(model, new_xyz, ...)          :730-778
    _prune_optimizer(model, mask)                       :656-677
    prune_points(model, mask)                           :679-695

Same arguments (with the model passed explicitly; `install(model_or_class)` binds them as methods), same resulting
state -- parameter values, `optimizer.param_groups[i]["params"][0]` replaced by a fresh nn.Parameter, its Adam state
(`step`, `exp_avg`, `exp_avg_sq`) moved over with zero moments for new rows, `_obj_prob`, `xyz_gradient_accum`,
`denom`, `max_radii_2d`, `kf_idx`, `nr_obs` -- bit for bit (tests/test_gpu_surgery.py against golden vectors the
reference functions produced, tests/golden/make_surgery_golden.py).

What differs is the cost.  The reference re-creates every tensor: 15 `torch.cat` per insertion that copy all P old rows
to append K, and ~19 boolean-mask indexings per prune (each a nonzero + gather + host sync).  Here the rows of every
array live in a CAPACITY-backed buffer (grow-only, 1.5x), parameters and moments are views `buf[:P]`:
  * append  = ONE launch of gsr_rows_append writing K rows per group (+ zero moments): O(K), no copy of old rows;
  * prune   = ONE stable compaction (gsr_rows_compact) of all device arrays by the keep mask into the alternate
              buffers, one host read of the surviving count.
The CUDA library is required (no fallback); CPU-resident bookkeeping tensors (`kf_idx`, `nr_obs`, which the reference
keeps on the CPU) are handled with the reference's own torch expressions.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch
from torch import nn

from . import _abi
from .diff_gaussian_rasterization import _C

GROUPS = ("xyz", "f_dc", "opacity", "scaling", "rotation")          # optimizer group names, training_setup :400-437
ATTRS = dict(xyz="_xyz", f_dc="_features_dc", opacity="_opacity", scaling="_scaling", rotation="_rotation")
STATS = ("xyz_gradient_accum", "denom", "max_radii_2d")


class _Store:
    """Capacity-backed storage of one row array; `other` is the compaction target (allocated on first prune)."""

    def __init__(self, like: torch.Tensor, cap: int):
        self.tail = tuple(like.shape[1:])
        self.buf = torch.empty((cap,) + self.tail, dtype=like.dtype, device=like.device)
        self.other: Optional[torch.Tensor] = None

    @property
    def cap(self) -> int:
        return int(self.buf.shape[0])

    @property
    def row_bytes(self) -> int:
        n = 1
        for d in self.tail:
            n *= int(d)
        return n * self.buf.element_size()

    def view(self, n: int) -> torch.Tensor:
        return self.buf[:n]

    def owns(self, t: torch.Tensor) -> bool:
        return t is not None and t.device == self.buf.device and t.dtype == self.buf.dtype and \
            t.numel() > 0 and t.data_ptr() == self.buf.data_ptr() and tuple(t.shape[1:]) == self.tail

    def adopt(self, t: torch.Tensor, n: int) -> None:
        if n and not self.owns(t):
            self.buf[:n].copy_(t.detach().reshape((n,) + self.tail))

    def grow(self, cap: int, n: int) -> None:
        new = torch.empty((cap,) + self.tail, dtype=self.buf.dtype, device=self.buf.device)
        new[:n].copy_(self.buf[:n])
        self.buf, self.other = new, None

    def swap(self) -> None:
        self.buf, self.other = self.other, self.buf

    def target(self) -> torch.Tensor:
        if self.other is None or self.other.shape != self.buf.shape:
            self.other = torch.empty_like(self.buf)
        return self.other


def _capacity(n: int, old: int = 0) -> int:
    """Grow-only row capacity: 25 % headroom over the request, at least 1.5x the previous capacity."""
    want = max(n + n // 4, old + old // 2, 4096)
    return (want + 4095) // 4096 * 4096


class _State:
    """Per-model storage: stores[name] for parameters, exp_avg/<name>, exp_avg_sq/<name>, obj_prob and the statistics."""

    def __init__(self):
        self.stores: Dict[str, _Store] = {}
        self.stats: Optional[torch.Tensor] = None   # f32[3, capacity]: xyz_gradient_accum | denom | max_radii_2d
        self.ws: Optional[torch.Tensor] = None
        self.count: Optional[torch.Tensor] = None

    def get(self, key: str, like: torch.Tensor, n: int, need: int) -> _Store:
        st = self.stores.get(key)
        if st is None or st.buf.device != like.device or st.tail != tuple(like.shape[1:]) or st.buf.dtype != like.dtype:
            st = _Store(like, _capacity(need))
            self.stores[key] = st
        if st.cap < need:
            if st.owns(like):
                st.grow(_capacity(need, st.cap), n)
            else:
                st = _Store(like, _capacity(need, st.cap))
                self.stores[key] = st
        st.adopt(like, n)
        return st


def _state(model) -> _State:
    s = getattr(model, "_gsr_surgery", None)
    if s is None:
        s = _State()
        model._gsr_surgery = s
    return s


def _groups(model):
    return {g["name"]: g for g in model.optimizer.param_groups}


def _ptr_table(ptrs):
    return (C.c_void_p * len(ptrs))(*[p if p else None for p in ptrs])


def _rebind(model, group, new_param, exp_avg=None, exp_avg_sq=None):
    """Replace the group's parameter object and move its Adam state over (gaussian_model.py:707-723 / :661-671)."""
    old = group["params"][0]
    stored = model.optimizer.state.get(old, None)
    if stored is not None:
        if exp_avg is not None:
            stored["exp_avg"], stored["exp_avg_sq"] = exp_avg, exp_avg_sq
        del model.optimizer.state[old]
        group["params"][0] = new_param
        model.optimizer.state[new_param] = stored
    else:
        group["params"][0] = new_param


def cat_tensors_to_optimizer(model, tensors_dict):
    """gaussian_model.py:697-728 -- append `tensors_dict[name]` ([K, ...]) to every optimizer group."""
    lib = _abi.load()
    st = _state(model)
    groups = _groups(model)
    names = [g["name"] for g in model.optimizer.param_groups]
    P = int(groups[names[0]]["params"][0].shape[0])
    K = int(tensors_dict[names[0]].shape[0])
    dev = groups[names[0]]["params"][0].device
    if dev.type != "cuda":
        raise RuntimeError("optimizer_surgery: the model must live on a CUDA device (no CPU path)")
    par, m1, m2, rows, cols, keep = [], [], [], [], [], []
    stores = {}
    for name in names:
        g = groups[name]
        assert len(g["params"]) == 1
        p = g["params"][0]
        ext = _C._f32(tensors_dict[name].detach(), name)
        like = (lambda t: t if P > 0 else tensors_dict[name].detach()[:0])   # an empty map has shape-[0] tensors
        sp = st.get(name, like(p.data), P, P + K)
        stored = model.optimizer.state.get(p, None)
        sm = sv = None
        if stored is not None:
            sm = st.get("exp_avg/" + name, like(stored["exp_avg"]), P, P + K)
            sv = st.get("exp_avg_sq/" + name, like(stored["exp_avg_sq"]), P, P + K)
        stores[name] = (sp, sm, sv)
        par.append(sp.buf.data_ptr()); m1.append(sm.buf.data_ptr() if sm else 0); m2.append(sv.buf.data_ptr() if sv else 0)
        rows.append(ext.data_ptr() if ext is not None else 0)
        cols.append(sp.row_bytes // 4)
        keep.append(ext)
    if K > 0:
        with _C._on_device(dev):
            rc = lib.gsr_rows_append(_C._stream(), len(names), _ptr_table(par), _ptr_table(m1), _ptr_table(m2),
                                     _ptr_table(rows), (C.c_int * len(cols))(*cols), P, K)
        _abi.check(rc, "gsr_rows_append")
    out = {}
    for name in names:
        sp, sm, sv = stores[name]
        new = nn.Parameter(sp.view(P + K).requires_grad_(True))
        _rebind(model, groups[name], new, sm.view(P + K) if sm else None, sv.view(P + K) if sv else None)
        out[name] = new
    return out


def densification_postfix(model, new_xyz, new_features_dc, new_opacities, new_scaling, new_rotation, new_obj_prob,
                          new_kf_idxs=None, new_nr_obs=None):
    """gaussian_model.py:730-778."""
    d = {"xyz": new_xyz, "f_dc": new_features_dc, "opacity": new_opacities, "scaling": new_scaling,
         "rotation": new_rotation}
    opt = cat_tensors_to_optimizer(model, d)
    for name, attr in ATTRS.items():
        setattr(model, attr, opt[name])
    model._obj_prob = _append_plain(model, "obj_prob", model._obj_prob, new_obj_prob)
    n = int(model._xyz.shape[0])
    dev = model._xyz.device
    # the three statistics restart from zero for ALL rows (:768-772): one memset of one [3, capacity] block
    st = _state(model)
    if st.stats is None or st.stats.shape[1] < n or st.stats.device != dev:
        st.stats = torch.empty(3, _capacity(n, 0 if st.stats is None else int(st.stats.shape[1])),
                               dtype=torch.float32, device=dev)
    st.stats[:, :n].zero_()
    model.xyz_gradient_accum = st.stats[0, :n].view(n, 1)
    model.denom = st.stats[1, :n].view(n, 1)
    model.max_radii_2d = st.stats[2, :n]
    if new_kf_idxs is not None:
        model.kf_idx = torch.cat((model.kf_idx, new_kf_idxs.to(model.kf_idx.device))).int()
    if new_nr_obs is not None:
        model.nr_obs = torch.cat((model.nr_obs, new_nr_obs.to(model.nr_obs.device))).int()


def _append_plain(model, key, old, new):
    """`torch.cat((old, new))` of a non-optimizer array (`_obj_prob`, :765) into its capacity buffer."""
    P, K = int(old.shape[0]), int(new.shape[0])
    if P == 0 and old.dim() != new.dim():
        old = new.detach()[:0]
    sp = _state(model).get(key, old.detach(), P, P + K)
    if K:
        sp.buf[P:P + K].copy_(new.detach())
    return sp.view(P + K)   # a plain tensor, like the reference's torch.cat result


def _compact(model, arrays: Dict[str, torch.Tensor], keep_mask: torch.Tensor):
    """Stable compaction of the device row arrays by keep_mask in one launch -> ({key: compacted view}, n_kept)."""
    lib = _abi.load()
    st = _state(model)
    P = int(keep_mask.shape[0])
    dev = keep_mask.device
    keys = list(arrays)
    stores = [st.get(k, arrays[k].detach(), P, P) for k in keys]
    src = [s.buf.data_ptr() for s in stores]
    dst = [s.target().data_ptr() for s in stores]
    rb = [s.row_bytes for s in stores]
    need = int(lib.gsr_surgery_ws_bytes(P))
    if st.ws is None or st.ws.numel() < need or st.ws.device != dev:
        st.ws = torch.empty(need, dtype=torch.uint8, device=dev)
        st.count = torch.zeros(1, dtype=torch.int32, device=dev)
    keep_u8 = keep_mask.contiguous().view(torch.uint8)
    with _C._on_device(dev):
        rc = lib.gsr_rows_compact(_C._stream(), len(keys), _ptr_table(src), _ptr_table(dst), (C.c_int * len(rb))(*rb),
                                  P, keep_u8.data_ptr(), st.count.data_ptr(), st.ws.data_ptr())
    _abi.check(rc, "gsr_rows_compact")
    n = int(st.count.item())          # the one host read of a prune (the reference syncs once per indexed tensor)
    out = {}
    for k, s in zip(keys, stores):
        s.swap()
        out[k] = s.view(n)
    return out, n


def _prune_gather(model, mask, extra: Dict[str, torch.Tensor]):
    groups = _groups(model)
    arrays = {}
    for g in model.optimizer.param_groups:
        p = g["params"][0]
        arrays[g["name"]] = p.data
        stored = model.optimizer.state.get(p, None)
        if stored is not None:
            arrays["exp_avg/" + g["name"]] = stored["exp_avg"]
            arrays["exp_avg_sq/" + g["name"]] = stored["exp_avg_sq"]
    arrays.update(extra)
    if mask.dtype != torch.bool:
        mask = mask != 0
    res, n = _compact(model, arrays, mask)
    out = {}
    for g in model.optimizer.param_groups:
        name = g["name"]
        new = nn.Parameter(res[name].requires_grad_(True))
        _rebind(model, groups[name], new, res.get("exp_avg/" + name), res.get("exp_avg_sq/" + name))
        out[name] = new
    return out, res, n


def _prune_optimizer(model, mask):
    """gaussian_model.py:656-677 -- keep the rows where `mask` is True in every optimizer group."""
    return _prune_gather(model, mask, {})[0]


def prune_points(model, mask):
    """gaussian_model.py:679-695 -- remove the Gaussians where `mask` is True."""
    valid = ~mask
    extra = {"xyz_gradient_accum_": model.xyz_gradient_accum, "obj_prob": model._obj_prob, "denom_": model.denom,
             "max_radii_2d_": model.max_radii_2d}
    opt, res, n = _prune_gather(model, valid, extra)
    for name, attr in ATTRS.items():
        setattr(model, attr, opt[name])
    model.xyz_gradient_accum = res["xyz_gradient_accum_"]
    model._obj_prob = res["obj_prob"]
    model.denom = res["denom_"]
    model.max_radii_2d = res["max_radii_2d_"]
    # the reference keeps these two on the CPU and indexes them there (:693-694)
    vk = valid.to(model.kf_idx.device)
    model.kf_idx = model.kf_idx[vk]
    model.nr_obs = model.nr_obs[vk]


def install(target):
    """Bind the four functions as methods of a GaussianModel instance or class, replacing the reference's."""
    import types
    fns = dict(cat_tensors_to_optimizer=cat_tensors_to_optimizer, densification_postfix=densification_postfix,
               _prune_optimizer=_prune_optimizer, prune_points=prune_points)
    for name, fn in fns.items():
        setattr(target, name, fn if isinstance(target, type) else types.MethodType(fn, target))
    return target
