"""Multi-GPU host logic of the sharded mapping window on CPU: world_size 2, gloo, 127.0.0.1.
The render backend is injected: here it is the CPU oracle (test infrastructure), so what is tested is the
partitioning, the flat gradient buffer, the sum / max exchange and that 2 ranks reproduce the 1-rank result."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT
from monogs_b200 import synth, window

N_KF = 5
P = 6000


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_render_fn(cams, params_np, bg):
    from oracle import cpu_oracle as co
    sl = window.column_slices()

    def render(k, params, flat, accumulate, chunks, chunk_done):
        if not accumulate:
            flat.zero_()
        cam = cams[k]
        view, proj, proj_raw, campos = [t.numpy() for t in cam.matrices()]
        n = lambda t: t.numpy()
        f = co.forward(n(params["means3D"]), n(params["scales"]), n(params["rotations"]), n(params["opacities"]),
                       n(params["colors"]), view, proj, campos, cam.W, cam.H, cam.tanfovx, cam.tanfovy, bg)
        g = torch.Generator().manual_seed(100 + k)
        dcol = torch.randn(3, cam.H, cam.W, generator=g).numpy() / (cam.H * cam.W)
        ddep = torch.randn(1, cam.H, cam.W, generator=g).numpy() / (cam.H * cam.W)
        b = co.backward(f, n(params["means3D"]), n(params["scales"]), n(params["rotations"]), view, proj, proj_raw,
                        cam.W, cam.H, cam.tanfovx, cam.tanfovy, bg, dcol, ddep)
        t = torch.from_numpy
        flat[:, sl["means3D"]] += t(b["means3D"])
        flat[:, sl["colors"]] += t(b["colors"])
        flat[:, sl["opacities"]] += t(b["opacity"])
        flat[:, sl["scales"]] += t(b["scales"])
        flat[:, sl["rotations"]] += t(b["rotations"])
        vis = t(f["radii"] > 0)
        flat[:, sl["grad2d_norm"]] += t(b["means2D"])[:, :2].norm(dim=1, keepdim=True) * vis[:, None]
        flat[:, sl["vis_count"]] += vis[:, None].float()
        if chunk_done is not None:
            for c in chunks:
                chunk_done(*c)
        return window.KeyframeResult(k, t(f["radii"]), t(f["n_touched"]), t(b["tau_sum"].astype(np.float32)))

    return render


def _scene():
    synth.INTRINSICS.setdefault("tiny", (160, 120, 133.85, 134.8, 80.03, 61.9))
    cams = synth.arc_cameras("tiny", N_KF)
    sc = synth.make_scene(P, cams[N_KF // 2], seed=3)
    return cams, sc


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cams, sc = _scene()
    bg = np.array([0.1, 0.2, 0.3], np.float32)
    w = window.MappingWindow(_oracle_render_fn(cams, None, bg), n_chunks=3)
    res = w.step(sc, N_KF)
    assert [r.index for r in res.local] == window.shard_keyframes(N_KF, world, rank)
    torch.save(dict(grads={k: v.clone() for k, v in res.grads.items()}, g2=res.grad2d_norm.clone(),
                    vc=res.vis_count.clone(), mr=res.max_radii.clone(),
                    taus={r.index: r.grad_tau for r in res.local}), f"{out_dir}/rank{rank}.pt")
    dist.barrier()
    dist.destroy_process_group()


def test_shard_keyframes():
    assert window.shard_keyframes(8, 1, 0) == list(range(8))
    assert window.shard_keyframes(8, 4, 1) == [1, 5]
    assert window.shard_keyframes(3, 8, 5) == []          # more GPUs than keyframes: idle rank
    got = sorted(k for r in range(3) for k in window.shard_keyframes(8, 3, r))
    assert got == list(range(8))                           # a partition
    with pytest.raises(ValueError):
        window.shard_keyframes(8, 2, 2)
    assert window.N_COLS == 16 and window.column_slices()["rotations"] == slice(10, 14)
    # ranges: 256-aligned, contiguous cover of [0, P), shrinking, identical on every rank
    for P in (1, 255, 4096, 100_001, 1_000_000):
        ch = window.window_chunks(P, 4)
        assert ch[0][0] == 0 and ch[-1][1] == P and all(a[1] == b[0] for a, b in zip(ch[:-1], ch[1:]))
        assert all(a % 256 == 0 for a, _ in ch)
    ch = window.window_chunks(1_000_000, 4)
    assert len(ch) == 4 and (ch[0][1] - ch[0][0]) > (ch[-1][1] - ch[-1][0])


def test_two_ranks_equal_one_rank(built, tmp_path):
    cams, sc = _scene()
    bg = np.array([0.1, 0.2, 0.3], np.float32)
    single = window.MappingWindow(_oracle_render_fn(cams, None, bg)).step(sc, N_KF)
    ref = dict(grads={k: v.clone() for k, v in single.grads.items()}, g2=single.grad2d_norm.clone(),
               vc=single.vis_count.clone(), mr=single.max_radii.clone())
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(tmp_path / "rank0.pt")
    r1 = torch.load(tmp_path / "rank1.pt")
    for k in ref["grads"]:
        assert torch.equal(r0["grads"][k], r1["grads"][k])                       # replicas identical
        scale = ref["grads"][k].abs().max().clamp_min(1e-30)
        assert ((r0["grads"][k] - ref["grads"][k]).abs().max() / scale) <= 1e-5   # summation order only
    assert torch.equal(r0["vc"], ref["vc"]) and torch.equal(r0["mr"], ref["mr"]) and torch.equal(r1["mr"], ref["mr"])
    assert torch.allclose(r0["g2"], ref["g2"], rtol=1e-5, atol=1e-12)
    assert sorted(list(r0["taus"]) + list(r1["taus"])) == list(range(N_KF))       # pose grads stay with their owner
    assert float(ref["vc"].max()) <= N_KF and float(ref["vc"].max()) >= 1


def test_row_buffer_capacity_is_grow_only():
    """the rows buffer must not be re-allocated on every change of P (keyframe insertion / densification
    change P every few iterations): capacity grows by >= 1.5x, rows views of a smaller or slightly larger map share
    the same storage, and a decided exchange mode survives."""
    from monogs_b200 import window
    ex = window.RowExchange()
    a = ex.allocate(5000, "cpu")
    assert a.shape == (5000, window.N_COLS) and ex._cap == 8192
    ptr = a.untyped_storage().data_ptr()
    b = ex.allocate(6000, "cpu")
    assert b.shape == (6000, window.N_COLS) and b.untyped_storage().data_ptr() == ptr and ex._cap == 8192
    c = ex.allocate(4000, "cpu")                        # pruning shrinks the map: still the same block
    assert c.untyped_storage().data_ptr() == ptr
    d = ex.allocate(9000, "cpu")                        # outgrown: one re-allocation to 1.5x
    assert ex._cap == 12288 and d.shape == (9000, window.N_COLS) and d.untyped_storage().data_ptr() != ptr
    assert ex.mode == "all_reduce" and ex.decided
    w = window.MappingWindow(lambda *a, **k: None)
    w._buffers(5000, "cpu")
    f0 = w._flat.untyped_storage().data_ptr()
    w._buffers(5100, "cpu")
    assert w._flat.shape[0] == 5100 and w._flat.untyped_storage().data_ptr() == f0 and w._radii.shape[0] == 5100
