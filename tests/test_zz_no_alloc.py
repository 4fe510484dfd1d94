"""gsr_insert_select_static (include/gsr_b200.h): the allocation-free variant of gsr_insert_select whose scratch is a
`static __device__` array inside the library.  No counterpart in the reference,
so the check is against the existing entry point: selected indices and count must be IDENTICAL (integer
work, the tolerance of tests/test_insert.py for indices is exact), on two calls in a row -- with a different image the
second time, so that stale scratch contents from the first call would show -- and the size limit must be refused with
GSR_ESTATICLIMIT before anything is launched.  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from monogs_b200 import _abi

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _images(W, H, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    gt = (0.2 + 4.0 * torch.rand(H, W, generator=g)) * (torch.rand(H, W, generator=g) > 0.1).float()
    rd = gt + 0.4 * torch.randn(H, W, generator=g)
    ro = torch.rand(H, W, generator=g)
    med = torch.tensor([0.004 * seed])
    return [t.to(DEV).contiguous() for t in (gt, rd, ro, med)]


def _run(lib, static, W, H, gt, rd, ro, med, init):
    sel = torch.full((W * H,), -1, dtype=torch.int32, device=DEV)
    count = torch.zeros(1, dtype=torch.int32, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    args = (st, W, H, gt.data_ptr(), rd.data_ptr(), ro.data_ptr(), int(init), med.data_ptr(), sel.data_ptr(),
            count.data_ptr())
    if static:
        rc = lib.gsr_insert_select_static(*args)
        _abi.check(rc, "gsr_insert_select_static")
    else:
        ws = torch.empty(int(lib.gsr_insert_ws_bytes(W, H)), dtype=torch.uint8, device=DEV)
        rc = lib.gsr_insert_select(*args, ws.data_ptr())
        _abi.check(rc, "gsr_insert_select")
    torch.cuda.synchronize()
    return sel, int(count.item())


def test_static_scratch_entry_point_matches_workspace_entry_point(built):
    lib = _abi.load()
    # two calls in a row: a larger image first, then a smaller, different one (fewer blocks than the scratch still
    # holds counts for), then the first again
    for k, (W, H, seed, init) in enumerate([(640, 480, 3, False), (333, 257, 5, False), (640, 480, 7, True),
                                            (640, 480, 3, False)]):
        imgs = _images(W, H, seed)
        want_sel, want_n = _run(lib, False, W, H, *imgs, init)
        sel = torch.full((W * H,), -1, dtype=torch.int32, device=DEV)
        count = torch.zeros(1, dtype=torch.int32, device=DEV)
        free_mid, allocated_mid = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        rc = lib.gsr_insert_select_static(st, W, H, imgs[0].data_ptr(), imgs[1].data_ptr(), imgs[2].data_ptr(),
                                          int(init), imgs[3].data_ptr(), sel.data_ptr(), count.data_ptr())
        _abi.check(rc, "gsr_insert_select_static")
        torch.cuda.synchronize()
        if k > 0:   # (the very first call may load the module image; later calls must not move device memory at all)
            assert torch.cuda.mem_get_info()[0] == free_mid and torch.cuda.memory_allocated() == allocated_mid
        n = int(count.item())
        assert n == want_n and 0 < n < W * H
        assert torch.equal(sel[:n], want_sel[:n])
        assert bool((sel[n:] == -1).all())                      # nothing written past the count
        assert bool((sel[1:n] > sel[:n - 1]).all())             # ascending flat indices (stable compaction)


def test_static_scratch_entry_point_refuses_sizes_beyond_its_array(built):
    lib = _abi.load()
    W = H = 2049                                                # 2049^2 > 4 Mi pixels
    assert W * H > _abi.INSERT_STATIC_MAX_PIXELS
    one = torch.zeros(1, device=DEV)
    cnt = torch.full((1,), 123, dtype=torch.int32, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # refused on the sizes alone, before any pointer is dereferenced or any kernel launched
    rc = lib.gsr_insert_select_static(st, W, H, one.data_ptr(), one.data_ptr(), one.data_ptr(), 0, one.data_ptr(),
                                      cnt.data_ptr(), cnt.data_ptr())
    assert rc == _abi.GSR_ESTATICLIMIT
    with pytest.raises(RuntimeError, match="GSR_ESTATICLIMIT"):
        _abi.check(rc, "gsr_insert_select_static")
    torch.cuda.synchronize()
    assert int(cnt.item()) == 123
    # the largest size it accepts: 2048 x 2048
    W = H = 2048
    imgs = _images(W, H, 11)
    want_sel, want_n = _run(lib, False, W, H, *imgs, False)
    sel, n = _run(lib, True, W, H, *imgs, False)
    assert n == want_n and torch.equal(sel[:n], want_sel[:n])
