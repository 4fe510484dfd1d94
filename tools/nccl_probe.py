"""tools/nccl_probe.py -- allreduce latency vs size on this box (torchrun), CUDA-event timed, max over ranks."""
import os, sys, json, torch, torch.distributed as dist
local = int(os.environ.get("LOCAL_RANK", 0)); torch.cuda.set_device(local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", local)
def timeit(fn, n=20):
    for _ in range(5): fn()
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / n], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
out = {}
for mb in (1, 4, 8, 16, 32, 56, 64, 128):
    x = torch.zeros(mb * 1024 * 1024 // 4, device=dev)
    out[f"allreduce_{mb}MB_ms"] = round(timeit(lambda: dist.all_reduce(x)), 4)
x = torch.zeros(64 * 1024 * 1024 // 4, device=dev)
def four():
    ws = [dist.all_reduce(x[a:b], async_op=True) for a, b in ((0, 6710886), (6710886, 11744051), (11744051, 15099494), (15099494, 16777216))]
    for w in ws: w.wait()
out["allreduce_64MB_in_4_async_pieces_ms"] = round(timeit(four), 4)
y = torch.zeros(64 * 1024 * 1024 // 4, device=dev)
def rs_ag():
    o = torch.empty(y.numel() // world, device=dev)
    dist.reduce_scatter_tensor(o, y); dist.all_gather_into_tensor(y, o)
out["reduce_scatter+all_gather_64MB_ms"] = round(timeit(rs_ag), 4)
try:
    import torch.distributed._symmetric_memory as symm_mem
    gname = dist.group.WORLD.group_name
    for mb in (16, 56, 64):
        t = symm_mem.empty(mb * 1024 * 1024 // 4, device=dev, dtype=torch.float32)
        hdl = symm_mem.rendezvous(t, gname)
        t.zero_()
        for opname in ("multimem_all_reduce_", "two_shot_all_reduce_", "one_shot_all_reduce"):
            op = getattr(torch.ops.symm_mem, opname, None)
            if op is None:
                out[f"symm_{opname}_{mb}MB_ms"] = "absent"; continue
            try:
                out[f"symm_{opname}_{mb}MB_ms"] = round(timeit(lambda: op(t, "sum", gname)), 4)
            except Exception as e:
                out[f"symm_{opname}_{mb}MB_ms"] = "err: " + str(e)[:80]
    out["multicast_ptr_nonzero"] = bool(getattr(hdl, "multicast_ptr", 0))
except Exception as e:
    out["symm_mem"] = "unavailable: " + str(e)[:200]
out["NCCL_ALGO"] = os.environ.get("NCCL_ALGO", "default")
if rank == 0: print(json.dumps({"world": world, **out}))
dist.destroy_process_group()
