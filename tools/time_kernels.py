"""tools/time_kernels.py -- CUDA-event timing of forward-plan / forward-render / backward through the C ABI."""
import argparse, sys, ctypes as C
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import synth, _abi
ap = argparse.ArgumentParser(); ap.add_argument("--P", type=int, default=1_000_000); ap.add_argument("--kind", default="tum")
ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
dev = "cuda"; lib = _abi.load()
cam = synth.make_camera(a.kind, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
sc = synth.make_scene(a.P, cam, seed=0, device=dev)
view, proj, proj_raw, campos = cam.matrices(dev)
bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
P, W, H = a.P, cam.W, cam.H
radii = torch.empty(P, dtype=torch.int32, device=dev)
geom = torch.empty(lib.gsr_geom_bytes(P, W, H), dtype=torch.uint8, device=dev)
img = torch.empty(lib.gsr_image_bytes(W, H), dtype=torch.uint8, device=dev)
R = C.c_int64(0); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
def plan():
    rc = lib.gsr_forward_plan(st, P, 0, 0, W, H, sc["means3D"].data_ptr(), None, sc["colors"].data_ptr(), sc["opacities"].data_ptr(),
        sc["scales"].data_ptr(), 1.0, sc["rotations"].data_ptr(), None, view.data_ptr(), proj.data_ptr(), campos.data_ptr(),
        cam.tanfovx, cam.tanfovy, 0, radii.data_ptr(), geom.data_ptr(), C.byref(R), 0)
    assert rc == 0, rc
def timeit(fn, n):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(n): fn()
    t1.record(); torch.cuda.synchronize(); return t0.elapsed_time(t1) / n * 1000
print("plan_us", round(timeit(plan, a.iters), 1), "R", R.value)
plan()
Rv = R.value
plist = torch.empty(lib.gsr_point_list_bytes(Rv), dtype=torch.uint8, device=dev)
binning = torch.empty(lib.gsr_binning_bytes(Rv), dtype=torch.uint8, device=dev)
col = torch.empty(3, H, W, device=dev); dep = torch.empty(1, H, W, device=dev); opa = torch.empty(1, H, W, device=dev)
nt = torch.zeros(P, dtype=torch.int32, device=dev)
def render():
    rc = lib.gsr_forward_render(st, P, W, H, bg.data_ptr(), radii.data_ptr(), geom.data_ptr(), img.data_ptr(), binning.data_ptr(),
        plist.data_ptr(), Rv, col.data_ptr(), dep.data_ptr(), opa.data_ptr(), nt.data_ptr(), 0)
    assert rc == 0, rc
print("render_phase_us", round(timeit(render, a.iters), 1))
g = torch.Generator().manual_seed(7)
dcol = (torch.randn(3, H, W, generator=g) / (H * W)).to(dev); ddep = (torch.randn(1, H, W, generator=g) / (H * W)).to(dev)
scratch = torch.empty(lib.gsr_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
o = {k: torch.empty(P, n, device=dev) for k, n in dict(m2=3, c=3, o=1, m3=3, s=3, r=4).items()}
tau = torch.zeros(6, device=dev)
def bwd():
    rc = lib.gsr_backward(st, P, 0, 0, W, H, bg.data_ptr(), sc["means3D"].data_ptr(), None, sc["colors"].data_ptr(), sc["scales"].data_ptr(), 1.0,
        sc["rotations"].data_ptr(), None, view.data_ptr(), proj.data_ptr(), proj_raw.data_ptr(), campos.data_ptr(), cam.tanfovx, cam.tanfovy,
        radii.data_ptr(), geom.data_ptr(), img.data_ptr(), plist.data_ptr(), Rv, dcol.data_ptr(), ddep.data_ptr(), scratch.data_ptr(),
        o["m2"].data_ptr(), o["c"].data_ptr(), o["o"].data_ptr(), o["m3"].data_ptr(), None, None, o["s"].data_ptr(), o["r"].data_ptr(), None,
        tau.data_ptr(), 0)
    assert rc == 0, rc
print("backward_us", round(timeit(bwd, a.iters), 1))

# --- per-CTA timeline of the render kernels (library variant built with -DGSR_CTA_TRACE) --------------------
if hasattr(lib, "gsr_debug_cta_trace_render_bwd") or "--trace" in sys.argv:
    import numpy as np
    raw = C.CDLL(str(_abi._LIB_PATH))
    sub = int(__import__("os").environ.get("TRACE_SUB", "1"))
    n_cta = ((W + 15) // 16) * ((H + 15) // 16) * sub
    for name, fn in (("render_fwd", render), ("render_bwd", bwd)):
        torch.cuda.synchronize(); fn(); torch.cuda.synchronize()
        buf = np.zeros(3 * n_cta, dtype=np.uint64)
        rc = getattr(raw, "gsr_debug_cta_trace_" + name)(buf.ctypes.data_as(C.c_void_p), n_cta); assert rc == 0, rc
        t0, t1, x = buf[0::3].astype(np.int64), buf[1::3].astype(np.int64), buf[2::3]
        smid = (x & np.uint64(0xffffffff)).astype(np.int64); waited = (x >> np.uint64(32)).astype(np.int64)
        base = t0.min(); span = t1.max() - base; dur = t1 - t0
        print(f"[{name}] ctas {n_cta} span_us {span/1e3:.1f} cta_dur_us mean {dur.mean()/1e3:.2f} p50 {np.median(dur)/1e3:.2f} "
              f"p90 {np.percentile(dur,90)/1e3:.2f} max {dur.max()/1e3:.2f} pdl_wait_us mean {waited.mean()/1e3:.2f} max {waited.max()/1e3:.2f}")
        # concurrency over time in 20 slices
        edges = np.linspace(0, span, 21)
        conc = [(np.minimum(t1 - base, e1) - np.maximum(t0 - base, e0)).clip(0).sum() / max(e1 - e0, 1) for e0, e1 in zip(edges[:-1], edges[1:])]
        print(f"[{name}] resident CTAs per 5% slice:", " ".join(f"{c:.0f}" for c in conc))
        busy = np.zeros(smid.max() + 1)
        np.add.at(busy, smid, dur)
        print(f"[{name}] per-SM sum of CTA time / span: min {busy.min()/span:.2f} mean {busy.mean()/span:.2f} max {busy.max()/span:.2f}; "
              f"last start at {100*(t0.max()-base)/span:.0f}% of span; first end at {100*(t1.min()-base)/span:.0f}%")
