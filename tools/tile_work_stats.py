"""tools/tile_work_stats.py -- where the render kernels' work sits: per-tile replay length (largest n_contrib) vs
list length and tile position, for a bench workload."""
import ctypes as C, json, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench
wl_name = sys.argv[1] if len(sys.argv) > 1 else "headline"
torch.cuda.set_device(0)
wl = bench.Workload(wl_name, 0, 1, torch.device("cuda", 0))
s = bench.OursDeviceStep(wl, 1)
s.step(); s.step(); torch.cuda.synchronize()
nc_p, ft_p, rg_p = C.c_void_p(), C.c_void_p(), C.c_void_p()
s.lib.gsr_debug_image_views(s.img.data_ptr(), wl.W, wl.H, C.byref(ft_p), C.byref(nc_p), C.byref(rg_p))
tx, ty = (wl.W + 15) // 16, (wl.H + 15) // 16
off = nc_p.value - s.img.data_ptr()
nc = s.img[off:off + 4 * wl.N].view(torch.int32).view(wl.H, wl.W).float()
work = torch.nn.functional.max_pool2d(nc[None, None], 16, ceil_mode=True)[0, 0]
off = rg_p.value - s.img.data_ptr()
rg = s.img[off:off + 8 * tx * ty].view(torch.int32).view(ty, tx, 2)
length = (rg[..., 1] - rg[..., 0]).float()
border = torch.zeros(ty, tx, dtype=torch.bool, device=work.device)
border[0] = border[-1] = True; border[:, 0] = border[:, -1] = True
def st(x): return {"mean": round(float(x.mean()), 1), "p50": float(x.median()), "p90": float(x.quantile(0.9)), "max": float(x.max())}
w = work.flatten(); order = torch.argsort(w, descending=True)
top = [(int(i) // tx, int(i) % tx, int(w[i]), int(length.flatten()[i])) for i in order[:24]]
corr = float(torch.corrcoef(torch.stack([w, length.flatten()]))[0, 1])
print(json.dumps({"workload": wl_name, "tiles": [tx, ty], "work_all": st(w), "work_border": st(work[border]),
                  "work_interior": st(work[~border]), "length_all": st(length.flatten()), "corr_work_length": corr,
                  "frac_work_in_border": float(work[border].sum() / w.sum()),
                  "unsaturated_tiles(work==length)": int((work >= length).sum()), "top24(y,x,work,len)": top}))
