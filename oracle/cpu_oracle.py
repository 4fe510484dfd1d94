"""oracle/cpu_oracle.py -- TEST INFRASTRUCTURE: numpy front end of oracle/liboracle_cpu.so (cpu_ref.cpp).

`forward()` / `backward()` restate the whole reference path on the host cores; `knn()` restates distCUDA2.
Never imported by the product package (monogs_b200/): only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg use it.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = _HERE / "liboracle_cpu.so"
_lib = None


def build():
    subprocess.run(["make", "-C", str(_HERE), "cpu"], check=True, capture_output=True)


def load():
    global _lib
    if _lib is None:
        if not _LIB.exists():
            build()
        _lib = C.CDLL(str(_LIB))
        _lib.oracle_preprocess.restype = C.c_int64
    return _lib


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def forward(means3D, scales, rotations, opacities, colors, view, proj, campos, W, H, tanx, tany, bg,
            scale_modifier=1.0, shs=None, sh_degree=0, cov3D_precomp=None):
    """Full reference forward on the CPU. Returns a dict with every intermediate the reference keeps."""
    lib = load()
    means3D, scales, rotations, opacities = _f(means3D), _f(scales), _f(rotations), _f(opacities)
    colors, view, proj, campos, bg, shs, cov3D_precomp = map(_f, (colors, view, proj, campos, bg, shs, cov3D_precomp))
    P = means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    o = dict(radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), np.float32), depths=np.zeros(P, np.float32),
             cov3D=np.zeros((P, 6), np.float32), conic_opacity=np.zeros((P, 4), np.float32),
             rgb=np.zeros((P, 3), np.float32), clamped=np.zeros((P, 3), np.uint8),
             tiles_touched=np.zeros(P, np.uint32))
    R = lib.oracle_preprocess(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _ptr(means3D), _ptr(scales),
                              C.c_float(scale_modifier), _ptr(rotations), _ptr(opacities), _ptr(shs),
                              _ptr(cov3D_precomp), _ptr(colors), _ptr(view), _ptr(proj), _ptr(campos), C.c_int(W),
                              C.c_int(H), C.c_float(tanx), C.c_float(tany), _ptr(o["radii"]), _ptr(o["means2D"]),
                              _ptr(o["depths"]), _ptr(o["cov3D"]), _ptr(o["conic_opacity"]), _ptr(o["rgb"]),
                              _ptr(o["clamped"]), _ptr(o["tiles_touched"]))
    R = int(R)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    o.update(num_rendered=R, keys=np.zeros(max(R, 1), np.uint64)[:R], point_list=np.zeros(max(R, 1), np.uint32)[:R],
             ranges=np.zeros((tiles, 2), np.uint32))
    keys_buf = np.zeros(max(R, 1), np.uint64)
    pl_buf = np.zeros(max(R, 1), np.uint32)
    lib.oracle_bin(C.c_int(P), C.c_int(W), C.c_int(H), _ptr(o["means2D"]), _ptr(o["depths"]), _ptr(o["radii"]),
                   _ptr(o["tiles_touched"]), C.c_int64(R), _ptr(keys_buf), _ptr(pl_buf), _ptr(o["ranges"]))
    o["keys"], o["point_list"] = keys_buf[:R], pl_buf[:R]
    feat = colors if colors is not None else o["rgb"]
    o.update(final_T=np.zeros(H * W, np.float32), n_contrib=np.zeros(H * W, np.uint32),
             color=np.zeros((3, H, W), np.float32), depth=np.zeros((1, H, W), np.float32),
             opacity=np.zeros((1, H, W), np.float32), n_touched=np.zeros(P, np.int32))
    lib.oracle_render(C.c_int(W), C.c_int(H), _ptr(o["ranges"]), _ptr(pl_buf), _ptr(o["means2D"]), _ptr(feat),
                      _ptr(o["conic_opacity"]), _ptr(o["depths"]), _ptr(bg), _ptr(o["final_T"]), _ptr(o["n_contrib"]),
                      _ptr(o["color"]), _ptr(o["depth"]), _ptr(o["opacity"]), _ptr(o["n_touched"]))
    o["_feat"] = feat
    o["_pl_buf"] = pl_buf
    return o


def backward(fwd, means3D, scales, rotations, view, proj, proj_raw, W, H, tanx, tany, bg, dL_dcolor, dL_ddepth,
             scale_modifier=1.0, cov3D_precomp=None):
    """Reference backward on the CPU from a `forward()` result. Returns dict of gradients (+ tau_sum[6])."""
    lib = load()
    means3D, scales, rotations, view, proj, proj_raw, bg = map(_f, (means3D, scales, rotations, view, proj, proj_raw, bg))
    dL_dcolor, dL_ddepth, cov3D_precomp = _f(dL_dcolor), _f(dL_ddepth), _f(cov3D_precomp)
    P = means3D.shape[0]
    g = dict(means2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 4), np.float32),
             opacity=np.zeros((P, 1), np.float32), colors=np.zeros((P, 3), np.float32),
             depths=np.zeros(P, np.float32), means3D=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32),
             scales=np.zeros((P, 3), np.float32), rotations=np.zeros((P, 4), np.float32),
             tau=np.zeros((P, 6), np.float32), tau_sum=np.zeros(6, np.float64))
    lib.oracle_render_backward(C.c_int(P), C.c_int(W), C.c_int(H), _ptr(fwd["ranges"]), _ptr(fwd["_pl_buf"]), _ptr(bg),
                               _ptr(fwd["means2D"]), _ptr(fwd["conic_opacity"]), _ptr(fwd["_feat"]),
                               _ptr(fwd["depths"]), _ptr(fwd["final_T"]), _ptr(fwd["n_contrib"]), _ptr(dL_dcolor),
                               _ptr(dL_ddepth), _ptr(g["means2D"]), _ptr(g["conic"]), _ptr(g["opacity"]),
                               _ptr(g["colors"]), _ptr(g["depths"]))
    cov = cov3D_precomp if cov3D_precomp is not None else fwd["cov3D"]
    lib.oracle_preprocess_backward(C.c_int(P), _ptr(means3D), _ptr(fwd["radii"]), _ptr(scales), _ptr(rotations),
                                   C.c_float(scale_modifier), _ptr(cov), C.c_int(0 if cov3D_precomp is not None else 1),
                                   _ptr(view), _ptr(proj), _ptr(proj_raw), C.c_int(W), C.c_int(H), C.c_float(tanx),
                                   C.c_float(tany), _ptr(g["means2D"]), _ptr(g["conic"]), _ptr(g["depths"]),
                                   _ptr(g["means3D"]), _ptr(g["cov3D"]), _ptr(g["scales"]), _ptr(g["rotations"]),
                                   _ptr(g["tau"]), _ptr(g["tau_sum"]))
    return g


def knn(points):
    lib = load()
    pts = _f(points)
    out = np.zeros(pts.shape[0], np.float32)
    lib.oracle_knn(C.c_int(pts.shape[0]), _ptr(pts), _ptr(out))
    return out


def mark_visible(means3D, view):
    lib = load()
    m, v = _f(means3D), _f(view)
    out = np.zeros(m.shape[0], np.uint8)
    lib.oracle_mark_visible(C.c_int(m.shape[0]), _ptr(m), _ptr(v), _ptr(out))
    return out.astype(bool)
