"""ctypes binding of the C ABI in include/gsr_b200.h.

The product path has NO fallback: if the CUDA library is missing or a symbol
cannot be resolved this module raises at import/first use.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import os

# GSR_B200_LIB selects an experimental build variant (monogs_b200/build.py --variant); default = the product library
_LIB_PATH = Path(os.environ.get("GSR_B200_LIB") or (Path(__file__).resolve().parent / "libgsr_b200.so"))

vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t

# name -> (restype, argtypes); mirrors include/gsr_b200.h one to one.
SIGNATURES = {
    "gsr_version": (C.c_char_p, []),
    "gsr_geom_bytes": (sz, [i32, i32, i32]),
    "gsr_image_bytes": (sz, [i32, i32]),
    "gsr_point_list_bytes": (sz, [i64]),
    "gsr_binning_bytes": (sz, [i64]),
    "gsr_backward_scratch_bytes": (sz, [i32]),
    "gsr_knn_bytes": (sz, [i32]),
    "gsr_forward_plan": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, f32, f32,
                               i32, vp, vp, C.POINTER(i64), i32]),
    "gsr_forward_render": (i32, [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, i32]),
    "gsr_forward": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, f32, f32, i32,
                          vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, C.POINTER(i64), i32]),
    "gsr_forward_enqueue": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, f32,
                                  f32, i32, vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, vp]),
    "gsr_backward": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, vp, f32, f32,
                           vp, vp, vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32]),
    "gsr_l1_loss_forward": (i32, [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, f32, vp, vp]),
    "gsr_l1_loss_grad": (i32, [vp, i32, i32, vp, vp, vp]),
    "gsr_backward_l1": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, vp, f32, f32,
                              vp, vp, vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32]),
    "gsr_backward_render": (i32, [vp, i32, i32, i32, vp, vp, vp, vp, i64, vp, vp, vp, i32]),
    "gsr_backward_render_l1": (i32, [vp, i32, i32, i32, vp, vp, vp, vp, i64, vp, vp, i32]),
    "gsr_backward_rows": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, f32, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, i32,
                                vp, vp, i32]),
    "gsr_backward_rows_marked": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, f32, vp, vp, vp, vp, f32, f32, vp, vp, vp,
                                       vp, i32, vp, vp, vp, vp, i32]),
    "gsr_rows_exchange_bytes": (C.c_size_t, [i32, i32]),
    "gsr_rows_exchange_layout": (i32, [i32, i32, C.POINTER(C.c_size_t), C.POINTER(i32), C.POINTER(i32)]),
    "gsr_rows_exchange": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, vp, C.c_uint32, vp, vp]),
    "gsr_debug_exchange_trace": (i32, [vp]),
    "gsr_insert_ws_bytes": (sz, [i32, i32]),
    "gsr_select_kth_f32": (i32, [vp, i64, vp, vp, i64, vp, vp]),
    "gsr_insert_select": (i32, [vp, i32, i32, vp, vp, vp, i32, vp, vp, vp, vp]),
    "gsr_insert_select_static": (i32, [vp, i32, i32, vp, vp, vp, i32, vp, vp, vp]),
    "gsr_insert_points": (i32, [vp, i32, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, f32, f32, f32, f32, vp, vp,
                                vp]),
    "gsr_insert_finish": (i32, [vp, i32, vp, vp, i32, vp, vp, vp]),
    "gsr_surgery_ws_bytes": (sz, [i64]),
    "gsr_rows_append": (i32, [vp, i32, vp, vp, vp, vp, vp, i64, i64]),
    "gsr_rows_compact": (i32, [vp, i32, vp, vp, vp, i64, vp, vp, vp]),
    "gsr_mark_visible": (i32, [vp, i32, vp, vp, vp, vp]),
    "gsr_knn_dist2": (i32, [vp, i32, vp, vp, vp]),
    "gsr_profile_enable": (i32, [i32]),
    "gsr_profile_collect": (i32, [C.POINTER(C.c_double), C.POINTER(i32)]),
    "gsr_profile_kernel_launches": (C.c_ulonglong, []),
    "gsr_sort_pairs_u64": (i32, [vp, i64, vp, vp, vp, vp, i32, vp]),
    "gsr_debug_geom_views": (i32, [vp, i32, i32, i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
    "gsr_debug_image_views": (i32, [vp, i32, i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
    "gsr_debug_sorted_keys": (i32, [vp, vp, vp, vp, i32, i64, i32, i32, vp]),
    "gsr_debug_set_binning": (i32, [i32]),
    "gsr_binning_needed": (i32, [i32, i32]),
    "gsr_densify_stats": (i32, [vp, i32, vp, vp, vp, vp, vp]),
    "gsr_visibility_prune": (i32, [vp, i32, i32, vp, vp, i32, i32, vp, vp]),
    "gsr_camera_setup": (i32, [vp, vp, i32, i32, vp, i32, vp, vp, vp, vp]),
}

class L1LossSpec(C.Structure):
    """gsr_l1_loss of include/gsr_b200.h."""
    _fields_ = [("kind", i32)] + [(n, vp) for n in ("color", "depth", "opacity", "gt_rgb", "gt_depth", "mask",
                                                     "grad_mask", "exposure_a", "exposure_b", "out", "upstream")]


LOSS_MAPPING, LOSS_TRACKING = 1, 2

STAGES = ("preprocess", "depth_sort", "scan", "emit", "tile_sort", "ranges", "render_fwd", "render_bwd",
          "preprocess_bwd")

GSR_ECAPACITY = -6
GSR_ESTATICLIMIT = -7
INSERT_STATIC_MAX_PIXELS = 4 * 1024 * 1024

ERRORS = {
    -1: "GSR_EINVAL: bad argument combination or size",
    -2: "GSR_ENOCOLOR: neither shs nor colors_precomp given",
    -3: "GSR_ENOCOV: neither scales+rotations nor cov3D_precomp given",
    -4: "GSR_EPREFILTER: point is filtered although prefiltered is set",
    -5: "GSR_EOVERFLOW: num_rendered exceeds 2^30 instances",
    -6: "GSR_ECAPACITY: num_rendered exceeds the capacity passed to gsr_forward",
    -7: "GSR_ESTATICLIMIT: the size exceeds the fixed scratch array of a *_static entry point "
        "(gsr_insert_select_static: width * height <= 4 Mi pixels); use the workspace variant",
}

_lib = None


def lib_path() -> Path:
    return _LIB_PATH


def load() -> C.CDLL:
    """Load libgsr_b200.so and attach the signatures. Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise RuntimeError(
            f"{_LIB_PATH} is not built. Run `python -m monogs_b200.build` (needs nvcc); "
            "there is no CPU or PyTorch fallback for the rasterizer.")
    lib = C.CDLL(str(_LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc == 0:
        return
    if rc < 0:
        raise RuntimeError(f"{what}: {ERRORS.get(rc, f'error {rc}')}")
    raise RuntimeError(f"{what}: CUDA error {rc}")
