"""CPU oracle (TEST INFRASTRUCTURE, never imported by the product) for the Gaussian-insertion path: a numpy
restatement of GaussianModel.create_viewpoint_pcd (/root/reference/gaussian_splatting/scene/gaussian_model.py:121-319)
with the random subsample passed in.  Pinned against tests/golden/insert.npz, which holds outputs of the reference's
own function (tests/golden/make_insert_golden.py); the kNN is oracle/cpu_ref.cpp's restatement of simple_knn.cu."""
import numpy as np

from . import cpu_oracle


def lower_median(v):
    """torch.median: the element of rank (n-1)//2 (gaussian_model.py:177,218)."""
    v = np.sort(np.asarray(v, np.float32).ravel())
    return v[(v.size - 1) // 2]


def select_mask(gt_depth, render_depth=None, render_opacity=None, init=False):
    """Flat indices (x*H + y, the reference's (W,H) flattening :180-186) of the pixels to densify (:191-226)."""
    d = np.asarray(gt_depth, np.float32).T.ravel()              # (W,H) order
    mask = d >= np.float32(1e-3)
    med_err = None
    if not init:
        rd = np.asarray(render_depth, np.float32).reshape(gt_depth.shape).T.ravel()
        low = (np.asarray(render_opacity, np.float32).reshape(gt_depth.shape).T.ravel() < 0.5
               if render_opacity is not None else np.ones_like(mask))
        err = np.abs(d - rd)
        med_err = lower_median(err)
        cond = (d < rd) & (err > np.float32(50.0) * med_err)
        mask = mask & (low | cond)
    return np.nonzero(mask)[0].astype(np.int32), med_err


def create_viewpoint_pcd(rgb, depth, segmentation, R, T, fx, fy, cx, cy, random_indices, render_depth=None,
                         render_opacity=None, exposure_a=None, exposure_b=None, init=False, isotropic=True):
    H, W = depth.shape
    rgb = np.asarray(rgb, np.float32)
    if not init:                                                                        # :133-139
        rgb = np.clip(np.exp(np.float32(exposure_a)) * rgb + np.float32(exposure_b), 0.0, 1.0).astype(np.float32)
    sel, _ = select_mask(depth, render_depth, render_opacity, init)
    idx = sel[np.asarray(random_indices, np.int64)]
    x, y = idx // H, idx % H
    Kmat = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    K_inv = np.linalg.inv(Kmat).astype(np.float32)                                      # :163
    w2c = np.eye(4, dtype=np.float32)
    w2c[:3, :3], w2c[:3, 3] = R, T
    c2w = np.linalg.inv(w2c).astype(np.float32)                                         # :164-165
    d = np.asarray(depth, np.float32)[y, x]
    pix = np.stack([x + 0.5, y + 0.5, np.ones_like(x, np.float32)], 1).astype(np.float32)  # :238-262
    cam = (pix @ K_inv.T) * d[:, None]                                                  # :266-272
    world = (cam @ c2w[:3, :3].T + c2w[:3, 3]).astype(np.float32)                       # :275-277
    feats = rgb[:, y, x].T.copy()
    ids = np.asarray(segmentation)[y, x]
    point_size = min(np.float32(0.05), np.float32(0.01) * lower_median(depth))          # :169-177
    dist2 = np.maximum(cpu_oracle.knn(world), np.float32(1e-7)) * point_size            # :294-303
    scales = np.log(np.sqrt(dist2))[:, None].astype(np.float32)
    if not isotropic:
        scales = np.repeat(scales, 3, 1)                                                # :305-306
    rots = np.zeros((len(idx), 4), np.float32)
    rots[:, 0] = 1                                                                      # :308-311
    opac = np.zeros((len(idx), 1), np.float32)                                          # logit(0.5), :313-318
    return world, feats, scales, rots, opac, ids
