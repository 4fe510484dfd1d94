"""Host-core PyTorch reference of the render() math (TEST INFRASTRUCTURE / reported CPU baseline; never imported by
the product): the rasterizer's forward written with vectorised torch ops so that torch AUTOGRAD derives the
Gaussian-parameter gradients independently of the analytic backward kernels (BASELINE.json's
"host-core PyTorch reference of the same render() math").

Restates, per stage (paths under /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer):
  projection + culling      forward.cu:157-258, auxiliary.h:58-77,139-164
  3D / 2D covariance        forward.cu:120-154, 76-115  (clamped t.x/t.z, t.y/t.z; +0.3 low-pass)
  conic, radius, tile rect  forward.cu:219-247, auxiliary.h:41-56
  tile lists                rasterizer_impl.cu:70-138 (ordered by (view depth, Gaussian id) per tile)
  alpha compositing         forward.cu:263-392 (alpha < 1/255 and power > 0 skipped, alpha <= 0.99,
                            stop before the Gaussian that would bring T below 1e-4)
Integer decisions (culling, radius, rects, order, skip / stop masks) are taken under no_grad and enter the
differentiable expression as constant masks, exactly like the reference's backward treats them.
Not restated: spherical harmonics, cov3D_precomp, and the pose gradient (its analytic form has deliberate quirks,
SURVEY 7 hard part 4: it is checked against oracle/cpu_ref.cpp, not against autograd).
"""
from __future__ import annotations

import math

import torch

TILE = 16


def project(means3D, scales, rotations, view, proj, W, H, tanfovx, tanfovy, scale_modifier=1.0):
    """-> dict(depth[P], xy[P,2], conic[P,3], radius[P] int, visible[P] bool, rect[P,4] int (x0,y0,x1,y1 tiles))."""
    P = means3D.shape[0]
    one = torch.ones(P, 1, dtype=means3D.dtype, device=means3D.device)
    hom = torch.cat([means3D, one], 1)
    pv = hom @ view                                  # view = W2C^T: p_view = [x y z 1] . V   (auxiliary.h:58-66)
    ph = hom @ proj
    depth = pv[:, 2]
    p_w = 1.0 / (ph[:, 3] + 0.0000001)
    ndc = ph[:, :2] * p_w[:, None]
    # computeCov3D
    s = scales * scale_modifier
    r, x, y, z = rotations.unbind(1)
    Rm = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                      2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                      2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(P, 3, 3)
    M = Rm * s[:, None, :]
    Sigma = M @ M.transpose(1, 2)
    # computeCov2D
    fx, fy = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    tz = pv[:, 2]
    cx = torch.clamp(pv[:, 0] / tz, -1.3 * tanfovx, 1.3 * tanfovx)
    cy = torch.clamp(pv[:, 1] / tz, -1.3 * tanfovy, 1.3 * tanfovy)
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tz, zero, -fx * cx / tz, zero, fy / tz, -fy * cy / tz], 1).view(P, 2, 3)
    Rw = view[:3, :3].transpose(0, 1)                # W2C rotation
    T = J @ Rw
    cov2 = T @ Sigma @ T.transpose(1, 2)
    a, b, c = cov2[:, 0, 0] + 0.3, cov2[:, 0, 1], cov2[:, 1, 1] + 0.3
    det = a * c - b * b
    with torch.no_grad():
        ok = (depth > 0.2) & (det != 0)
    det_s = torch.where(ok, det, torch.ones_like(det))
    conic = torch.stack([c / det_s, -b / det_s, a / det_s], 1)
    xy = ((ndc + 1.0) * torch.tensor([W, H], dtype=ndc.dtype, device=ndc.device) - 1.0) * 0.5   # ndc2Pix
    with torch.no_grad():
        mid = 0.5 * (a + c)
        sq = torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
        lam = torch.maximum(mid + sq, mid - sq)
        radius = torch.ceil(3.0 * torch.sqrt(lam)).to(torch.int64)
        gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
        rf = radius.to(xy.dtype)
        x0 = ((xy[:, 0] - rf) / TILE).to(torch.int64).clamp(0, gx)
        y0 = ((xy[:, 1] - rf) / TILE).to(torch.int64).clamp(0, gy)
        x1 = ((xy[:, 0] + rf + TILE - 1) / TILE).to(torch.int64).clamp(0, gx)
        y1 = ((xy[:, 1] + rf + TILE - 1) / TILE).to(torch.int64).clamp(0, gy)
        visible = ok & ((x1 - x0) * (y1 - y0) > 0)
        radius = torch.where(visible, radius, torch.zeros_like(radius))
    return dict(depth=depth, xy=xy, conic=conic, radius=radius, visible=visible,
                rect=torch.stack([x0, y0, x1, y1], 1))


def render(means3D, scales, rotations, opacities, colors, view, proj, W, H, tanfovx, tanfovy, bg,
           scale_modifier=1.0):
    """-> (color[3,H,W], depth[1,H,W], opacity[1,H,W], radii[P] int64, n_touched[P] int64); differentiable in the
    five Gaussian tensors."""
    g = project(means3D, scales, rotations, view, proj, W, H, tanfovx, tanfovy, scale_modifier)
    dev, dt = means3D.device, means3D.dtype
    color = torch.zeros(3, H, W, dtype=dt, device=dev)
    depth = torch.zeros(1, H, W, dtype=dt, device=dev)
    opacity = torch.zeros(1, H, W, dtype=dt, device=dev)
    n_touched = torch.zeros(means3D.shape[0], dtype=torch.int64, device=dev)
    with torch.no_grad():
        order = torch.argsort(g["depth"].detach(), stable=True)          # (depth, id): ids ascending within ties
        order = order[g["visible"][order]]
        rect = g["rect"][order]
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    ys, xs = torch.meshgrid(torch.arange(TILE, device=dev), torch.arange(TILE, device=dev), indexing="ij")
    out_c, out_d, out_o = [], [], []
    for ty in range(gy):
        row_c, row_d, row_o = [], [], []
        in_row = (rect[:, 1] <= ty) & (rect[:, 3] > ty)
        for tx in range(gx):
            ids = order[in_row & (rect[:, 0] <= tx) & (rect[:, 2] > tx)]
            px = (xs + tx * TILE).reshape(-1).to(dt)
            py = (ys + ty * TILE).reshape(-1).to(dt)
            inside = (px < W) & (py < H)                                   # pixels beyond a ragged edge never run
            if ids.numel() == 0:
                Tf = torch.ones(TILE * TILE, dtype=dt, device=dev)
                C = bg[:, None] * Tf[None]
                row_c.append(C); row_d.append(torch.zeros_like(Tf)); row_o.append(1 - Tf)
                continue
            xy, con, op = g["xy"][ids], g["conic"][ids], opacities[ids].reshape(-1)
            dx = xy[None, :, 0] - px[:, None]                              # [256, n]
            dy = xy[None, :, 1] - py[:, None]
            power = -0.5 * (con[None, :, 0] * dx * dx + con[None, :, 2] * dy * dy) - con[None, :, 1] * dx * dy
            alpha = torch.clamp(op[None] * torch.exp(power), max=0.99)
            with torch.no_grad():
                valid = (power <= 0) & (alpha >= 1.0 / 255.0)
            a_eff = torch.where(valid, alpha, torch.zeros_like(alpha))
            one_m = 1.0 - a_eff
            T_excl = torch.cumprod(torch.cat([torch.ones_like(one_m[:, :1]), one_m[:, :-1]], 1), 1)
            with torch.no_grad():
                test_T = T_excl * one_m
                stop = valid & (test_T < 0.0001)
                applied = valid & (torch.cumsum(stop.to(torch.int32), 1) == 0) & inside[:, None]
                n_touched.index_add_(0, ids, (applied & (test_T > 0.5)).sum(0))
            w = a_eff * T_excl * applied.to(dt)
            Tf = torch.prod(torch.where(applied, one_m, torch.ones_like(one_m)), 1)
            row_c.append(colors[ids].t() @ w.t() + bg[:, None] * Tf[None])
            row_d.append((w * g["depth"][ids][None]).sum(1))
            row_o.append(1.0 - Tf)
        out_c.append(torch.cat([t.view(3, TILE, TILE) for t in row_c], 2))
        out_d.append(torch.cat([t.view(TILE, TILE) for t in row_d], 1))
        out_o.append(torch.cat([t.view(TILE, TILE) for t in row_o], 1))
    color = torch.cat(out_c, 1)[:, :H, :W]
    depth = torch.cat(out_d, 0)[None, :H, :W]
    opacity = torch.cat(out_o, 0)[None, :H, :W]
    return color, depth, opacity, g["radius"], n_touched


def tracking_step(scene, view, proj, W, H, tanfovx, tanfovy, bg, dL_dcolor, dL_ddepth, threads=None):
    """One fwd + bwd (autograd) of config C1's shape on the host cores; returns (outputs, grads dict)."""
    if threads:
        torch.set_num_threads(int(threads))
    leaves = {k: scene[k].detach().clone().requires_grad_(True)
              for k in ("means3D", "scales", "rotations", "opacities", "colors")}
    color, depth, opacity, radii, n_touched = render(leaves["means3D"], leaves["scales"], leaves["rotations"],
                                                     leaves["opacities"], leaves["colors"], view, proj, W, H,
                                                     tanfovx, tanfovy, bg)
    ((color * dL_dcolor).sum() + (depth * dL_ddepth).sum()).backward()
    return (color.detach(), depth.detach(), opacity.detach(), radii, n_touched), {k: v.grad for k, v in leaves.items()}
