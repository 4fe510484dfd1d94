"""Synthetic scenes + cameras in the conventions MonoGS feeds the rasterizer.

Shared by the tests, the benchmark and the oracle checks so that every implementation sees
bit-identical inputs.  Conventions (reference file:line):
  * viewmatrix   = W2C^T                     (utils/camera_utils.py:172-174)
  * projmatrix   = (P * W2C)^T = W2C^T @ P^T (utils/camera_utils.py:224-231)
  * projmatrix_raw = P^T, OpenCV-style P     (gaussian_splatting/utils/graphics_utils.py:68-89, znear .01 zfar 100)
  * scales post-exp, rotations normalised w-first, opacities post-sigmoid [P,1]
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

# name -> (W, H, fx, fy, cx, cy)
INTRINSICS = {
    "tum": (640, 480, 535.4, 539.2, 320.1, 247.6),       # configs/mono/tum/fr3_office.yaml:6-16
    "replica": (1200, 680, 600.0, 600.0, 599.5, 339.5),  # configs/rgbd/replica/base_config.yaml:18-28
    "tum_half": (320, 240, 267.7, 269.6, 160.05, 123.8),
    "hd": (1920, 1080, 1606.2, 1617.6, 960.3, 557.1),
}


def projection_matrix(znear, zfar, cx, cy, fx, fy, W, H) -> torch.Tensor:
    """OpenCV-intrinsics projection used by MonoGS (row-major P, before the transpose)."""
    left = ((2 * cx - W) / W - 1.0) * W / 2.0
    right = ((2 * cx - W) / W + 1.0) * W / 2.0
    top = ((2 * cy - H) / H + 1.0) * H / 2.0
    bottom = ((2 * cy - H) / H - 1.0) * H / 2.0
    left, right = znear / fx * left, znear / fx * right
    top, bottom = znear / fy * top, znear / fy * bottom
    P = torch.zeros(4, 4, dtype=torch.float32)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def so3_exp(theta: torch.Tensor) -> torch.Tensor:
    th = theta.double()
    angle = th.norm()
    K = torch.tensor([[0, -th[2], th[1]], [th[2], 0, -th[0]], [-th[1], th[0], 0]], dtype=torch.float64)
    if angle < 1e-8:
        return (torch.eye(3, dtype=torch.float64) + K).float()
    return (torch.eye(3, dtype=torch.float64) + torch.sin(angle) / angle * K +
            (1 - torch.cos(angle)) / angle ** 2 * (K @ K)).float()


@dataclass
class Camera:
    W: int
    H: int
    fx: float
    fy: float
    cx: float
    cy: float
    w2c: torch.Tensor  # [4,4] row-major world-to-camera

    @property
    def tanfovx(self) -> float:
        return math.tan(math.atan(self.W / (2 * self.fx)))

    @property
    def tanfovy(self) -> float:
        return math.tan(math.atan(self.H / (2 * self.fy)))

    def matrices(self, device="cpu"):
        view = self.w2c.t().contiguous()
        P = projection_matrix(0.01, 100.0, self.cx, self.cy, self.fx, self.fy, self.W, self.H)
        proj_raw = P.t().contiguous()
        full = (view.unsqueeze(0).bmm(proj_raw.unsqueeze(0))).squeeze(0).contiguous()
        campos = torch.linalg.inv(view)[3, :3].contiguous()
        return tuple(t.to(device) for t in (view, full, proj_raw, campos))


def make_camera(kind="tum", rot=(0.0, 0.0, 0.0), trans=(0.0, 0.0, 0.0)) -> Camera:
    W, H, fx, fy, cx, cy = INTRINSICS[kind]
    w2c = torch.eye(4, dtype=torch.float32)
    w2c[:3, :3] = so3_exp(torch.tensor(rot, dtype=torch.float32))
    w2c[:3, 3] = torch.tensor(trans, dtype=torch.float32)
    return Camera(W, H, fx, fy, cx, cy, w2c)


def arc_cameras(kind: str, n: int, radius=0.3):
    """n cameras on a `radius`-metre arc looking roughly down +z (mapping-window keyframes, config C3/C4)."""
    cams = []
    for k in range(n):
        a = (k - (n - 1) / 2) / max(n - 1, 1)
        cams.append(make_camera(kind, rot=(0.0, -0.08 * a, 0.0), trans=(-radius * a, 0.01 * math.sin(3 * a), 0.0)))
    return cams


def make_scene(P: int, cam: Camera, seed=0, layered=False, anisotropic=False, frac_outlier=0.01,
               frac_behind=0.05, device="cpu"):
    """Gaussians mimicking create_viewpoint_pcd (gaussian_model.py:121-319): pixels sampled uniformly,
    depth U(0.5,6) (or 3 shells when `layered`), back-projected with the camera; screen-space sigma
    0.7-2 px (radius 2-6 px), `frac_outlier` large splats (sigma 17-67 px), `frac_behind` points behind the
    camera / outside the frustum; colour U(0,1); opacity sigmoid(N(0,1.5)); rotation = normalised N(0,1)^4."""
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(P, generator=g) * cam.W
    v = torch.rand(P, generator=g) * cam.H
    if layered:
        shell = torch.randint(0, 3, (P,), generator=g).float()
        z = 1.0 + 1.5 * shell + 0.05 * torch.randn(P, generator=g)
    else:
        z = 0.5 + 5.5 * torch.rand(P, generator=g)
    xc = (u - cam.cx) / cam.fx * z
    yc = (v - cam.cy) / cam.fy * z
    pts_c = torch.stack([xc, yc, z], 1)
    nb = int(P * frac_behind)
    if nb > 0:
        idx = torch.randperm(P, generator=g)[:nb]
        half = nb // 2
        pts_c[idx[:half], 2] = -pts_c[idx[:half], 2]                       # behind the camera
        pts_c[idx[half:], 0] = pts_c[idx[half:], 0] * 4.0 + 3.0 * z[idx[half:]]  # far outside the frustum
    c2w = torch.linalg.inv(cam.w2c)
    means = (pts_c @ c2w[:3, :3].t() + c2w[:3, 3]).contiguous()

    sigma_px = torch.exp(torch.log(torch.tensor(0.7)) + torch.rand(P, generator=g) * math.log(2.0 / 0.7))
    no = int(P * frac_outlier)
    if no > 0:
        idx = torch.randperm(P, generator=g)[:no]
        sigma_px[idx] = 17.0 + 50.0 * torch.rand(no, generator=g)
    sigma_w = sigma_px * z.abs() / cam.fx
    if anisotropic:
        scales = sigma_w[:, None] * torch.exp(torch.randn(P, 3, generator=g) * 0.7)
    else:
        scales = sigma_w[:, None].repeat(1, 3)
    rots = torch.randn(P, 4, generator=g)
    rots = rots / rots.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5)
    cols = torch.rand(P, 3, generator=g)
    out = dict(means3D=means.float(), scales=scales.float().contiguous(), rotations=rots.float().contiguous(),
               opacities=opac.float().contiguous(), colors=cols.float().contiguous())
    return {k: t.to(device) for k, t in out.items()}


def make_targets(cam: Camera, seed=0, device="cpu"):
    """Synthetic ground truth: smooth random image + depth with 10% invalid (zero) pixels."""
    g = torch.Generator().manual_seed(seed + 12345)
    low = torch.rand(1, 3, cam.H // 16 + 2, cam.W // 16 + 2, generator=g)
    img = torch.nn.functional.interpolate(low, size=(cam.H, cam.W), mode="bilinear", align_corners=False)[0]
    dlow = 0.5 + 5.5 * torch.rand(1, 1, cam.H // 16 + 2, cam.W // 16 + 2, generator=g)
    depth = torch.nn.functional.interpolate(dlow, size=(cam.H, cam.W), mode="bilinear", align_corners=False)[0]
    depth = depth * (torch.rand(1, cam.H, cam.W, generator=g) > 0.1).float()
    return img.contiguous().to(device), depth.contiguous().to(device)
