"""Golden vectors for the callers' losses (rows a17 / f1): outputs of the REFERENCE's own
get_loss_mapping / get_loss_tracking (/root/reference/utils/slam_utils.py:58-146), imported in this container and run
on CPU tensors, with the gradient images, exposure gradients and loss value taken by torch autograd THROUGH that
code.  The inputs are regenerated from the seed by `case_inputs` (torch's CPU generator is deterministic), so only the
outputs are stored.  Run from the repo root, where /root/reference exists:

    python tests/golden/make_loss_golden.py        -> tests/golden/loss.npz

Cases: mapping (mask + exposure), mapping_init (init=True: no exposure), tracking (mask * grad_mask * opacity > 0.99),
tracking_sparse (few opaque pixels), each with an upstream gradient != 1.
"""
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
REF = Path("/root/reference")
H, W = 64, 96
CASES = {  # name -> (kind, seed, init, upstream)
    "mapping": ("mapping", 1, False, 1.0),
    "mapping_init": ("mapping", 2, True, 3.0),
    "tracking": ("tracking", 3, False, 1.0),
    "tracking_sparse": ("tracking", 4, False, 0.5),
}


def case_inputs(name):
    """-> dict of CPU tensors: rendered image / depth / opacity and the viewpoint's fields."""
    kind, seed, init, upstream = CASES[name]
    g = torch.Generator().manual_seed(1000 + seed)
    image = torch.rand(3, H, W, generator=g)
    depth = 0.5 + 4.0 * torch.rand(1, H, W, generator=g)
    if name == "tracking_sparse":
        opacity = torch.rand(1, H, W, generator=g)                       # ~1 % of the pixels above 0.99
    else:
        opacity = 1.0 - 0.03 * torch.rand(1, H, W, generator=g) ** 2      # mostly above 0.99
    gt_rgb = torch.rand(3, H, W, generator=g)
    gt_depth = (0.5 + 4.0 * torch.rand(H, W, generator=g)) * (torch.rand(H, W, generator=g) > 0.15).float()
    mask = torch.rand(H, W, generator=g) > 0.2
    grad_mask = torch.rand(1, H, W, generator=g) > 0.5
    a = torch.tensor([0.05 * seed])
    b = torch.tensor([-0.01 * seed])
    return dict(kind=kind, init=init, upstream=upstream, image=image, depth=depth, opacity=opacity, gt_rgb=gt_rgb,
                gt_depth=gt_depth, mask=mask, grad_mask=grad_mask, exposure_a=a, exposure_b=b)


def main():
    sys.path.insert(0, str(REF))
    from utils import slam_utils  # the reference's own code (torch only)
    out = {}
    for name in CASES:
        c = case_inputs(name)
        image = c["image"].clone().requires_grad_(True)
        depth = c["depth"].clone().requires_grad_(True)
        opacity = c["opacity"].clone().requires_grad_(True)
        vp = types.SimpleNamespace(rgb=c["gt_rgb"], depth=c["gt_depth"], mask=c["mask"], grad_mask=c["grad_mask"],
                                   exposure_a=c["exposure_a"].clone().requires_grad_(True),
                                   exposure_b=c["exposure_b"].clone().requires_grad_(True))
        if c["kind"] == "mapping":
            loss = slam_utils.get_loss_mapping(image, depth, vp, init=c["init"], invert_depth=False)
        else:
            loss = slam_utils.get_loss_tracking(image, depth, opacity, vp, invert_depth=False)
        (c["upstream"] * loss).backward()
        z = lambda t, like: np.zeros_like(like.detach().numpy()) if t is None else t.numpy()
        out[f"{name}/loss"] = np.float64(loss.item())
        out[f"{name}/d_image"] = z(image.grad, image)
        out[f"{name}/d_depth"] = z(depth.grad, depth)
        out[f"{name}/d_opacity"] = z(opacity.grad, opacity)   # the rasterizer drops it (backward.cu), stored for reference
        out[f"{name}/d_a"] = z(vp.exposure_a.grad, vp.exposure_a)
        out[f"{name}/d_b"] = z(vp.exposure_b.grad, vp.exposure_b)
        print(name, float(loss), float(np.abs(out[f"{name}/d_image"]).max()), float(np.abs(out[f"{name}/d_depth"]).max()))
    np.savez_compressed(ROOT / "tests" / "golden" / "loss.npz", **out)


if __name__ == "__main__":
    main()
