"""Generate tests/golden/loss_*.npz by running the REAL reference ``utils/loss_utils.py`` (l1_loss, l2_loss, ssim,
s3im_fun: :16-87) on the CPU through the expressions of ``train_ZJU.py:107-131`` (masked gathers, cv2.boundingRect crop,
loss weights), with autograd for dL/dimage and dL/dalpha.  LPIPS (a VGG checkpoint) and the pose NLL are left out: they
are not part of the fused op.

Run in the build container only (needs /root/reference and cv2; the GPU box never reads either):

    python tests/golden/make_loss_golden.py
"""
import importlib.util
import os
import sys

import cv2
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def synth_case(W, H, seed, box, empty_bound=False):
    """A rendered-looking image pair: smooth blobs + noise; bound_mask = an irregular region inside `box`; bkgd_mask = a
    smaller silhouette; alpha = a blurred version of it."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    def blobs(n):
        img = np.zeros((3, H, W), np.float32)
        for _ in range(n):
            cx, cy, s = rng.uniform(0, W), rng.uniform(0, H), rng.uniform(3, 0.25 * max(W, H))
            col = rng.uniform(0, 1, 3).astype(np.float32)
            img += col[:, None, None] * np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * s * s))[None]
        return img
    gt = np.clip(blobs(12) * 0.6 + 0.05 * rng.standard_normal((3, H, W)).astype(np.float32), 0, 1)
    image = np.clip(gt + 0.15 * blobs(6) - 0.1 + 0.08 * rng.standard_normal((3, H, W)).astype(np.float32), 0, 1)
    x0, y0, x1, y1 = box
    bound = np.zeros((1, H, W), np.float32)
    if not empty_bound:
        cx, cy = 0.5 * (x0 + x1), 0.5 * (y0 + y1)
        ell = ((xx - cx) / (0.5 * (x1 - x0) + 0.5)) ** 2 + ((yy - cy) / (0.5 * (y1 - y0) + 0.5)) ** 2 <= 1.0
        bound[0][ell] = 1.0
        bound[0, y0, x0] = 1.0; bound[0, y1, x1] = 1.0            # the box corners: the rectangle is exactly `box`
    sil = ((xx - 0.5 * (x0 + x1)) ** 2 / (0.3 * (x1 - x0) + 1) ** 2 + (yy - 0.5 * (y0 + y1)) ** 2 / (0.4 * (y1 - y0) + 1) ** 2 <= 1.0)
    bkgd = sil.astype(np.float32)[None]
    alpha = np.clip(bkgd * 0.9 + 0.1 * rng.uniform(0, 1, (1, H, W)).astype(np.float32), 0, 1)
    image[:, :2, :] = gt[:, :2, :]                                 # some exactly equal pixels: sign(0) = 0 in the L1 gradient
    return image, alpha, gt, bkgd, bound


def main():
    spec = importlib.util.spec_from_file_location("ref_loss_utils", os.path.join(REF, "utils", "loss_utils.py"))
    LU = importlib.util.module_from_spec(spec)          # the real reference module, unmodified
    spec.loader.exec_module(LU)
    cases = {
        # name: (W, H, seed, box (x0,y0,x1,y1) inclusive)
        "mid": (160, 128, 1, (37, 11, 121, 117)),       # a person-sized box, not tile aligned, several tiles each way
        "ragged": (200, 120, 2, (3, 0, 190, 119)),      # box touches three image borders, ragged tiles
        "small": (70, 45, 3, (31, 9, 40, 20)),          # crop narrower than the window
        "full": (96, 64, 4, (0, 0, 95, 63)),            # bound covers everything
    }
    for name, (W, H, seed, box) in cases.items():
        image, alpha, gt, bkgd, bound = (torch.from_numpy(a) for a in synth_case(W, H, seed, box))
        image.requires_grad_(True); alpha.requires_grad_(True)
        gt_image, bkgd_mask, bound_mask = gt, bkgd, bound
        # ---- train_ZJU.py:111-123 on the real module -----------------------------------------------------------------------
        Ll1 = LU.l1_loss(image.permute(1, 2, 0)[bound_mask[0] == 1], gt_image.permute(1, 2, 0)[bound_mask[0] == 1])
        mask_loss = LU.l2_loss(alpha[bound_mask == 1], bkgd_mask[bound_mask == 1])
        x, y, w, h = cv2.boundingRect(bound_mask[0].cpu().numpy().astype(np.uint8))
        img_pred = image[:, y:y + h, x:x + w].unsqueeze(0)
        img_gt = gt_image[:, y:y + h, x:x + w].unsqueeze(0)
        ssim_loss = LU.ssim(img_pred, img_gt)
        s3im_loss = LU.s3im_fun(img_pred, img_gt)
        loss = Ll1 + 0.5 * mask_loss + 0.2 * (1.0 - ssim_loss) + 0.3 * s3im_loss        # :131 without LPIPS / NLL
        # per-term gradients as well: a term with a small weight must not hide behind the others
        terms = dict(l1=Ll1, mask=mask_loss, ssim=ssim_loss, s3im=s3im_loss, total=loss)
        out = dict(W=W, H=H, rect=np.array([x, y, w, h], np.int32), count=int((bound_mask == 1).sum()),
                   image=image.detach().numpy(), alpha=alpha.detach().numpy(), gt=gt.numpy(), bkgd=bkgd.numpy(), bound=bound.numpy())
        for k, t in terms.items():
            gi, ga = torch.autograd.grad(t, [image, alpha], retain_graph=True, allow_unused=True)
            out[f"{k}"] = np.float32(t.item())
            out[f"d_{k}_dimage"] = (gi if gi is not None else torch.zeros_like(image)).numpy()
            if k in ("mask", "total"):
                out[f"d_{k}_dalpha"] = ga.numpy()
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), f"loss_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "->", path, {k: (v.shape if hasattr(v, "shape") and v.shape else v) for k, v in out.items() if not k.startswith("d_")})


if __name__ == "__main__":
    main()
