"""Host-side mirror of the image-space losses of MOSS's training step, over the fused sm_100a kernel.

What ``train_ZJU.py:107-131`` does with ``utils/loss_utils.py`` (l1_loss, l2_loss, ssim, s3im_fun) --

    Ll1 = l1_loss(image.permute(1,2,0)[bound_mask[0]==1], gt_image.permute(1,2,0)[bound_mask[0]==1])
    mask_loss = l2_loss(alpha[bound_mask==1], bkgd_mask[bound_mask==1])
    x, y, w, h = cv2.boundingRect(bound_mask[0].cpu().numpy().astype(np.uint8))
    ssim_loss = ssim(image[:, y:y+h, x:x+w][None], gt_image[:, y:y+h, x:x+w][None])
    s3im_loss = s3im_fun(img_pred, img_gt)
    loss = Ll1 + 0.5 * mask_loss + 0.2 * (1.0 - ssim_loss) + 0.3 * s3im_loss  (+ LPIPS + pose NLL)

-- becomes

    bounds = mask_bounds(bound_mask)                       # once per camera: rectangle + pixel count, stays on the device
    loss, terms = image_losses(image, alpha, gt_image, bkgd_mask, bound_mask, bounds)
    (loss + 0.5 * lpips_loss + 0.06 * nll_loss).backward()

ONE launch forward (``moss_image_losses``: the four loss values AND dL/dimage, dL/dalpha), one tiny launch backward
(the upstream scale, read on the device).  No boolean-mask gather, no host sync, no CPU path.
"""
import ctypes as C

import torch

from . import _abi

WEIGHTS = (1.0, 0.5, 0.2, 0.3)          # Ll1, mask_loss, (1 - ssim), s3im (train_ZJU.py:131)
_scratch = {}                           # (device index, stream handle) -> zero-initialised scratch (self-cleaning)


def _f32(t, name, shape=None):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (moss_b200 has no CPU path)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32")
    if shape is not None and tuple(t.shape[-len(shape):]) != tuple(shape):
        raise RuntimeError(f"{name} must have trailing dimensions {tuple(shape)}")
    return t.contiguous()


def _p(t):
    return None if t is None else t.data_ptr()


def _scratch_for(dev, stream, nbytes):
    key = (dev.index, stream)
    t = _scratch.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.zeros((nbytes,), dtype=torch.uint8, device=dev)
        _scratch[key] = t
    return t


def mask_bounds(mask):
    """int32 [8] on the device: (x, y, w, h) of the bounding rectangle of the non-zero pixels of ``mask`` ([1,H,W] or
    [H,W]) -- what ``cv2.boundingRect(bound_mask[0].cpu().numpy().astype(np.uint8))`` returns (train_ZJU.py:115),
    without the device-to-host copy -- then the number of pixels equal to 1 (the size of the gathers at :111-112).
    Constant per camera: compute once, keep."""
    lib = _abi.load()
    if mask.dtype != torch.float32:
        mask = mask.float()
    m = _f32(mask, "mask")
    H, W = m.shape[-2:]
    if m.numel() != H * W:
        raise RuntimeError("mask must be [H,W] or [1,H,W]")
    with _abi.on_device(m.device):
        out = torch.empty((8,), dtype=torch.int32, device=m.device)
        st = lib.moss_mask_bounds(_abi.stream_ptr(m.device), W, H, _p(m), _p(out))
    _abi.check(st, "moss_mask_bounds")
    return out


class _ImageLosses(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, alpha, gt_image, bkgd_mask, bound_mask, bounds, weights):
        lib = _abi.load()
        image = _f32(image, "image")
        if image.ndim != 3 or image.shape[0] != 3:
            raise RuntimeError("image must have dimensions (3, H, W)")
        H, W = image.shape[1:]
        dev = image.device
        gt_image = _f32(gt_image, "gt_image", (3, H, W))
        alpha, bkgd_mask = _f32(alpha, "alpha", (H, W)), _f32(bkgd_mask, "bkgd_mask", (H, W))
        bound_mask = _f32(bound_mask, "bound_mask", (H, W))
        if bounds.dtype != torch.int32 or not bounds.is_cuda or bounds.numel() < 5:
            raise RuntimeError("bounds must be the int32 CUDA tensor mask_bounds() returns")
        w = [float(x) for x in weights]
        if w[1] != 0.0 and (alpha is None or bkgd_mask is None):
            raise RuntimeError("the mask loss needs alpha and bkgd_mask")
        with _abi.on_device(dev):
            stream = _abi.stream_ptr(dev)
            scratch = _scratch_for(dev, stream, lib.moss_image_losses_scratch_bytes(W, H))
            out = torch.empty((8 + 4 * H * W,), dtype=torch.float32, device=dev)
            losses, d_image, d_alpha = out[:5], out[8:8 + 3 * H * W].view(3, H, W), out[8 + 3 * H * W:].view(1, H, W)
            k = _abi.MossImageLoss()
            k.W, k.H = W, H
            k.image, k.alpha, k.gt_image, k.bkgd_mask, k.bound_mask = _p(image), _p(alpha), _p(gt_image), _p(bkgd_mask), _p(bound_mask)
            k.rect = _p(bounds)
            k.w_l1, k.w_mask, k.w_ssim, k.w_s3im = w
            k.scratch, k.scratch_bytes = _p(scratch), scratch.numel()
            st = lib.moss_image_losses(stream, C.byref(k), _p(losses), _p(d_image), _p(d_alpha) if alpha is not None else None)
        _abi.check(st, "moss_image_losses")
        ctx.has_alpha = alpha is not None
        ctx.alpha_shape = None if alpha is None else tuple(alpha.shape)
        ctx.save_for_backward(d_image, d_alpha)
        terms = losses[:4]
        ctx.mark_non_differentiable(terms)
        return losses[4], terms

    @staticmethod
    def backward(ctx, g_total, _g_terms):
        lib = _abi.load()
        d_image, d_alpha = ctx.saved_tensors
        dev = d_image.device
        g = _f32(g_total.reshape(1), "grad")
        with _abi.on_device(dev):
            # in place on buffers this node owns; a no-op on the device when the upstream gradient is exactly 1
            st = lib.moss_scale_gradients(_abi.stream_ptr(dev), _p(g), _p(d_image), d_image.numel(),
                                          _p(d_alpha) if ctx.has_alpha else None, d_alpha.numel() if ctx.has_alpha else 0)
        _abi.check(st, "moss_scale_gradients")
        return d_image, (d_alpha.view(ctx.alpha_shape) if ctx.has_alpha else None), None, None, None, None, None


def image_losses(image, alpha, gt_image, bkgd_mask, bound_mask, bounds=None, weights=WEIGHTS):
    """-> (loss, terms): ``loss`` = w0 Ll1 + w1 mask_loss + w2 (1 - ssim) + w3 s3im as train_ZJU.py:131 weighs them
    (differentiable w.r.t. ``image`` [3,H,W] and ``alpha`` [1,H,W]); ``terms`` float32 [4] on the device = (Ll1, mask_loss,
    ssim, s3im) for logging.  ``bounds`` = mask_bounds(bound_mask), computed here when absent (cache it per camera).
    A zero weight skips that term's work.  Backward once (the gradient buffers are scaled in place)."""
    if bounds is None:
        bounds = mask_bounds(bound_mask)
    return _ImageLosses.apply(image, alpha, gt_image, bkgd_mask, bound_mask, bounds, tuple(weights))


def ssim(img1, img2, window_size=11, size_average=True):
    """utils/loss_utils.py:57 for one image pair ([3,H,W] or [1,3,H,W]): mean SSIM with the 11x11 sigma-1.5 window."""
    if window_size != 11 or not size_average:
        raise RuntimeError("moss_b200.losses.ssim implements window_size=11, size_average=True (what MOSS calls)")
    a = img1.reshape(img1.shape[-3:]) if img1.ndim == 4 and img1.shape[0] == 1 else img1
    b = img2.reshape(img2.shape[-3:]) if img2.ndim == 4 and img2.shape[0] == 1 else img2
    H, W = a.shape[-2:]
    bounds = torch.tensor([0, 0, W, H, W * H, 0, 0, 0], dtype=torch.int32, device=a.device)
    total, _ = _ImageLosses.apply(a, None, b.detach(), None, None, bounds, (0.0, 0.0, -1.0, 0.0))
    return total + 1.0          # w_ssim = -1: total = -(1 - ssim)
