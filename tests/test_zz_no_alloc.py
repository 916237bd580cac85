"""moss_image_losses_static(): the image-loss kernel with its scratch memory in a static device array of the library
(include/moss_splat.h; no counterpart in the reference) against
moss_image_losses() with a caller-owned scratch buffer, both through the C ABI.  The kernel adds its per-tile partial sums
in a fixed order, so the two entry points must agree BIT FOR BIT (tolerance 0, tighter than GRAD_TOL of
tests/test_loss_gpu.py), on two calls in a row -- the second one finds the first one's contents in the shared array.

Collected last on purpose (pytest -x would hide every test behind a failure here)."""
import ctypes as C

import pytest
import torch

from moss_b200 import _abi

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _inputs(W, H, seed):
    from tests.golden.make_loss_golden import synth_case
    box = (W // 5, H // 7, W // 2 + 3, H // 2 + 5)
    image, alpha, gt, bkgd, bound = (torch.from_numpy(a).to(DEV) for a in synth_case(W, H, seed, box))
    lib = _abi.load()
    rect = torch.zeros(8, dtype=torch.int32, device=DEV)
    _abi.check(lib.moss_mask_bounds(torch.cuda.current_stream().cuda_stream, W, H, _p(bound.reshape(H, W)), _p(rect)), "moss_mask_bounds")
    return image.contiguous(), alpha.reshape(H, W).contiguous(), gt.contiguous(), bkgd.reshape(H, W).contiguous(), bound.reshape(H, W).contiguous(), rect


def _struct(W, H, image, alpha, gt, bkgd, bound, rect, scratch=None):
    k = _abi.MossImageLoss()
    k.W, k.H = W, H
    k.image, k.alpha, k.gt_image, k.bkgd_mask, k.bound_mask, k.rect = _p(image), _p(alpha), _p(gt), _p(bkgd), _p(bound), _p(rect)
    k.w_l1, k.w_mask, k.w_ssim, k.w_s3im = 1.0, 0.5, 0.2, 0.3           # train_ZJU.py:131
    if scratch is not None:
        k.scratch, k.scratch_bytes = _p(scratch), scratch.numel()
    return k


@pytest.mark.parametrize("W,H", [(512, 512), (1000, 1002), (2048, 2048)])
def test_static_scratch_entry_point_equals_the_caller_scratch_one_twice_in_a_row(W, H):
    lib = _abi.load()
    stream = torch.cuda.current_stream().cuda_stream
    outs = []
    for seed in (3, 4):                                  # two calls in a row, different images
        t = _inputs(W, H, seed)
        scratch = torch.zeros(lib.moss_image_losses_scratch_bytes(W, H), dtype=torch.uint8, device=DEV)
        ref = [torch.full((8,), 7.0, device=DEV), torch.full((3, H, W), 7.0, device=DEV), torch.full((H, W), 7.0, device=DEV)]
        new = [torch.full((8,), -7.0, device=DEV), torch.full((3, H, W), -7.0, device=DEV), torch.full((H, W), -7.0, device=DEV)]
        k_ref, k_new = _struct(W, H, *t, scratch=scratch), _struct(W, H, *t)
        _abi.check(lib.moss_image_losses(stream, C.byref(k_ref), *[_p(x) for x in ref]), "moss_image_losses")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0] if outs else None      # from the second call on: not a byte of device memory taken
        _abi.check(lib.moss_image_losses_static(stream, C.byref(k_new), *[_p(x) for x in new]), "moss_image_losses_static")
        torch.cuda.synchronize()
        if free0 is not None:
            assert torch.cuda.mem_get_info()[0] == free0
        for a, b, what in zip(ref, new, ("losses", "dL_dimage", "dL_dalpha")):
            a, b = (a[:5], b[:5]) if what == "losses" else (a, b)
            assert torch.isfinite(b).all(), what
            assert torch.equal(a, b), (what, float((a - b).abs().max()))
        assert float(new[0][4]) > 0 and float(new[1].abs().max()) > 0
        outs.append(new[0][:5].clone())
    assert not torch.equal(outs[0], outs[1])              # the second call computed its own image, not the array's leftovers


def test_static_scratch_entry_point_rejects_what_does_not_fit_and_bad_arguments():
    lib = _abi.load()
    stream = torch.cuda.current_stream().cuda_stream
    W, H = 2080, 2048                                    # 65 x 64 = 4160 tiles > 4096
    img = torch.zeros(3, H, W, device=DEV)
    rect = torch.zeros(8, dtype=torch.int32, device=DEV)
    losses = torch.full((8,), 5.0, device=DEV)
    k = _struct(W, H, img, None, img, None, None, rect)
    k.w_mask = 0.0
    st = lib.moss_image_losses_static(stream, C.byref(k), _p(losses), None, None)
    assert st == _abi.MOSS_ERR_STATIC_SCRATCH_LIMIT
    assert "4096 tiles" in lib.moss_error_string(st).decode()
    with pytest.raises(RuntimeError, match="static scratch"):
        _abi.check(st, "moss_image_losses_static")
    torch.cuda.synchronize()
    assert torch.equal(losses, torch.full((8,), 5.0, device=DEV))        # nothing was launched
    k = _struct(64, 64, img, None, img, None, None, rect)                 # mask loss asked for without alpha / bkgd_mask
    assert lib.moss_image_losses_static(stream, C.byref(k), _p(losses), None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_image_losses_static(stream, None, _p(losses), None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
