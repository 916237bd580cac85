"""Fused AdamW for the per-Gaussian parameter groups.

``FusedAdamW`` is a ``torch.optim.Optimizer`` with the constructor, ``param_groups`` and per-parameter ``state`` layout
(``step``, ``exp_avg``, ``exp_avg_sq``) of ``torch.optim.AdamW`` -- MOSS's densification code edits exactly those
entries (``cat_tensors_to_optimizer`` / ``_prune_optimizer`` / ``replace_tensor_to_optimizer``,
scene/gaussian_model.py) -- so it replaces ``torch.optim.AdamW(l, lr=0.0, eps=1e-15)`` (scene/gaussian_model.py:225)
for the groups that hold float32 CUDA tensors.  ``step()`` updates up to 8 such tensors per launch through the C ABI
(``moss_adamw_step``); parameters the kernel does not take (other dtypes, e.g. an MLP in half precision) are refused
loudly rather than silently updated by another code path.
"""
import ctypes as C

import torch

from . import _abi


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        if lr < 0.0 or eps < 0.0 or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or weight_decay < 0.0:
            raise ValueError("invalid AdamW hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        """One launch per 8 parameter tensors that share betas / eps / weight_decay: MOSS's six per-Gaussian groups
        differ only in ``lr`` (scene/gaussian_model.py:205-212), so a whole ``optimizer.step()`` is ONE kernel."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _abi.load()
        batches = {}
        for group in self.param_groups:
            key = (float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]), float(group["weight_decay"]))
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdamW takes contiguous float32 CUDA parameters (there is no CPU path)")
                g = p.grad
                if g.is_sparse or g.dtype != torch.float32 or g.device != p.device:
                    raise RuntimeError("FusedAdamW needs dense float32 gradients on the parameter's device")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                # the reference's densification code replaces these with torch.cat results: keep them dense
                st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"].contiguous(), st["exp_avg_sq"].contiguous()
                st["step"] = int(st["step"]) + 1
                batches.setdefault((p.device,) + key, []).append((p, g.contiguous(), st, float(group["lr"])))
        for (dev, beta1, beta2, eps, wd), items in batches.items():
            for i in range(0, len(items), 8):
                part = items[i:i + 8]
                arr = (_abi.MossAdamGroup * len(part))()
                for a, (p, g, st, lr) in zip(arr, part):
                    a.param, a.grad = p.data_ptr(), g.data_ptr()
                    a.exp_avg, a.exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                    a.numel, a.lr, a.step = p.numel(), lr, st["step"]
                with _abi.on_device(dev):
                    status = lib.moss_adamw_step(_abi.stream_ptr(dev), len(part), arr, beta1, beta2,
                                                 eps, wd)
                _abi.check(status, "moss_adamw_step")
        return loss
