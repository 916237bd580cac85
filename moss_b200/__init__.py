"""moss_b200 -- B200-native (sm_100a) drop-in for MOSS's differentiable Gaussian-splat rasterizer.

Public surface (mirrors DGR/diff_gaussian_rasterization/__init__.py):
    GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians
plus the fused front end of gaussian_renderer/__init__.py:57-85: skinning.frame_tables, skinning.skin_and_covariance
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians,  # noqa: F401
                         forward_raw, backward_raw, forward_batch_raw, backward_batch_raw, rasterize_batch, set_gradient_sink, public_api_kind, export_state, set_segmented_backward,
                         add_densification_stats)

from . import skinning  # noqa: F401,E402  (fused LBS + covariance front end: skinning.skin_and_covariance)
from . import losses, fused  # noqa: F401,E402  (image losses behind the rasterizer; fused skin + rasterize)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]
