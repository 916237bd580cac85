"""moss_b200 -- B200-native (sm_100a) drop-in for MOSS's differentiable Gaussian-splat rasterizer.

Public surface (mirrors DGR/diff_gaussian_rasterization/__init__.py):
    GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians,  # noqa: F401
                         forward_raw, backward_raw, export_state)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]
