"""Drop-in package: put ``<repo>/moss_b200/dropin`` (and ``<repo>``) on PYTHONPATH and MOSS's
``from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer``
(gaussian_renderer/__init__.py:16) resolves to the B200-native implementation."""
from moss_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                  rasterize_gaussians, _RasterizeGaussians)
