"""Drop-in `diff_gauss` package: same import surface as the reference's
submodules/diff-gaussian-rasterization/diff_gauss (gaussian_renderer/__init__.py:15
imports GaussianRasterizationSettings and GaussianRasterizer from here), backed
by the B200-native libigs_raster.so.  See instag_b200/rasterizer.py."""
from instag_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    _RasterizeGaussians,
    cpu_deep_copy_tuple,
    rasterize_gaussians,
)
