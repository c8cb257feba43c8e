"""Import-name alias so that `from diff_gaussian_rasterization import GaussianRasterizationSettings,
GaussianRasterizer` (gaussian_renderer/__init__.py:14 of the reference) resolves to the B200-native
implementation without touching the caller."""
from neuralgaussiansplatting_b200.diff_gaussian_rasterization import *  # noqa: F401,F403
from neuralgaussiansplatting_b200.diff_gaussian_rasterization import _RasterizeGaussians  # noqa: F401
from neuralgaussiansplatting_b200 import _C  # noqa: F401
