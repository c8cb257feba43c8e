"""Import-name alias so that `from rasterizer2 import GaussianRasterizationSettings2, GaussianRasterizer2`
(gaussian_renderer/__init__.py:15 of the reference) resolves to the B200-native implementation."""
from neuralgaussiansplatting_b200.rasterizer2 import *  # noqa: F401,F403
from neuralgaussiansplatting_b200.rasterizer2 import _GetIdxMap2  # noqa: F401
from neuralgaussiansplatting_b200 import _C  # noqa: F401
