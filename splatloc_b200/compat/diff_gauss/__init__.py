"""Name shim: `import diff_gauss` -> splatloc_b200.diff_gauss."""
from splatloc_b200.diff_gauss import *  # noqa: F401,F403
from splatloc_b200.diff_gauss import (BLOCK_X, BLOCK_Y, GaussianRasterizationSettings,  # noqa: F401
                                      GaussianRasterizer, _RasterizeGaussians, interpret_binningBuffer,
                                      interpret_geomBuffer, interpret_imgBuffer, mark_visible,
                                      rasterize_gaussians)
from splatloc_b200 import _C  # noqa: F401
