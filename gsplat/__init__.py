"""`import gsplat` shim: resolves the three gsplat v0.1.11 entry points that SplatFormer uses
(/root/reference/utils/gs_utils.py:3,78,82,96) to the B200-native implementation, so that the
reference's `utils/gs_utils.py` and `train.py` run unchanged with this repository on PYTHONPATH."""
from splatformer_b200.api import (  # noqa: F401
    project_gaussians,
    rasterize_gaussians,
    spherical_harmonics,
)
from splatformer_b200.sh import deg_from_sh, num_sh_bases  # noqa: F401
from splatformer_b200.utils import (  # noqa: F401
    bin_and_sort_gaussians,
    compute_cov2d_bounds,
    compute_cumulative_intersects,
    get_tile_bin_edges,
    map_gaussian_to_intersects,
)
from . import sh, utils  # noqa: F401

__version__ = "0.1.11"
__all__ = ["project_gaussians", "rasterize_gaussians", "spherical_harmonics", "num_sh_bases",
           "deg_from_sh", "bin_and_sort_gaussians", "compute_cumulative_intersects",
           "get_tile_bin_edges", "map_gaussian_to_intersects", "compute_cov2d_bounds", "__version__"]
