from splatformer_b200.api import spherical_harmonics  # noqa: F401
from splatformer_b200.sh import deg_from_sh, num_sh_bases  # noqa: F401
