"""B200-native (sm_100a) implementation of DPVO's per-iteration update hot path:
patch-to-feature-map correlation lookup (`altcorr`), sparse patch-graph bundle
adjustment (`fastba`) and the SE3 operators they use (`lietorch`), behind the
reference's own Python operator surface.

The directory name contains hyphens; import it with
    importlib.import_module("deep-visual-odometry-slam_b200")
or put `deep-visual-odometry-slam_b200/dropin` on sys.path to get drop-in
`cuda_corr`, `cuda_ba` and `lietorch_backends` modules for the unmodified
reference Python code.
"""
from . import _lib          # noqa: F401  (fails loudly if libdpvo_b200.so is missing)
from . import cuda_corr     # noqa: F401
from . import altcorr       # noqa: F401
from . import cuda_ba       # noqa: F401
from . import fastba        # noqa: F401
from . import lietorch_backends  # noqa: F401
from . import projective_ops     # noqa: F401
