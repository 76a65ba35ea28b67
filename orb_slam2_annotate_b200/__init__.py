"""orb_slam2_annotate_b200 — B200-native ORB feature front end (extractor + Hamming matcher)
behind the ORB-SLAM2 ``ORBextractor`` / ``ORBmatcher`` interface.  The product is the C-ABI
library ``liborbfe.so`` (include/orbfe.h); this package is its Python host side."""
from ._lib import KP_DTYPE, OrbfeError, build, lib   # noqa: F401
from .extractor import ORBextractor, pinned_empty, pinned_free   # noqa: F401
from .matcher import BowSide, Frame, ORBmatcher, TriSide, project_last_frame, project_map_points   # noqa: F401
from .distributed import Comm, KNN_RESULT_DTYPE, shard_rows, unpack_results   # noqa: F401
