"""orbslam2_b200 — B200-native ORB front end (extraction + Hamming matching) behind the reference's class API.

Python host mirror of include/orbx.h; all compute runs in hand-written sm_100a CUDA kernels (liborbx.so)."""
from .extractor import ORBextractor, ExtractorRing, KP_DTYPE, compute_stereo_matches  # noqa: F401
from .matcher import ORBmatcher  # noqa: F401
from ._lib import OrbxError  # noqa: F401
from .frame import Frame  # noqa: F401
from .vocabulary import Vocabulary  # noqa: F401,E402
