"""dcvc_b200 -- B200-native (sm_100a) implementation of the DCVC-DC per-frame coding path behind
the reference's Python module API.  See DESIGN.md; the C ABI is include/dcvc_b200.h."""
from .image_model import IntraNoAR  # noqa: F401
from .video_model import DMC  # noqa: F401
