"""DCVC-HEM (ACM MM 2022) inference models on the dcvc_b200 kernel set: the second consumer of
the per-frame coding path (BASELINE.json north_star).  Same module API as the reference's
DCVC-HEM/src/models: `IntraNoAR(N=192)`, `DMC()`."""
from .models import DMC, IntraNoAR  # noqa: F401
