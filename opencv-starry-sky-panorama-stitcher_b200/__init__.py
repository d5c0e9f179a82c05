"""B200-native pairwise image-matching engine for opencv-starry-sky-panorama-stitcher.

The directory name carries the reference's name (hyphens included), so the package is imported
through the alias module at the repo root:  `import ssp_b200`.

  ssp_b200.Engine                   numpy face of the C ABI (include/ssp_b200.h)
  ssp_b200.BFMatcher                drop-in for cv.BFMatcher(cv.NORM_HAMMING[, crossCheck=True])
  ssp_b200.BFPolygonMatcher         mirror of the reference's star-polygon matcher
  ssp_b200.CustomBruteForceMatcher  mirror of the reference's all-pairs matcher
"""
from . import _native
from .engine import Engine, EngineError, HammingResult, PolygonMatchResult
from .bf_matcher import BFMatcher, shared_engine
from .custom_bf_polygon_matcher import BFPolygonMatcher
from .custom_bruteforce_matcher import CustomBruteForceMatcher

__all__ = ["Engine", "EngineError", "HammingResult", "PolygonMatchResult", "BFMatcher",
           "BFPolygonMatcher", "CustomBruteForceMatcher", "shared_engine"]
