"""oracle/reference_shims.py — TEST INFRASTRUCTURE, used only in the build container.

Lets the UNCHANGED reference modules (/root/reference/custom_bruteforce_matcher.py,
custom_bf_polygon_matcher.py, image_processors.py) be imported where five of their third-party
packages are missing (skimage, shapely, alphashape, matplotlib, descartes), so that
tests/golden/make_golden.py can record the reference's own outputs as fixtures.  Nothing on
the GPU box reads /root/reference; only the recorded fixtures travel.

Stub behaviour:
  matplotlib.*, descartes.*            no-ops (debug plots only)
  skimage.exposure.equalize_adapthist  OpenCV CLAHE (upstream of the path: only debug JPEGs)
  alphashape.alphashape(points, 0)     convex hull polygon; alphashape 1.3.1 returns
                                       MultiPoint(points).convex_hull for alpha <= 0.  The hull
                                       here comes from scipy.spatial.ConvexHull (Qhull), i.e.
                                       independent of oracle/polygon_oracle.py's hull.
  shapely.algorithms.cga.signed_area   shoelace formula as in shapely 2.0.0
"""
import os
import sys
import types

import numpy as np

REFERENCE_DIR = "/root/reference"


class _Ring:
    def __init__(self, coords):
        self.coords = [tuple(c) for c in coords]  # closed ring

    @property
    def xy(self):
        return [c[0] for c in self.coords], [c[1] for c in self.coords]

    def __bool__(self):
        return len(self.coords) > 0


class _HullPolygon:
    geom_type = "Polygon"

    def __init__(self, ring_coords):
        self.exterior = _Ring(ring_coords)
        self.boundary = self.exterior
        self.interiors = []


def _alphashape(points, alpha):
    from scipy.spatial import ConvexHull
    assert alpha == 0, "the reference only calls alphashape(points, 0)"
    pts = np.asarray(points, dtype=np.float64)
    hull = ConvexHull(pts)  # 2-D: vertices in counter-clockwise order
    order = list(hull.vertices)
    # GEOS sorts the points before scanning, so its ring does not depend on the input order;
    # Qhull's start vertex does.  Start at the lowest point (min y, then min x) like the pivot of
    # GEOS' Graham scan, otherwise the reference's dict keyed by hull order (cbpm:593) could not
    # merge a star set that two regions list in different orders.
    k = min(range(len(order)), key=lambda j: (pts[order[j]][1], pts[order[j]][0]))
    order = order[k:] + order[:k]
    verts = [tuple(points[i]) for i in order]
    return _HullPolygon(verts + [verts[0]])


def _signed_area(ring):
    xs, ys = ring.xy
    xs = list(xs) + [xs[1]]
    ys = list(ys) + [ys[1]]
    return sum(xs[i] * (ys[i + 1] - ys[i - 1]) for i in range(1, len(ring.coords))) / 2.0


class _Anything:
    """Absorbs any attribute access / call (matplotlib, descartes)."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        return _Anything()

    def __iter__(self):
        return iter([_Anything(), _Anything()])

    def __getitem__(self, i):
        return _Anything()

    def __mul__(self, other):
        return self

    __rmul__ = __mul__


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__getattr__ = lambda attr: _Anything()  # type: ignore[attr-defined]
    sys.modules[name] = m
    return m


def _equalize_adapthist(img, clip_limit=0.01, **_):
    import cv2
    clahe = cv2.createCLAHE(clipLimit=max(1.0, clip_limit * 256), tileGridSize=(8, 8))
    return clahe.apply(img).astype(np.float64) / 255.0


def install():
    """Registers the stub modules and puts /root/reference on sys.path (idempotent)."""
    if not os.path.isdir(REFERENCE_DIR):
        raise FileNotFoundError(f"{REFERENCE_DIR} is only present in the build container")
    missing = []
    for name in ("skimage", "shapely", "alphashape", "matplotlib", "descartes"):
        try:
            __import__(name)
        except ImportError:
            missing.append(name)
    if "matplotlib" in missing:
        _module("matplotlib", pyplot=_module("matplotlib.pyplot"), path=_module("matplotlib.path", Path=_Anything()),
                patches=_module("matplotlib.patches"))
    if "descartes" in missing:
        _module("descartes", patch=_module("descartes.patch", Polygon=_Anything()))
    if "skimage" in missing:
        _module("skimage", exposure=_module("skimage.exposure", equalize_adapthist=_equalize_adapthist))
    if "alphashape" in missing:
        _module("alphashape", alphashape=_alphashape)
    if "shapely" in missing:
        cga = _module("shapely.algorithms.cga", signed_area=_signed_area)
        alg = _module("shapely.algorithms", cga=cga)
        _module("shapely", algorithms=alg)
    if REFERENCE_DIR not in sys.path:
        sys.path.insert(0, REFERENCE_DIR)
    return missing
