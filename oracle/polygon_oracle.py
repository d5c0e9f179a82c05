"""oracle/polygon_oracle.py — TEST INFRASTRUCTURE, not product code.

CPU restatement (plain Python + numpy, float64) of the reference's star-polygon path.  Every
function cites the lines of /root/reference/custom_bf_polygon_matcher.py ("cbpm") or
image_processors.py it follows.

Third-party arithmetic that is NOT under /root/reference and is restated from its published
behaviour:
  * convex hull: alphashape 1.3.1 `alphashape(points, 0)` returns
    `shapely.geometry.MultiPoint(points).convex_hull` for alpha <= 0 (shapely 2.0.0 / GEOS:
    Graham scan with exact orientation predicates; interior and collinear points are not hull
    vertices).  Restated here as Andrew's monotone chain with exact rational orientation.
  * `shapely.algorithms.cga.signed_area(ring)`: shoelace sum, positive for counter-clockwise
    rings in (x, y) axes.  Only its sign is used (cbpm:537).

Parity pin: tests/test_oracle_polygon.py checks this file against (a) the known-answer vector
of SURVEY.md §8(c) (the reference's own functions run on its debug points cbpm:504-505) and
(b) tests/golden/polygon_golden.npz, produced by tests/golden/make_golden.py, which imports the
UNCHANGED reference module in the build container (under stub modules for the five packages
missing there) and records its outputs.  Ring start vertex on exact minimum-angle ties is
GEOS-internal and not pinned (documented near-tie exception).
"""
import math
from collections import defaultdict
from fractions import Fraction

import numpy as np


class PolygonOracle:
    def __init__(self, img_shape, focal_length_pinhole):
        # cbpm:34-59: img_shape is work_scale_img.shape = (H, W, 3)
        self.img_shape = tuple(img_shape)
        self.f = focal_length_pinhole
        # cbpm:102-103
        self.half_w = int(self.img_shape[1] / 2)
        self.half_h = int(self.img_shape[0] / 2)

    # ---- geometry (cbpm:70-135) ---------------------------------------------------------
    @staticmethod
    def pixel_distance(p1, p2):
        # cbpm:70-72
        return math.sqrt((p2[0] - p1[0]) ** 2 + (p2[1] - p1[1]) ** 2)

    def ray(self, p):
        # cbpm:105 + unit_vector cbpm:74-76
        v = np.array((p[0] - self.half_w, p[1] - self.half_h, self.f))
        return v / np.linalg.norm(v)

    def angle_distance(self, p1, p2):
        # cbpm:78-84, 86-108
        return float(np.arccos(np.clip(np.dot(self.ray(p1), self.ray(p2)), -1.0, 1.0)))

    def sphere_angle(self, p1, p2, p3):
        # cbpm:110-135: spherical law of cosines at vertex p2; math.acos raises ValueError
        # outside [-1, 1] exactly as in the reference.
        b = self.angle_distance(p1, p2)
        c = self.angle_distance(p2, p3)
        a = self.angle_distance(p3, p1)
        return math.acos((math.cos(a) - math.cos(b) * math.cos(c)) / (math.sin(b) * math.sin(c)))

    # ---- convex hull restatement --------------------------------------------------------
    @staticmethod
    def _orient(a, b, c):
        """Exact sign of the orientation of (a, b, c): > 0 counter-clockwise in (x, y) axes."""
        ax, ay, bx, by, cx, cy = (Fraction(v) for v in (a[0], a[1], b[0], b[1], c[0], c[1]))
        s = (bx - ax) * (cy - ay) - (by - ay) * (cx - ax)
        return (s > 0) - (s < 0)

    @classmethod
    def hull_ccw(cls, pts):
        """Strict convex hull vertices (indices into pts), counter-clockwise in (x, y) axes."""
        order = sorted(range(len(pts)), key=lambda i: (pts[i][0], pts[i][1]))
        if len(order) < 3:
            return order

        def chain(seq):
            out = []
            for i in seq:
                while len(out) >= 2 and cls._orient(pts[out[-2]], pts[out[-1]], pts[i]) <= 0:
                    out.pop()
                out.append(i)
            return out

        lower = chain(order)
        upper = chain(order[::-1])
        return lower[:-1] + upper[:-1]

    def canonical_ring(self, pts):
        """Ring order of cbpm:537-545 as indices into pts, or None when not all points are hull
        vertices (cbpm:524).  Orientation: negative shoelace area.  Start vertex: the lowest
        point (min y, then min x) — GEOS' own start vertex is not observable here and only
        matters for exact ties of the minimum angle."""
        hull = self.hull_ccw(pts)
        if len(hull) != len(pts):
            return None
        start = min(hull, key=lambda i: (pts[i][1], pts[i][0]))
        k = hull.index(start)
        ccw = hull[k:] + hull[:k]
        return [ccw[0]] + ccw[1:][::-1]

    # ---- per-polygon measure (cbpm:500-600) ---------------------------------------------
    def measure(self, pts, min_side_px=5):
        """Returns None (rejected) or dict(order, angles, sides) in canonical rotated order;
        `order` indexes into pts."""
        n = len(pts)
        # cbpm:515
        for i in range(n):
            for j in range(i + 1, n):
                if self.pixel_distance(pts[i], pts[j]) < min_side_px:
                    return None
        ring = self.canonical_ring(pts)
        if ring is None:
            return None
        # cbpm:547-555: vertex angle from (previous, current, next)
        angles = [self.sphere_angle(pts[ring[i - 1]], pts[ring[i]], pts[ring[(i + 1) % n]])
                  for i in range(n)]
        # cbpm:569-576: side i joins ring[i] and ring[i+1]
        sides = [self.angle_distance(pts[ring[i]], pts[ring[(i + 1) % n]]) for i in range(n)]
        # cbpm:578-591: rotate so the (first) smallest angle leads
        r = angles.index(min(angles))
        rot = lambda seq: list(seq[r:]) + list(seq[:r])
        return {"order": rot(ring), "angles": rot(angles), "sides": rot(sides)}

    def characteristics(self, star_xy, star_brightness, candidates, min_side_px=5):
        """get_polygons_characteristics (cbpm:445-634) + normalisation (cbpm:722-740) over a
        candidate list, in candidate order.  Returns dict of arrays like the reference's
        polygons_from_ftX plus the un-normalised measures and a per-candidate valid flag."""
        idx_l, ang_l, side_l, bri_l, valid = [], [], [], [], []
        seen = set()
        for cand in candidates:
            pts = [tuple(float(v) for v in star_xy[i]) for i in cand]
            m = self.measure(pts, min_side_px)
            # cbpm:593: results are stored in a dict keyed by the hull-ordered keypoint tuple, so
            # a star set that is listed twice (two regions, different tuple order) is kept once,
            # at the position of its first occurrence
            key = None if m is None else frozenset(int(c) for c in cand)
            fresh = m is not None and key not in seen
            valid.append(fresh)
            if not fresh:
                continue
            seen.add(key)
            kp = [int(cand[o]) for o in m["order"]]
            idx_l.append(kp)
            ang_l.append(m["angles"])
            side_l.append(m["sides"])
            bri_l.append([float(star_brightness[k]) for k in kp])
        idx = np.array(idx_l, dtype=np.int64).reshape(-1, 5)
        ang = np.array(ang_l, dtype=np.float64).reshape(-1, 5)
        side = np.array(side_l, dtype=np.float64).reshape(-1, 5)
        bri = np.array(bri_l, dtype=np.float64).reshape(-1, 5)
        feat = self.normalise(ang, side, bri)
        return {"star_idx": idx, "angle": ang, "side": side, "brightness": bri, "feat": feat,
                "valid": np.array(valid, dtype=bool)}

    @staticmethod
    def normalise(angle, side, brightness):
        # cbpm:723, 727-729, 738-740
        a = angle / math.pi
        s = (side.T / np.amax(side, axis=1)).T if len(side) else side
        b = (brightness.T / np.amax(brightness, axis=1)).T if len(brightness) else brightness
        return np.concatenate([a, s, b], axis=1) if len(a) else np.zeros((0, 15))

    # ---- compare / select / vote --------------------------------------------------------
    @staticmethod
    def compare(feat1, feat2):
        """polygon_similarity_matrix of cbpm:992-1034, shape (P1, P2); same numpy expressions
        (tile / repeat / square / sum / sqrt) so rounding is numpy's own."""
        p1, p2 = feat1.shape[0], feat2.shape[0]
        parts = []
        for g in range(3):
            a = feat1[:, 5 * g:5 * g + 5]
            b = feat2[:, 5 * g:5 * g + 5]
            parts.append(np.sum(np.square(np.tile(a, reps=(p2, 1)) - np.repeat(b, repeats=p1, axis=0)),
                                axis=1))
        overall = np.sqrt(parts[0] + parts[1] + parts[2])
        return np.reshape(overall, (p2, p1)).T

    @staticmethod
    def compare_min_over_shifts(feat1, feat2):
        """Non-parity mode of kernel 3: minimum over the 5 cyclic vertex shifts of polygon 2."""
        best = None
        for s in range(5):
            f2 = np.concatenate([np.roll(feat2[:, 5 * g:5 * g + 5], -s, axis=1) for g in range(3)], axis=1)
            d2 = PolygonOracle._dist2_seq(feat1, f2)
            best = d2 if best is None else np.minimum(best, d2)
        return np.sqrt(best)

    @staticmethod
    def _dist2_seq(feat1, feat2):
        tot = None
        for g in range(3):
            d = feat1[:, None, 5 * g:5 * g + 5] - feat2[None, :, 5 * g:5 * g + 5]
            s = d * d
            acc = s[..., 0]
            for k in range(1, 5):
                acc = acc + s[..., k]
            tot = acc if tot is None else tot + acc
        return tot

    @staticmethod
    def select(matrix, threshold=0.1566):
        # cbpm:1063-1067 and :1145-1152
        if matrix.shape[0] == 0 or matrix.shape[1] == 0:
            z = np.zeros(0, dtype=np.int64)
            return z, z, np.zeros(0)
        cols = np.argmin(matrix, axis=1)
        vals = matrix.min(axis=1)
        keep = vals < threshold
        return np.arange(matrix.shape[0])[keep], cols[keep], vals[keep]

    @staticmethod
    def vote(rows, cols, dists, idx1, idx2, n_orb1=0, n_orb2=0):
        """cbpm:1160-1221: returns [(queryIdx, trainIdx, distance)] in the reference's order
        (most-voted first; ties in reverse insertion order)."""
        votes = defaultdict(list)
        for r, c, d in zip(rows, cols, dists):
            for k1, k2 in zip(list(idx1[r]), list(idx2[c])):
                votes[(int(k1), int(k2))].append(float(d))
        ordered = sorted([(k, len(v), v) for k, v in votes.items()], key=lambda x: x[1])[::-1]
        return [(k[0] + n_orb1, k[1] + n_orb2, sum(v) / n) for k, n, v in ordered]


def star_brightness(gray, pt, radius=10, top_k=10):
    """get_star_brightness (image_processors.py:222-254) on an already grey image: filled
    circle mask at int(pt), sum of the top_k largest masked values.  Uses cv2.circle for the
    mask exactly like the reference."""
    import cv2
    mask = np.zeros(gray.shape, np.uint8)
    cv2.circle(mask, tuple(int(v) for v in pt), int(radius), (255, 255, 255), thickness=-1)
    masked = cv2.bitwise_and(gray, gray, mask=mask)
    return int(np.sum(np.sort(masked.flatten())[-top_k:]))
