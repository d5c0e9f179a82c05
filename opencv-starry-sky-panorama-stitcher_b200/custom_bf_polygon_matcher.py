"""`BFPolygonMatcher` — host-side mirror of the reference class of the same name
(custom_bf_polygon_matcher.py:19-1221) with the arithmetic on the B200:

  get_star_brightness (image_processors.py:222-254)      -> ssp_star_brightness   (row N1)
  per-polygon measure + normalise (cbpm:500-600, 722-740) -> ssp_polygon_measure   (kernel 2)
  all-pairs compare + argmin + threshold (cbpm:992-1148)  -> ssp_polygon_match_pairs (kernel 3)

What stays on the host, exactly as in the reference because it is tiny and order-sensitive:
the 12 border regions and Ward clustering that pick candidate stars (cbpm:261-443, sklearn)
and the keypoint voting that turns polygon matches into cv2.DMatch (cbpm:1160-1221).

Same constructor, same `match(...)` signature, same `polygon_data_store` layout, same
DMatch order as the reference, so it drops into custom_bruteforce_matcher.py:416-436.
"""
import itertools
import math
import os
from collections import defaultdict

import cv2 as cv
import numpy as np

from .bf_matcher import shared_engine
from .engine import Engine

_SESSION_KEY = "__ssp_b200_session__"


def _debug_images_enabled(flag):
    if flag is not None:
        return bool(flag)
    return os.environ.get("SSP_B200_DEBUG_IMAGES", "1") not in ("0", "false", "False")


def disc_offsets(radius: int = 10):
    """Pixel offsets (dx, dy) of cv.circle(center, radius, thickness=-1): OpenCV's own raster."""
    size = 2 * radius + 21
    c = size // 2
    mask = np.zeros((size, size), np.uint8)
    cv.circle(mask, (c, c), int(radius), (255, 255, 255), thickness=-1)
    ys, xs = np.nonzero(mask)
    return np.stack([xs - c, ys - c], axis=1).astype(np.int32)


class _Session:
    """Star tables of the images one engine has seen (lives inside polygon_data_store)."""

    def __init__(self, engine: Engine):
        self.engine = engine
        self.star_xy = {}
        self.star_brightness = {}
        self.shape_focal = None
        self.on_device = set()  # image ids whose polygon bank is resident in the engine

    def set_stars(self, img_id, xy, brightness, img_shape, focal):
        self.star_xy[img_id] = np.asarray(xy, dtype=np.float64).reshape(-1, 2)
        self.star_brightness[img_id] = np.asarray(brightness, dtype=np.float64)
        self.shape_focal = (int(img_shape[1]), int(img_shape[0]), float(focal))
        n = max(self.star_xy) + 1
        xy_l = [self.star_xy.get(i, np.zeros((0, 2))) for i in range(n)]
        br_l = [self.star_brightness.get(i, np.zeros(0)) for i in range(n)]
        w, h, f = self.shape_focal
        self.engine.upload_star_bank(xy_l, br_l, w, h, f, owner=self)

    def resident(self, img_id):
        """True while the engine still holds THIS session's polygon bank of `img_id` (another
        polygon_data_store on the same context may have measured its own image of that id)."""
        return img_id in self.on_device and self.engine.poly_owner.get(img_id) is self


class BFPolygonMatcher:
    # constants the reference hard-codes
    N_SIDES = 5                       # cbpm:704
    MIN_SIDE_PX = 5                   # cbpm:515
    MAX_POLYGONS = 15000              # cbpm:264
    VALID_DISTANCE = 0.1566           # cbpm:1047
    START_STARS_PER_REGION = 15       # cbpm:379
    BRIGHTNESS_RADIUS = 10            # image_processors.py:246
    BRIGHTNESS_TOP_K = 10             # image_processors.py:253

    def __init__(self, output_dir_polygon_matching, filename_img1, filename_img2, idx_img1, idx_img2,
                 img_shape, focal_length_pinhole, engine: Engine = None, write_debug_images=None,
                 min_over_shifts=False, fp32=False, candidate_generation="host"):
        self.filename_img1 = filename_img1
        self.filename_img2 = filename_img2
        self.idx_img1 = idx_img1
        self.idx_img2 = idx_img2
        self.focal_length_pinhole = focal_length_pinhole
        self.img_shape = img_shape
        self.engine = engine
        self.write_debug_images = _debug_images_enabled(write_debug_images)
        self.min_over_shifts = min_over_shifts
        self.fp32 = fp32
        # "host": the reference's own itertools / set() enumeration (parity: polygon ORDER decides
        # DMatch order).  "device": regions, per-cluster selection and the C(k, 5) enumeration run
        # on the GPU (row N2); the candidate SET is the reference's, its order is region-major.
        if candidate_generation not in ("host", "device"):
            raise ValueError("candidate_generation must be 'host' or 'device'")
        self.candidate_generation = candidate_generation
        self.output_dir = os.path.join(output_dir_polygon_matching, f"{filename_img1}___{filename_img2}")
        if self.write_debug_images:
            os.makedirs(self.output_dir, exist_ok=True)
        self._brightness_cache = {}

    # ---- small host-side geometry helpers (API parity with cbpm:70-135) ------------------
    def dist_between_2_points(self, p1, p2):
        return math.sqrt((p2[0] - p1[0]) ** 2 + (p2[1] - p1[1]) ** 2)

    def unit_vector(self, vector):
        return vector / np.linalg.norm(vector)

    def angle_between(self, v1, v2):
        return np.arccos(np.clip(np.dot(self.unit_vector(v1), self.unit_vector(v2)), -1.0, 1.0))

    def angle_distance(self, p1, p2):
        hw, hh = int(self.img_shape[1] / 2), int(self.img_shape[0] / 2)
        f = self.focal_length_pinhole
        return self.angle_between(np.array((p1[0] - hw, p1[1] - hh, f)), np.array((p2[0] - hw, p2[1] - hh, f)))

    def sphere_angle(self, p1_px, p2_px, p3_px):
        b = self.angle_distance(p1_px, p2_px)
        c = self.angle_distance(p2_px, p3_px)
        a = self.angle_distance(p3_px, p1_px)
        return math.acos((math.cos(a) - math.cos(b) * math.cos(c)) / (math.sin(b) * math.sin(c)))

    # ---- engine plumbing ----------------------------------------------------------------
    def _engine(self):
        return self.engine if self.engine is not None else shared_engine()

    def _session(self, polygon_data_store) -> _Session:
        sess = polygon_data_store.get(_SESSION_KEY)
        if sess is None or sess.engine is not self._engine():
            sess = _Session(self._engine())
            polygon_data_store[_SESSION_KEY] = sess
        return sess

    def star_brightnesses(self, keypoints, workscale_image):
        """get_star_brightness for every keypoint of one image in one kernel launch."""
        key = (id(workscale_image), len(keypoints))
        hit = self._brightness_cache.get(key)
        if hit is not None:
            return hit
        gray = cv.cvtColor(workscale_image, cv.COLOR_BGR2GRAY)
        centers = np.array([[int(kp.pt[0]), int(kp.pt[1])] for kp in keypoints], dtype=np.int32).reshape(-1, 2)
        vals = self._engine().star_brightness(gray, centers, disc_offsets(self.BRIGHTNESS_RADIUS),
                                              self.BRIGHTNESS_TOP_K)
        self._brightness_cache[key] = vals
        return vals

    # ---- candidate stars (host, cbpm:137-443) -------------------------------------------
    def reduce_amount_of_keypoints_by_clustering(self, indexed_keypoints, img_filename, work_scale_img,
                                                 n_clusters=30, brightness_by_id=None):
        """Ward clustering into n_clusters groups, keep the brightest star of each group
        (cbpm:137-259).  Returns {kp_id: keypoint} in cluster-label order like the reference."""
        from sklearn.cluster import AgglomerativeClustering
        ids = list(indexed_keypoints.keys())
        kps = list(indexed_keypoints.values())
        coords = np.array([kp.pt for kp in kps])
        labels = AgglomerativeClustering(n_clusters=n_clusters).fit(coords).labels_
        if brightness_by_id is None:
            vals = self.star_brightnesses(kps, work_scale_img)
            bright = {i: vals[i] for i in range(len(kps))}
        else:
            bright = {i: brightness_by_id[ids[i]] for i in range(len(kps))}
        survivors = {}
        for label in set(list(labels)):
            members = [i for i in range(len(kps)) if labels[i] == label]
            # ascending stable sort then reversed: the last of equally bright stars wins (cbpm:241)
            best = sorted(((i, bright[i]) for i in members), key=lambda x: x[1])[::-1][0][0]
            survivors[ids[best]] = kps[best]
        return survivors

    def _regions(self, keypoints, img_w, img_h):
        """The 12 overlapping border regions of cbpm:328-375 (strict comparisons)."""
        vh, vw, hh, hw = 0.70, 0.33, 0.33, 0.50
        x = [kp.pt[0] for kp in keypoints]
        y = [kp.pt[1] for kp in keypoints]
        ids = range(len(keypoints))
        left = lambda i: x[i] < vw * img_w
        right = lambda i: x[i] > (1 - vw) * img_w
        v_top = lambda i: y[i] < vh * img_h
        v_bot = lambda i: y[i] > (1 - vh) * img_h
        v_mid = lambda i: abs(y[i] - 0.5 * img_h) < vh * img_h / 2
        top = lambda i: y[i] < hh * img_h
        bot = lambda i: y[i] > (1 - hh) * img_h
        h_left = lambda i: x[i] < hw * img_w
        h_right = lambda i: x[i] > (1 - hw) * img_w
        h_mid = lambda i: abs(x[i] - 0.5 * img_w) < hw * img_w / 2
        spec = [("left_top", left, v_top), ("left_bot", left, v_bot), ("left_mid", left, v_mid),
                ("right_top", right, v_top), ("right_bot", right, v_bot), ("right_mid", right, v_mid),
                ("top_left", top, h_left), ("top_right", top, h_right), ("top_mid", top, h_mid),
                ("bot_left", bot, h_left), ("bot_right", bot, h_right), ("bot_mid", bot, h_mid)]
        return {name: [i for i in ids if a(i) and b(i)] for name, a, b in spec}

    def reduce_amount_of_keypoints_and_create_polygon_spanning_tuples_of_keypoints(
            self, keypoints, img_filename, workscale_img, n_polygon_sides,
            maximum_number_of_polygons_per_image_region=15000, _brightness=None):
        """Candidate 5-tuples (cbpm:261-443): per region keep at most k stars (k = 15, 14, ...)
        until the union of all C(region, 5) has between 1 and the maximum number of tuples.
        `_brightness` (not in the reference) supplies get_star_brightness values already known,
        as the synthetic C4 workload does."""
        vals = self.star_brightnesses(keypoints, workscale_img) if _brightness is None else _brightness
        brightness = {i: float(vals[i]) for i in range(len(keypoints))}
        regions = self._regions(keypoints, workscale_img.shape[1], workscale_img.shape[0])
        polygons = []
        k = self.START_STARS_PER_REGION
        while len(polygons) == 0 or len(polygons) > maximum_number_of_polygons_per_image_region:
            for name in regions:
                members = {i: keypoints[i] for i in regions[name]}
                if k >= len(members):
                    continue
                kept = self.reduce_amount_of_keypoints_by_clustering(
                    members, f"{img_filename}__{name}.jpg", workscale_img, n_clusters=k,
                    brightness_by_id=brightness)
                regions[name] = list(kept.keys())
            per_region = [list(itertools.combinations(regions[name], n_polygon_sides)) for name in regions]
            polygons = set([t for sub in per_region for t in sub])
            k -= 1
        return polygons

    def device_candidates(self, sess, img_id, keypoints, workscale_img, n_polygon_sides=5,
                          maximum_number_of_polygons_per_image_region=15000):
        """cbpm:261-443 with everything but the Ward clustering on the GPU (row N2).  The star bank of
        `img_id` must be on the engine (`sess.set_stars`).  Same reduction loop as the reference:
        k = 15, 14, ... stars per region until the union of all C(region, 5) has between 1 and the
        maximum number of tuples.  Returns (n candidates, device address of the int32 (n, 5) table,
        stars kept per region)."""
        from sklearn.cluster import AgglomerativeClustering
        if n_polygon_sides != self.N_SIDES:
            raise NotImplementedError("the candidate kernels enumerate 5-combinations (cbpm:704)")
        eng = sess.engine
        masks = eng.star_regions(img_id)
        coords = np.array([kp.pt for kp in keypoints]).reshape(-1, 2)
        regions = [np.nonzero(masks & (1 << r))[0].tolist() for r in range(12)]   # ascending ids, like filter()
        k = self.START_STARS_PER_REGION
        while True:
            if k < 1:
                raise ValueError("no image region is left with five stars: no polygon can be spanned")
            todo = [r for r in range(12) if len(regions[r]) > k]
            if todo:
                members, labels, off = [], [], [0]
                for r in todo:
                    members += regions[r]
                    labels += AgglomerativeClustering(n_clusters=k).fit(coords[regions[r]]).labels_.tolist()
                    off.append(len(members))
                surv = eng.cluster_brightest(img_id, members, labels, off, k)
                for g, r in enumerate(todo):
                    used = sorted(set(labels[off[g]:off[g + 1]]))        # `for cluster_id in unique_labels`
                    regions[r] = [int(surv[g, lab]) for lab in used]
            # the loop condition is on len(set of TUPLES) as in the reference (cbpm:383); the table holds
            # one tuple per distinct star set
            n_tuples, n, ptr = eng.polygon_candidates(regions, maximum_number_of_polygons_per_image_region)
            if ptr:
                return n, ptr, regions
            k -= 1

    # ---- kernel 2 -----------------------------------------------------------------------
    def get_polygons_characteristics(self, keypoints, workscale_image, img_title, n_sides,
                                     all_possible_polygons, _img_id=None, _session=None):
        """Measures all candidate polygons on the GPU; returns the reference's dict of arrays
        (cbpm:461-470), un-normalised.  Candidates are measured in iteration order of
        `all_possible_polygons`; the same star set listed twice is kept once (the reference's
        dict keyed by hull order does the same, cbpm:593)."""
        if n_sides != self.N_SIDES:
            raise NotImplementedError("the polygon-measure kernel is built for 5-gons (cbpm:704)")
        sess = _session if _session is not None else _Session(self._engine())
        img_id = 0 if _img_id is None else _img_id
        xy = np.array([kp.pt for kp in keypoints], dtype=np.float64).reshape(-1, 2)
        brightness = self.star_brightnesses(keypoints, workscale_image).astype(np.float64)
        sess.set_stars(img_id, xy, brightness, workscale_image.shape, self.focal_length_pinhole)
        seen = set()
        cands = []
        for tpl in all_possible_polygons:
            key = tuple(sorted(tpl))
            if key not in seen:
                seen.add(key)
                cands.append(tpl)
        cand_arr = np.array(cands, dtype=np.int32).reshape(-1, 5)
        eng = sess.engine
        eng.polygon_measure(img_id, cand_arr, min_side_px=float(self.MIN_SIDE_PX), owner=sess)
        sess.on_device.add(img_id)
        got = eng.polygon_fetch(img_id, raw=True)
        idx = got["star_idx"].astype(np.int64)
        out = {
            "keypoint_indices_rotated_list": idx,
            "keypoint_x_coordinates_list": xy[idx, 0] if len(idx) else np.zeros((0, 5)),
            "keypoint_y_coordinates_list": xy[idx, 1] if len(idx) else np.zeros((0, 5)),
            "keypoint_angles_rotated_list": got["angle"],
            "keypoint_brightnesses_rotated_list": got["brightness"],
            "side_lengths_rotated_list": got["side"],
        }
        out["_normalised_features"] = got["feat"]
        return out

    def _measure_with_device_candidates(self, sess, img_id, stars, img):
        """Row N2 + kernel 2 without a host round trip of the candidate table."""
        xy = np.array([kp.pt for kp in stars], dtype=np.float64).reshape(-1, 2)
        brightness = self.star_brightnesses(stars, img).astype(np.float64)
        sess.set_stars(img_id, xy, brightness, img.shape, self.focal_length_pinhole)
        eng = sess.engine
        n, ptr, _ = self.device_candidates(sess, img_id, stars, img, self.N_SIDES, self.MAX_POLYGONS)
        eng.poly_owner[img_id] = sess
        eng.polygon_measure_resident(img_id, ptr, n, min_side_px=float(self.MIN_SIDE_PX))
        eng.poly_owner[img_id] = sess      # polygon_measure_resident resets the owner
        sess.on_device.add(img_id)
        got = eng.polygon_fetch(img_id, raw=True)
        idx = got["star_idx"].astype(np.int64)
        polys = {
            "keypoint_indices_rotated_list": idx,
            "keypoint_x_coordinates_list": xy[idx, 0] if len(idx) else np.zeros((0, 5)),
            "keypoint_y_coordinates_list": xy[idx, 1] if len(idx) else np.zeros((0, 5)),
            "keypoint_angles_rotated_list": got["angle"],
            "keypoint_brightnesses_rotated_list": got["brightness"],
            "side_lengths_rotated_list": got["side"],
            "_normalised_features": got["feat"],
        }
        cands = set(map(tuple, eng.copy_to_host(ptr, (n, 5), np.int32).tolist()))
        return cands, polys

    def _prepare_image(self, sess, store, counter, img_id, keypoints, n_orb, img, filename, title):
        if img_id in store:
            entry = store[img_id]
            if not sess.resident(img_id):
                # polygon data computed elsewhere (a resumed store), or the bank of this image id
                # was replaced by another store sharing the context: load it into the engine
                p = entry["polygons_from_ftX"]
                feat = np.concatenate([p["keypoint_angles_rotated_list"], p["side_lengths_rotated_list"],
                                       p["keypoint_brightnesses_rotated_list"]], axis=1)
                sess.engine.polygon_upload_features(img_id, feat, p["keypoint_indices_rotated_list"], owner=sess)
                sess.on_device.add(img_id)
            return entry
        counter.append(img_id)
        stars = keypoints[n_orb:]  # ORB keypoints come first (cbpm:710)
        if self.candidate_generation == "device":
            cands, polys = self._measure_with_device_candidates(sess, img_id, stars, img)
        else:
            cands = self.reduce_amount_of_keypoints_and_create_polygon_spanning_tuples_of_keypoints(
                stars, filename, img, self.N_SIDES, self.MAX_POLYGONS)
            polys = self.get_polygons_characteristics(stars, img, title, self.N_SIDES, cands,
                                                      _img_id=img_id, _session=sess)
        feat = polys.pop("_normalised_features")
        # the reference normalises in place (cbpm:722-740); kernel 2 already did: angle / pi,
        # side / row max, brightness / row max
        polys["keypoint_angles_rotated_list"] = feat[:, 0:5].copy()
        polys["side_lengths_rotated_list"] = feat[:, 5:10].copy()
        polys["keypoint_brightnesses_rotated_list"] = feat[:, 10:15].copy()
        entry = {"ftX_kp_orb_removed": stars, "all_possible_polygons_ftX": cands, "polygons_from_ftX": polys}
        store[img_id] = entry
        return entry

    # ---- the public entry point (cbpm:636-1221) -----------------------------------------
    def match(self, ft1_kp, ft2_kp, ft1_kp_orb_len, ft2_kp_orb_len, workscale_img_1, workscale_img_2,
              polygon_data_store, img_id_1, img_id_2, fts_calculated_counter):
        sess = self._session(polygon_data_store)
        e1 = self._prepare_image(sess, polygon_data_store, fts_calculated_counter, img_id_1, ft1_kp,
                                 ft1_kp_orb_len, workscale_img_1, self.filename_img1, "img_title_1")
        e2 = self._prepare_image(sess, polygon_data_store, fts_calculated_counter, img_id_2, ft2_kp,
                                 ft2_kp_orb_len, workscale_img_2, self.filename_img2, "img_title_2")
        p1, p2 = e1["polygons_from_ftX"], e2["polygons_from_ftX"]

        if self.write_debug_images:
            self._dump_keypoints(e1, e2, ft1_kp_orb_len, ft2_kp_orb_len, workscale_img_1, workscale_img_2)

        # kernel 3: nearest polygon of image 2 for every polygon of image 1, thresholded
        res = sess.engine.polygon_match_pairs([(img_id_1, img_id_2)], threshold=self.VALID_DISTANCE,
                                              min_over_shifts=self.min_over_shifts, fp32=self.fp32)
        rows, cols, dists = res.pair(0)
        self.last_polygon_matches = (rows.copy(), cols.copy(), dists.copy())

        if self.write_debug_images:
            self._dump_polygon_matches(rows, cols, dists, e1, e2, ft1_kp_orb_len, ft2_kp_orb_len,
                                       workscale_img_1, workscale_img_2)

        # keypoint voting (cbpm:1160-1221): every matched polygon pair votes for its 5 vertex pairs
        idx1 = p1["keypoint_indices_rotated_list"]
        idx2 = p2["keypoint_indices_rotated_list"]
        votes = defaultdict(list)
        for r, c, d in zip(rows.tolist(), cols.tolist(), dists.tolist()):
            for pair in zip(list(idx1[r]), list(idx2[c])):
                votes[pair].append(d)
        ranked = sorted([(pair, len(ds), ds) for pair, ds in votes.items()], key=lambda x: x[1])[::-1]
        out = []
        for (kp1, kp2), n, ds in ranked:
            m = cv.DMatch()
            m.distance = sum(ds) / n
            m.imgIdx = 0
            m.queryIdx = int(kp1) + ft1_kp_orb_len
            m.trainIdx = int(kp2) + ft2_kp_orb_len
            out.append(m)
        return tuple(out)

    # ---- optional debug output (cbpm:804-895, :1069-1142, :1223-1368) ---------------------
    @staticmethod
    def _contrast(img):
        gray = cv.cvtColor(img, cv.COLOR_BGR2GRAY)
        try:
            from skimage import exposure
            return (exposure.equalize_adapthist(gray, clip_limit=0.0125) * 255).astype(np.uint8)
        except ImportError:
            return cv.createCLAHE(clipLimit=3.2, tileGridSize=(8, 8)).apply(gray)

    def imwrite_keypoints_of_2_images_to_disk(self, kps1, kps2, img1, img2, filename, circle_color,
                                              text_color, scale_circle=1, radii1=None, radii2=None,
                                              caption=None):
        canvas = cv.cvtColor(self._contrast(cv.hconcat([img1, img2])), cv.COLOR_GRAY2BGR)
        for side, (kps, radii) in enumerate(((kps1, radii1), (kps2, radii2))):
            shift = 0 if side == 0 else img1.shape[1]
            name = self.filename_img1 if side == 0 else self.filename_img2
            cv.putText(canvas, str(name), (20 + shift, img1.shape[0] - 30), cv.FONT_HERSHEY_SIMPLEX, 1,
                       (255, 255, 222), 1)
            for i, item in enumerate(kps):
                kp_id, kp = item[0], item[1]
                removed = len(item) == 3 and item[2] == 0
                if radii is not None:
                    radius = int(radii[i] * 30)
                else:
                    radius = max(1, int(10 / 24 * kp.size + 5)) * scale_circle
                col = (55, 155, 251) if removed else circle_color
                tcol = (0, 68, 135) if removed else text_color
                cv.circle(canvas, (int(kp.pt[0] + shift), int(kp.pt[1])), radius, col, 1, cv.LINE_AA)
                cv.putText(canvas, str(kp_id), (int(kp.pt[0] + shift + 5 + radius), int(kp.pt[1] - 5 - radius)),
                           cv.FONT_HERSHEY_SIMPLEX, 0.5, tcol, 1)
        if caption:
            cv.putText(canvas, caption, (25, img1.shape[0] - 70), cv.FONT_HERSHEY_SIMPLEX, 1, (0, 244, 154), 1)
        cv.imwrite(os.path.join(self.output_dir, filename), canvas)

    def _dump_keypoints(self, e1, e2, n_orb1, n_orb2, img1, img2):
        s1, s2 = e1["ftX_kp_orb_removed"], e2["ftX_kp_orb_removed"]
        all1 = [(i + n_orb1, kp) for i, kp in enumerate(s1)]
        all2 = [(i + n_orb2, kp) for i, kp in enumerate(s2)]
        red, dark = (0, 0, 255), (0, 0, 150)
        self.imwrite_keypoints_of_2_images_to_disk(
            all1, all2, img1, img2,
            f"00_all_star_keypoints_before_clustering_{self.filename_img1}___{self.filename_img2}"
            f"__{len(all1)}kps___{len(all2)}kps.jpg", red, dark, 3)
        keep1 = set(i for t in e1["all_possible_polygons_ftX"] for i in t)
        keep2 = set(i for t in e2["all_possible_polygons_ftX"] for i in t)
        rem1 = [(i + n_orb1, s1[i]) for i in keep1]
        rem2 = [(i + n_orb2, s2[i]) for i in keep2]
        self.imwrite_keypoints_of_2_images_to_disk(
            rem1, rem2, img1, img2,
            f"02_remaining_keypoints_after_clustering__{self.filename_img1}___{self.filename_img2}"
            f"__{len(rem1)}kps___{len(rem2)}kps.jpg", red, dark, 3)
        lab1 = [[i + n_orb1, kp, 1 if i in keep1 else 0] for i, kp in enumerate(s1)]
        lab2 = [[i + n_orb2, kp, 1 if i in keep2 else 0] for i, kp in enumerate(s2)]
        self.imwrite_keypoints_of_2_images_to_disk(
            lab1, lab2, img1, img2,
            f"01_all_star_keypoints_highlighted_if_removed_by_clustering_{self.filename_img1}___"
            f"{self.filename_img2}__{len(lab1)}kps___{len(lab2)}kps.jpg", red, dark, 3)

    def _dump_polygon_matches(self, rows, cols, dists, e1, e2, n_orb1, n_orb2, img1, img2, limit=100):
        p1, p2 = e1["polygons_from_ftX"], e2["polygons_from_ftX"]
        s1, s2 = e1["ftX_kp_orb_removed"], e2["ftX_kp_orb_removed"]
        order = np.argsort(dists, kind="stable")[:limit]
        for k in order.tolist():
            r, c, d = int(rows[k]), int(cols[k]), float(dists[k])
            ids1 = tuple(int(v) for v in p1["keypoint_indices_rotated_list"][r])
            ids2 = tuple(int(v) for v in p2["keypoint_indices_rotated_list"][c])
            ok = d < self.VALID_DISTANCE
            self.imwrite_keypoints_of_2_images_to_disk(
                tuple((i + n_orb1, s1[i]) for i in ids1), tuple((i + n_orb2, s2[i]) for i in ids2), img1, img2,
                "dist={:.4f}__polygon-{}_and_polygon{}__spanned_by_kps({})_and({}).jpg".format(
                    d, r, c, "-".join(str(i + n_orb1) for i in ids1), "-".join(str(i + n_orb2) for i in ids2)),
                (0, 244, 154), (0, 74, 47),
                radii1=p1["keypoint_brightnesses_rotated_list"][r],
                radii2=p2["keypoint_brightnesses_rotated_list"][c],
                caption="[{}] Euclidean distance = {:.6f} {} threshold ({:.6f})".format(
                    "ACCEPTED" if ok else "OMiTTED", d, "<" if ok else ">", self.VALID_DISTANCE))
