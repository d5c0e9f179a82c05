"""`CustomBruteForceMatcher` — host-side mirror of the reference class of the same name
(custom_bruteforce_matcher.py:13-1073) with the pairwise arithmetic on the B200.

Same keyword-only constructor, same public methods, same return contract: a row-major N x N
sequence of cv2.detail.MatchesInfo as cv.detail_BestOf2NearestMatcher.apply2 returns it, with
the (-1, -1) dummy on the diagonal and for rejected pairs.  What changes underneath:

  * the Hamming matching of ALL image pairs (reference: one cv.BFMatcher call per pair inside
    the loop :615-620) is one C-ABI call into CUDA kernel 1 issued in front of the pair loop;
  * the star-polygon fallback (:416-436) uses the engine-backed BFPolygonMatcher.

RANSAC homographies, confidences, the mirrored entries, the clamps and the bundle-adjuster
pruning stay on OpenCV / in Python exactly as the reference has them, with
its quirks Q1-Q6 preserved so every MatchesInfo field is identical.

Deliberate deviations (documented in DESIGN.md): the blocking `input()` of :700 honours
`disable_prompts`; debug JPEG writing can be switched off (SSP_B200_DEBUG_IMAGES=0 or
write_debug_images=False) so benchmarks time matching rather than JPEG encoding.
"""
import itertools
import os
import traceback
from collections import defaultdict

import cv2 as cv
import numpy as np

from .bf_matcher import as_descriptor_array, dmatches_from_arrays
from .custom_bf_polygon_matcher import BFPolygonMatcher, _debug_images_enabled
from .engine import Engine


def adjust_black_and_white_point(img, black_and_white_point_tpl):
    """GIMP-like levels (image_processors.py:32-41), used for the debug JPEGs only."""
    if not black_and_white_point_tpl:
        return img
    lo, hi = black_and_white_point_tpl
    return ((np.clip(img, lo, hi) - lo) * (255 / (hi - lo))).astype(np.uint8)


class CustomBruteForceMatcher:
    RANSAC_REPROJ_THRESHOLD_PX = 5   # custom_bruteforce_matcher.py:240
    MIN_MATCHES_FOR_HOMOGRAPHY = 6   # :215
    LOWE_RATIO = 0.75                # :164 (dead branch; selectable here)

    def __init__(self, *, imgs_features_orb_and_stars, imgs_features_len_orb, conf_thresh, img_names,
                 enforced_matches=[], enforced_no_matches=[], work_scale_imgs,
                 pairwise_matches_output_dir, black_and_white_point_adjustment_for_imshow=(0, 255),
                 disable_prompts=False, enforce_star_polygon_matcher=False,
                 output_dir_polygon_matching, predefined_overlaps, focal_length_pinhole,
                 engine: Engine = None, hamming_mode="crosscheck", hamming_variant="auto",
                 write_debug_images=None, verbose=True, polygon_candidate_generation="host",
                 ransac="opencv", ransac_hypotheses=4096, ransac_seed=0):
        self.imgs_features_orb_and_stars = imgs_features_orb_and_stars
        self.imgs_features_len_orb = imgs_features_len_orb
        self.conf_thresh = conf_thresh
        self.img_names = img_names
        self.enforced_matches = [set(m) for m in enforced_matches]
        self.enforced_no_matches = [set(m) for m in enforced_no_matches]
        self.confidence_when_enforced = 10
        self.work_scale_imgs = work_scale_imgs
        self.pairwise_matches_output_dir = pairwise_matches_output_dir
        self.black_and_white_point_adjustment_for_imshow = black_and_white_point_adjustment_for_imshow
        self.disable_prompts = disable_prompts
        self.output_dir_polygon_matching = output_dir_polygon_matching
        self.predefined_overlaps = [set(m) for m in predefined_overlaps]
        self.enforce_star_polygon_matcher = enforce_star_polygon_matcher
        self.kind_of_pairwise_matches = dict()
        self.polygon_data_store = dict()
        self.fts_calculated_counter = list()
        self.focal_length_pinhole = focal_length_pinhole
        # engine side: one context per matcher instance (it replaces the per-matcher state of
        # custom_bruteforce_matcher.py:57-59) unless the caller shares one on purpose; a shared
        # context is safe too, see `_upload_descriptor_bank`
        self.engine = engine if engine is not None else Engine()
        self.hamming_mode = hamming_mode
        self.hamming_variant = hamming_variant
        self.write_debug_images = _debug_images_enabled(write_debug_images)
        self.verbose = verbose
        self.polygon_candidate_generation = polygon_candidate_generation  # "device": row N2 (BFPolygonMatcher)
        # "opencv": cv.findHomography per pair exactly as the reference (parity).  "device": row N3,
        # one batched RANSAC launch for all ORB pairs right behind kernel 1 (same model, not OpenCV's
        # random numbers: inlier sets agree closely, bits do not); the star path stays on OpenCV.
        if ransac not in ("opencv", "device"):
            raise ValueError("ransac must be 'opencv' or 'device'")
        self.ransac = ransac
        self.ransac_hypotheses = ransac_hypotheses
        self.ransac_seed = ransac_seed
        self._orb_models = dict()   # (src, dst) -> (H or None, inlier mask list) from the batched RANSAC
        self._bank_generation = None  # Engine.bank_generation of MY upload
        self._orb_matches = dict()  # (src, dst) -> tuple of cv2.DMatch, filled by prefetch

    def _log(self, *a):
        if self.verbose:
            print(*a)

    # ---- user overrides (custom_bruteforce_matcher.py:63-85) ------------------------------
    def _named_pairs_as_index_sets(self, named):
        return [set((self.img_names.index(a), self.img_names.index(b))) for a, b in named]

    def combination_is_enforced_match(self, idx_tpl):
        return set(idx_tpl) in self._named_pairs_as_index_sets(self.enforced_matches)

    def combination_is_enforced_no_match(self, idx_tpl):
        return set(idx_tpl) in self._named_pairs_as_index_sets(self.enforced_no_matches)

    def get_orb_keypoint_indices_only(self):
        pass

    def get_star_keypoint_indices_only(self):
        pass

    # ---- kernel 1 in front of the pair loop -----------------------------------------------
    def _upload_descriptor_bank(self):
        # somebody else (another matcher, the BFMatcher drop-in) may have replaced the bank of a
        # shared context since my upload: the generation token tells
        if self._bank_generation is not None and self._bank_generation == self.engine.bank_generation:
            return
        rows = []
        for ft, n_orb in zip(self.imgs_features_orb_and_stars, self.imgs_features_len_orb):
            d = as_descriptor_array(ft.descriptors)
            rows.append(d[0:n_orb, :])  # ORB rows only, star descriptors are ignored (Q3, :197-198)
        self._bank_generation = self.engine.upload_descriptor_bank(rows)

    def prefetch_hamming_matches(self, pairs):
        """One engine call for a whole list of (src, dst) pairs; results are cached as DMatch
        tuples in OpenCV's order for match_2_features."""
        pairs = [tuple(p) for p in pairs if tuple(p) not in self._orb_matches]
        if not pairs:
            return
        self._upload_descriptor_bank()
        if self.hamming_mode == "knn2_ratio":
            for s, d in pairs:
                if self.imgs_features_len_orb[d] < 2:
                    # the reference's `for m, n in knnMatch(...)` cannot unpack a single neighbour
                    raise ValueError("not enough values to unpack (expected 2, got 1)")
        res = self.engine.hamming_match_pairs(pairs, mode=self.hamming_mode, ratio=self.LOWE_RATIO,
                                              variant=self.hamming_variant)
        for k, p in enumerate(pairs):
            self._orb_matches[p] = dmatches_from_arrays(*res.pair(k))
        if self.ransac == "device":
            kp_xy = [np.array([kp.pt for kp in ft.keypoints[0:n_orb]], dtype=np.float32).reshape(-1, 2)
                     for ft, n_orb in zip(self.imgs_features_orb_and_stars, self.imgs_features_len_orb)]
            H, mask, inl = self.engine.ransac_homography_pairs(
                kp_xy, pairs, int(res.offsets[-1]), reproj_threshold=self.RANSAC_REPROJ_THRESHOLD_PX,
                n_hypotheses=self.ransac_hypotheses, seed=self.ransac_seed,
                min_matches=self.MIN_MATCHES_FOR_HOMOGRAPHY)
            for k, p in enumerate(pairs):
                lo, hi = int(res.offsets[k]), int(res.offsets[k + 1])
                if inl[k] < 0:
                    continue                       # fewer than 6 matches: the host branch handles it (Q2)
                model = None if np.isnan(H[k]).any() else H[k].copy()
                self._orb_models[p] = (model, mask[lo:hi].astype(np.int64).tolist())

    def _hamming_matches(self, src_idx, dst_idx):
        key = (src_idx, dst_idx)
        if key not in self._orb_matches:
            self.prefetch_hamming_matches([key])
        return self._orb_matches[key]

    # ---- MatchesInfo assembly (host, OpenCV) ------------------------------------------------
    def _forward_info(self, matches, ft1, ft2, src_idx, dst_idx, model=None):
        """RANSAC homography + confidence for one direction (:210-266 / :437-495).  `model` = (H, mask)
        of the batched device RANSAC for these very matches, when that mode is on."""
        if model is not None and len(matches) >= self.MIN_MATCHES_FOR_HOMOGRAPHY:
            H, mask_list = model
        elif len(matches) >= self.MIN_MATCHES_FOR_HOMOGRAPHY:
            src_pts = np.float32([ft1.keypoints[m.queryIdx].pt for m in matches]).reshape(-1, 1, 2)
            dst_pts = np.float32([ft2.keypoints[m.trainIdx].pt for m in matches]).reshape(-1, 1, 2)
            H, mask = cv.findHomography(src_pts, dst_pts, method=cv.RANSAC,
                                        ransacReprojThreshold=self.RANSAC_REPROJ_THRESHOLD_PX)
            mask_list = mask.ravel().tolist()
        else:
            H, mask_list = None, [0]  # Q2: length-1 mask regardless of the match count
        info = cv.detail.MatchesInfo()
        info.dst_img_idx = dst_idx
        info.src_img_idx = src_idx
        info.matches = tuple(matches)
        info.H = H
        info.inliers_mask = tuple(mask_list)
        info.num_inliers = int(sum(mask_list))
        # M. Brown, D. Lowe, "Automatic Panoramic Image Stitching using Invariant Features"
        info.confidence = info.num_inliers / (8 + 0.3 * len(matches))
        return info, mask_list

    @staticmethod
    def _mirrored_info(fwd, matches, mask_list, src_idx, dst_idx):
        """The (dst, src) entry: swapped DMatch indices, inverse homography (:268-291)."""
        back = cv.detail.MatchesInfo()
        back.dst_img_idx = src_idx
        back.src_img_idx = dst_idx
        flipped = []
        for m in matches:
            w = cv.DMatch()
            w.imgIdx = 0
            w.distance = m.distance
            w.queryIdx = m.trainIdx
            w.trainIdx = m.queryIdx
            flipped.append(w)
        back.matches = tuple(flipped)
        back.H = np.linalg.inv(fwd.H) if fwd.H is not None else None
        back.confidence = fwd.confidence
        back.inliers_mask = tuple(mask_list)
        back.num_inliers = int(sum(mask_list))
        return back

    def match_2_features(self, src_idx, dst_idx):
        """Pairwise match of two images: (MatchesInfo src->dst, MatchesInfo dst->src)
        (custom_bruteforce_matcher.py:93-570)."""
        ft1 = self.imgs_features_orb_and_stars[src_idx]
        ft2 = self.imgs_features_orb_and_stars[dst_idx]
        thr = self.conf_thresh

        matches = self._hamming_matches(src_idx, dst_idx)
        self._log(f"\n[{src_idx:>3}] {self.img_names[src_idx][:24]:>25} <-> [{dst_idx:>3}] "
                  f"{self.img_names[dst_idx][:24]:>25}: {len(matches):>4} ORB matches ({self.hamming_mode})")
        fwd, mask_list = self._forward_info(matches, ft1, ft2, src_idx, dst_idx,
                                            model=self._orb_models.get((src_idx, dst_idx)))
        back = self._mirrored_info(fwd, matches, mask_list, src_idx, dst_idx)
        pair = [fwd, back]

        if fwd.num_inliers > 0:
            self.plot_pairwise_match(pair, "no window title", prompt_dialog=False, write_to_disk=True,
                                     kind="ORB",
                                     omitted="omitted__" if (fwd.confidence < thr and fwd.num_inliers < 50) else "")
        if fwd.num_inliers > 50:
            # Q1: only the forward entry is lifted, the mirror keeps the raw confidence (:307-310)
            fwd.confidence = thr * 2

        if self.combination_is_enforced_match((src_idx, dst_idx)):
            if fwd.confidence <= thr:
                before = fwd.confidence
                fwd.confidence = self.confidence_when_enforced
                back.confidence = self.confidence_when_enforced
                self._log(f"enforced match [{src_idx}] [{dst_idx}]: confidence {before} -> {fwd.confidence}")
                self.plot_pairwise_match(
                    pair, f"Enforced match [{src_idx}] {self.img_names[src_idx]} / [{dst_idx}] "
                          f"{self.img_names[dst_idx]}: confidence {before} -> {fwd.confidence}",
                    prompt_dialog=(not self.disable_prompts), write_to_disk=False,
                    omitted="omitted__" if fwd.confidence < thr else "")
            else:
                self._log(f"enforced match [{src_idx}] [{dst_idx}]: confidence {fwd.confidence} is high enough")

        wants_stars = self.enforce_star_polygon_matcher or (
            (fwd.confidence < thr or fwd.num_inliers < 12) and set((src_idx, dst_idx)) in self.predefined_overlaps)
        if not wants_stars:
            self.kind_of_pairwise_matches[(src_idx, dst_idx)] = "orb"
            return tuple(pair)

        # ---- star-polygon fallback (:360-552) ------------------------------------------------
        self._log("ORB: confidence {:.3f}, {} inliers of {} matches ({:.1f} %) => star polygon matcher".format(
            fwd.confidence, fwd.num_inliers, len(matches), fwd.num_inliers / len(matches) * 100))
        orb_conf, orb_inl = fwd.confidence, fwd.num_inliers
        if not np.array_equal(self.work_scale_imgs[src_idx].shape, self.work_scale_imgs[dst_idx].shape):
            raise NotImplementedError("Images with different shapes are NOT supported yet.")
        poly = BFPolygonMatcher(
            output_dir_polygon_matching=self.output_dir_polygon_matching,
            filename_img1=self.img_names[src_idx], filename_img2=self.img_names[dst_idx],
            idx_img1=src_idx, idx_img2=dst_idx, img_shape=self.work_scale_imgs[src_idx].shape,
            focal_length_pinhole=self.focal_length_pinhole, engine=self.engine,
            write_debug_images=self.write_debug_images, candidate_generation=self.polygon_candidate_generation)
        matches = poly.match(ft1.keypoints, ft2.keypoints,
                             self.imgs_features_len_orb[src_idx], self.imgs_features_len_orb[dst_idx],
                             self.work_scale_imgs[src_idx], self.work_scale_imgs[dst_idx],
                             self.polygon_data_store, src_idx, dst_idx, self.fts_calculated_counter)
        fwd, mask_list = self._forward_info(matches, ft1, ft2, src_idx, dst_idx)
        self._log("stars: confidence {:.3f}, {} inliers (ORB had {:.3f}, {})".format(
            fwd.confidence, fwd.num_inliers, orb_conf, orb_inl))
        if fwd.num_inliers >= 6 and fwd.confidence > 0.01:
            fwd.confidence = int(thr * 3)  # Q4: integer clamp, applied before mirroring (:508-511)
        self.kind_of_pairwise_matches[(src_idx, dst_idx)] = "stars"
        back = self._mirrored_info(fwd, matches, mask_list, src_idx, dst_idx)
        pair = [fwd, back]
        if fwd.num_inliers > 0:
            self.plot_pairwise_match(pair, "no window title", prompt_dialog=False, write_to_disk=True,
                                     kind="STARS", omitted="omitted__" if fwd.confidence < thr else "")
        return tuple(pair)

    # ---- all pairs (custom_bruteforce_matcher.py:572-762) -----------------------------------
    def match_n_features(self):
        n = len(self.imgs_features_orb_and_stars)
        combos = list(itertools.combinations(range(n), 2))

        dummy = cv.detail.MatchesInfo()
        # The reference assigns H = None (:603).  Under cv2 4.13 a Python-built MatchesInfo whose
        # H was never given an array makes detail_HomographyBasedEstimator.apply segfault
        # (cv2's own apply2() dummies do not); an empty float64 array is the same "no homography"
        # (H.empty() in C++) and is safe.  Host glue only, outside the CUDA path.
        dummy.H = np.zeros((0, 0), np.float64)
        dummy.confidence = 0.0
        dummy.dst_img_idx = -1
        dummy.inliers_mask = tuple()
        dummy.matches = tuple()
        dummy.num_inliers = 0
        dummy.src_img_idx = -1

        # kernel 1 for every pair at once, in front of the sequential host loop
        self.prefetch_hamming_matches(combos)
        raw = {c: self.match_2_features(*c) for c in combos}

        if self.verbose:
            print("\npairwise matches by confidence:")
            for (s, d), v in sorted(raw.items(), key=lambda kv: kv[1][0].confidence, reverse=True):
                m = len(v[0].matches)
                print("{:>8}{:>6}{:>6}{:>30}{:>30}{:8.2f}{:>8}{:>8}{:>8.2f}".format(
                    "[omit]" if v[0].confidence <= self.conf_thresh else "", s, d, self.img_names[s][:29],
                    self.img_names[d][:29], v[0].confidence, v[0].num_inliers, m,
                    v[0].num_inliers / m if m > 0 else v[0].num_inliers))

        per_image = defaultdict(list)
        for (s, d), v in raw.items():
            per_image[s].append(v[0].confidence)
            per_image[d].append(v[0].confidence)
        lonely = [i for i, confs in per_image.items() if not any(c > self.conf_thresh for c in confs)]
        for i in lonely:
            print(f"no match above conf_thresh={self.conf_thresh} for image [{i}] {self.img_names[i]}: "
                  "the bundle adjuster will fail; consider removing it from the inputs")
        if lonely and not self.disable_prompts:
            input("Press [ENTER] to continue anyway.")

        table = {}
        for s, d in combos:
            fwd = raw[(s, d)][0]
            keep = (not self.combination_is_enforced_no_match((s, d))
                    and fwd.confidence > self.conf_thresh and fwd.num_inliers > 5)  # Q5
            table[(s, d)], table[(d, s)] = raw[(s, d)] if keep else (dummy, dummy)
        table[(-1, -1)] = dummy
        return tuple(table[key] for key in self.strict_order_for_pairwise_matches_list())

    def strict_order_for_pairwise_matches_list(self):
        """Row-major (i, j) order with (-1, -1) on the diagonal (:764-811)."""
        n = len(self.imgs_features_orb_and_stars)
        return [(i, j) if i != j else (-1, -1) for i in range(n) for j in range(n)]

    # ---- bundle-adjuster feasibility pruning (:813-1006) -----------------------------------
    def get_valid_pairwise_matches_that_will_pass_bundle_adjuster_wo_crashes(self):
        infos = list(self.match_n_features())
        valid = set(tuple(sorted((m.src_img_idx, m.dst_img_idx))) for m in infos
                    if m.src_img_idx != -1 and m.dst_img_idx != -1)
        by_id = {(m.src_img_idx, m.dst_img_idx): m for m in infos}
        for n_removed in range(0, len(valid)):
            self._log(f"trying all ways of removing {n_removed} match(es)")
            kept_sets = list(itertools.combinations(valid, len(valid) - n_removed))
            trials = [(kept, list(set(valid).difference(set(kept)))) for kept in kept_sets]
            # cheapest removals (lowest summed confidence) first
            for kept, removed in sorted(trials, key=lambda kr: sum(by_id[i].confidence for i in kr[1])):
                dummy = infos[0]
                kept_as_sets = [set(k) for k in kept]
                pruned = [m if set([m.dst_img_idx, m.src_img_idx]) in kept_as_sets else dummy for m in infos]
                if self.test_camera_adjustment(self.imgs_features_orb_and_stars, pruned):
                    for bad in removed:
                        self._log(f"possible bad match [{bad[0]}] {self.img_names[bad[0]]} / [{bad[1]}] "
                                  f"{self.img_names[bad[1]]}: confidence = {by_id[bad].confidence}")
                    self._log("polygons were computed for images:", self.fts_calculated_counter)
                    return pruned
        raise ValueError("No success in finding a camera adjustment.")

    def test_camera_adjustment(self, features, pairwise_matches):
        estimator = cv.detail_HomographyBasedEstimator()
        try:
            ok, cameras = estimator.apply(features, pairwise_matches, None)
        except Exception:
            traceback.print_exc()
            return False
        if not ok:
            print("Homography estimation failed.")
            exit()
        for cam in cameras:
            cam.R = cam.R.astype(np.float32)
        adjuster = cv.detail_BundleAdjusterRay()
        adjuster.setConfThresh(self.conf_thresh)
        refine = np.zeros((3, 3), np.uint8)
        refine[0, 0] = refine[0, 1] = refine[0, 2] = refine[1, 1] = refine[1, 2] = 1  # "xxxxx"
        adjuster.setRefinementMask(refine)
        ok, cameras = adjuster.apply(features, pairwise_matches, cameras)
        return ok

    # ---- debug output (:1008-1073) ----------------------------------------------------------
    def plot_pairwise_match(self, pairwise_match, title_text, prompt_dialog=False, write_to_disk=False,
                            kind="", omitted=""):
        if not (prompt_dialog or (write_to_disk and self.write_debug_images)):
            return
        p = pairwise_match[0]
        a, b = self.work_scale_imgs[p.src_img_idx], self.work_scale_imgs[p.dst_img_idx]
        canvas = np.empty((max(a.shape[0], b.shape[0]), a.shape[1] + b.shape[1], 3), dtype=np.uint8)
        levels = self.black_and_white_point_adjustment_for_imshow
        cv.drawMatches(adjust_black_and_white_point(a, levels),
                       self.imgs_features_orb_and_stars[p.src_img_idx].keypoints,
                       adjust_black_and_white_point(b, levels),
                       self.imgs_features_orb_and_stars[p.dst_img_idx].keypoints,
                       p.matches, canvas, matchesMask=p.inliers_mask, matchesThickness=1,
                       matchColor=(255, 255, 255), singlePointColor=(255, 0, 0),
                       flags=cv.DrawMatchesFlags_NOT_DRAW_SINGLE_POINTS)
        if write_to_disk and self.write_debug_images:
            os.makedirs(self.pairwise_matches_output_dir, exist_ok=True)
            name = "{}___{}___conf={:.5f}___num_inliers={}___{}___{}.jpg".format(
                kind, omitted, p.confidence, p.num_inliers, self.img_names[p.src_img_idx],
                self.img_names[p.dst_img_idx])
            cv.imwrite(os.path.join(self.pairwise_matches_output_dir, name), canvas)
        if prompt_dialog:
            cv.imshow(f"{title_text}. Combining {p.src_img_idx} and {p.dst_img_idx}", canvas)
            cv.waitKey()
