"""Generates the committed fixtures in tests/golden/ by running the REFERENCE ITSELF.

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py

  orb_forest_c1.npz        BASELINE config 1 input: ORB(1000, edgeThreshold 10, patchSize 30)
                           descriptors + keypoints of the 21 forest JPEGs at work_megapix 1.2,
                           computed exactly like stitching_detailed_enhanced.py:744-793 (CLAHE via
                           the shim: skimage is absent here; it only decides WHICH descriptors
                           exist, not how they are matched).
  hamming_golden.npz       cv2.BFMatcher outputs (knnMatch k=2 and crossCheck match) on seeded
                           tie-heavy / ragged descriptor sets — the third-party arithmetic the
                           reference calls at custom_bruteforce_matcher.py:160-162 and :193-199.
  reference_matchesinfo_c1.npz   every MatchesInfo field the unchanged reference
                           CustomBruteForceMatcher.match_n_features() returns for the first 6
                           forest images (15 pairs).
  eisenberg_c2.npz         BASELINE config 2 on the 4 eisenberg frames the reference ships: stars
                           from the unchanged compute_star_features, DMatches and polygon banks
                           of the unchanged BFPolygonMatcher.match for all 6 pairs.
  reference_matchesinfo_c1_full.npz   the same for ALL 21 forest images (441 entries), every entry
                           additionally pushed through the reference's REAL result container
                           cv2_pickleable.detail.MatchesInfo(x).to_cv2() (cv2_pickleable/detail.py:
                           154-211, which validates the round trip itself) before it is recorded.
  eisenberg_c2_stage.npz   BASELINE config 2 as a WHOLE STAGE on the 4 eisenberg frames the reference
                           ships: ORB + star features merged as stitching_detailed_enhanced.py:784-
                           834 merges them, the unchanged CustomBruteForceMatcher.match_n_features()
                           with conf_thresh = 1.0 and example_06's predefined_overlaps re-indexed to
                           the 4 frames (run "overlaps": ORB -> RANSAC -> fallback decision -> star
                           MatchesInfo with the int(3 * conf) clamp and the mirror), a second run with
                           enforced_matches / enforced_no_matches ("enforced") and a third with
                           enforce_star_polygon_matcher=True ("allstars").  Every MatchesInfo of the
                           returned tuple AND the raw (forward, mirror) pair every match_2_features
                           call returned are recorded.
  polygon_golden.npz       unchanged reference BFPolygonMatcher on two synthetic overlapping
                           star views: get_polygons_characteristics arrays, the stored normalised
                           polygon_data_store arrays and the final DMatches of match().
"""
import builtins
import os
import sys
import tempfile

import cv2 as cv
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.dirname(os.path.abspath(__file__))

from oracle import reference_shims  # noqa: E402

reference_shims.install()
import custom_bf_polygon_matcher as ref_poly  # noqa: E402  (the unchanged reference)
import custom_bruteforce_matcher as ref_bf  # noqa: E402
import image_processors as ref_ip  # noqa: E402

import ssp_b200.synthetic as synth  # noqa: E402  (generators only, no engine)

FOREST = "/root/reference/img_autumn_forest_a_8+8+4+1_shots"


def forest_features(names):
    """stitching_detailed_enhanced.py:744-793 for the ORB-only configuration of example_01b."""
    finder = cv.ORB_create(nfeatures=1000, edgeThreshold=10, patchSize=30)
    feats, imgs, scale = [], [], None
    for name in names:
        full = cv.imread(os.path.join(FOREST, name))
        if scale is None:
            scale = min(1.0, np.sqrt(1.2 * 1e6 / (full.shape[0] * full.shape[1])))
        img = cv.resize(src=full, dsize=None, fx=scale, fy=scale, interpolation=cv.INTER_AREA)
        imgs.append(img)
        feats.append(cv.detail.computeImageFeatures2(finder, ref_ip.optimize_img_for_feature_detection(img)))
    return feats, imgs


def make_orb_fixture():
    names = sorted(n for n in os.listdir(FOREST) if n.endswith(".jpg"))
    feats, imgs = forest_features(names)
    desc = [f.descriptors.get() for f in feats]
    pts = [np.array([kp.pt for kp in f.keypoints], dtype=np.float32) for f in feats]
    offs = np.zeros(len(names) + 1, dtype=np.int64)
    np.cumsum([d.shape[0] for d in desc], out=offs[1:])
    np.savez_compressed(os.path.join(OUT, "orb_forest_c1.npz"), names=np.array(names),
                        descriptors=np.concatenate(desc), points=np.concatenate(pts), offsets=offs,
                        img_shape=np.array(imgs[0].shape), cv2_version=cv.__version__)
    print("orb_forest_c1:", [d.shape[0] for d in desc])
    return names, feats, imgs


def bf_knn2(q, t):
    out = np.full((q.shape[0], 4), -1, dtype=np.int32)
    if q.shape[0] == 0 or t.shape[0] == 0:
        return out
    for i, row in enumerate(cv.BFMatcher(cv.NORM_HAMMING).knnMatch(q, t, k=2)):
        for k, m in enumerate(row):
            out[i, 2 * k], out[i, 2 * k + 1] = m.trainIdx, int(m.distance)
    return out


def bf_cross(q, t):
    if q.shape[0] == 0 or t.shape[0] == 0:
        return np.zeros((0, 3), dtype=np.int32)
    ms = cv.BFMatcher(cv.NORM_HAMMING, crossCheck=True).match(cv.UMat(q), cv.UMat(t))
    return np.array([[m.queryIdx, m.trainIdx, int(m.distance)] for m in ms], dtype=np.int32).reshape(-1, 3)


def make_hamming_golden():
    cases = {}
    tie = synth.tie_stress_bank(2, 300, seed=11)
    cases["tie"] = (tie[0], tie[1][:257])
    planted = synth.descriptor_bank(2, 700, seed=12, world=1000)
    cases["planted"] = (planted[0], planted[1][:513])
    rng = np.random.default_rng(13)
    cases["one_train"] = (rng.integers(0, 256, (40, 32), dtype=np.uint8), rng.integers(0, 256, (1, 32), dtype=np.uint8))
    cases["one_query"] = (rng.integers(0, 256, (1, 32), dtype=np.uint8), rng.integers(0, 256, (90, 32), dtype=np.uint8))
    dup = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    cases["identical_sets"] = (dup, dup.copy())
    cases["all_equal_rows"] = (np.zeros((33, 32), np.uint8), np.full((65, 32), 255, np.uint8))
    store = {"cv2_version": cv.__version__, "case_names": np.array(list(cases))}
    for name, (q, t) in cases.items():
        store[f"{name}_q"], store[f"{name}_t"] = q, t
        store[f"{name}_knn2"] = bf_knn2(q, t)
        store[f"{name}_cross"] = bf_cross(q, t)
    np.savez_compressed(os.path.join(OUT, "hamming_golden.npz"), **store)
    print("hamming_golden:", {k: (v[0].shape[0], v[1].shape[0]) for k, v in cases.items()})


def matchesinfo_to_arrays(prefix, mi, store):
    store[f"{prefix}_ids"] = np.array([mi.src_img_idx, mi.dst_img_idx, mi.num_inliers], dtype=np.int64)
    store[f"{prefix}_conf"] = np.array([mi.confidence], dtype=np.float64)
    store[f"{prefix}_H"] = np.zeros((0, 0)) if mi.H is None or np.size(mi.H) == 0 else np.asarray(mi.H, dtype=np.float64)
    store[f"{prefix}_H_is_none"] = np.array([mi.H is None])
    store[f"{prefix}_mask"] = np.asarray(mi.inliers_mask, dtype=np.uint8).ravel()
    store[f"{prefix}_matches"] = np.array([[m.queryIdx, m.trainIdx, m.imgIdx] for m in mi.matches],
                                          dtype=np.int32).reshape(-1, 3)
    store[f"{prefix}_dist"] = np.array([m.distance for m in mi.matches], dtype=np.float32)


def make_reference_matchesinfo(names, feats, imgs, n=6):
    tmp = tempfile.mkdtemp(prefix="ssp_golden_")
    matcher = ref_bf.CustomBruteForceMatcher(
        imgs_features_orb_and_stars=feats[:n], imgs_features_len_orb=[len(f.keypoints) for f in feats[:n]],
        conf_thresh=0.2, img_names=list(names[:n]), work_scale_imgs=imgs[:n],
        pairwise_matches_output_dir=tmp, disable_prompts=True, output_dir_polygon_matching=tmp,
        predefined_overlaps=[], focal_length_pinhole=1135)
    builtins.input = lambda *a: ""  # custom_bruteforce_matcher.py:700 is not gated by disable_prompts
    infos = matcher.match_n_features()
    store = {"n_images": np.array([n]), "conf_thresh": np.array([0.2]), "cv2_version": cv.__version__}
    for k, mi in enumerate(infos):
        matchesinfo_to_arrays(f"m{k}", mi, store)
    np.savez_compressed(os.path.join(OUT, "reference_matchesinfo_c1.npz"), **store)
    kept = sum(1 for mi in infos if mi.src_img_idx >= 0)
    print(f"reference_matchesinfo_c1: {len(infos)} entries, {kept} kept")


def make_polygon_golden():
    w, h, f = 1342, 894, 1135.0
    rng = np.random.default_rng(21)
    # two views 20 degrees apart; ~130 stars per view, sub-pixel float32 coordinates like
    # real contour centroids (image_processors.py:43-219)
    xy_l, b_l, ids = synth.star_fields(2, 130, seed=21, img_w=w, img_h=h, focal=f, integer_pixels=False)
    # views 0 and 1 of camera_poses are 45 degrees apart: re-project view 1 closer by hand
    v = rng.normal(size=(2200, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    bright = rng.uniform(70, 255, size=len(v))
    views = []
    for yaw in (0.0, 18.0):
        cam = v @ synth._rot_yaw_pitch(yaw, -35.0)
        front = cam[:, 2] > 1e-6
        x = f * cam[:, 0] / np.where(front, cam[:, 2], 1) + w // 2
        y = f * cam[:, 1] / np.where(front, cam[:, 2], 1) + h // 2
        vis = front & (x >= 12) & (x < w - 12) & (y >= 12) & (y < h - 12)
        k = np.nonzero(vis)[0]
        pts = np.stack([x[k], y[k]], axis=1).astype(np.float32)
        views.append((pts, bright[k], k))
    store = {"img_shape": np.array([h, w, 3]), "focal": np.array([f]), "cv2_version": cv.__version__}
    kps, imgs = [], []
    for i, (pts, br, sky) in enumerate(views):
        img = synth.render_star_image(pts, br, w, h, radius=2)
        imgs.append(img)
        kps.append(tuple(cv.KeyPoint(float(p[0]), float(p[1]), 4.0) for p in pts))
        store[f"v{i}_xy"] = np.array([kp.pt for kp in kps[-1]], dtype=np.float64)
        store[f"v{i}_sky"] = sky
        store[f"v{i}_brightness"] = np.array([ref_ip.get_star_brightness(kp, img) for kp in kps[-1]], dtype=np.int64)
        store[f"v{i}_gray"] = cv.cvtColor(img, cv.COLOR_BGR2GRAY)  # for the brightness kernel test
    tmp = tempfile.mkdtemp(prefix="ssp_golden_poly_")
    bf = ref_poly.BFPolygonMatcher(tmp, "view0", "view1", 0, 1, imgs[0].shape, 1135)
    pstore, counter = {}, []
    dms = bf.match(kps[0], kps[1], 0, 0, imgs[0], imgs[1], pstore, 0, 1, counter)
    store["dmatches"] = np.array([[m.queryIdx, m.trainIdx] for m in dms], dtype=np.int32).reshape(-1, 2)
    store["dmatch_dist"] = np.array([m.distance for m in dms], dtype=np.float32)
    for i in (0, 1):
        cands = list(pstore[i]["all_possible_polygons_ftX"])  # set iteration order = measure order
        store[f"v{i}_candidates"] = np.array(cands, dtype=np.int32).reshape(-1, 5)
        p = pstore[i]["polygons_from_ftX"]
        store[f"v{i}_norm_idx"] = p["keypoint_indices_rotated_list"].astype(np.int32)
        store[f"v{i}_norm_angle"] = p["keypoint_angles_rotated_list"]
        store[f"v{i}_norm_side"] = p["side_lengths_rotated_list"]
        store[f"v{i}_norm_bright"] = p["keypoint_brightnesses_rotated_list"]
        raw = bf.get_polygons_characteristics(kps[i], imgs[i], "t", 5, pstore[i]["all_possible_polygons_ftX"])
        store[f"v{i}_raw_idx"] = raw["keypoint_indices_rotated_list"].astype(np.int32)
        store[f"v{i}_raw_angle"] = raw["keypoint_angles_rotated_list"]
        store[f"v{i}_raw_side"] = raw["side_lengths_rotated_list"]
        store[f"v{i}_raw_bright"] = raw["keypoint_brightnesses_rotated_list"]
    np.savez_compressed(os.path.join(OUT, "polygon_golden.npz"), **store)
    print("polygon_golden: stars", [len(k) for k in kps], "candidates",
          [len(store[f'v{i}_candidates']) for i in (0, 1)], "polygons",
          [len(store[f'v{i}_raw_idx']) for i in (0, 1)], "dmatches", len(dms))


EISENBERG = "/root/reference/img_eisenberg_18h06m_ISO1600_10s"


def make_eisenberg_c2():
    """BASELINE config 2 on the real frames the reference ships (4 of the 21 eisenberg JPEGs):
    the unchanged star finder (image_processors.compute_star_features, Canny + contours) and the
    unchanged BFPolygonMatcher.match on all 6 pairs, with the polygon cache shared across pairs
    exactly as custom_bruteforce_matcher.py:57-59 shares it."""
    names = sorted(n for n in os.listdir(EISENBERG) if n.endswith(".jpg"))
    tmp = tempfile.mkdtemp(prefix="ssp_golden_c2_")
    store = {"names": np.array(names), "focal": np.array([1135.0]), "cv2_version": cv.__version__}
    kps, imgs, scale = [], [], None
    for i, name in enumerate(names):
        full = cv.imread(os.path.join(EISENBERG, name))
        if scale is None:
            scale = min(1.0, np.sqrt(1.2 * 1e6 / (full.shape[0] * full.shape[1])))
        img = cv.resize(src=full, dsize=None, fx=scale, fy=scale, interpolation=cv.INTER_AREA)
        feats = ref_ip.compute_star_features(img, name, tmp, n_brightest_stars=1000)
        imgs.append(img)
        kps.append(tuple(feats.keypoints))
        store[f"i{i}_xy"] = np.array([kp.pt for kp in kps[-1]], dtype=np.float32).reshape(-1, 2)
        store[f"i{i}_size"] = np.array([kp.size for kp in kps[-1]], dtype=np.float32)
        store[f"i{i}_gray"] = cv.cvtColor(img, cv.COLOR_BGR2GRAY)
    store["img_shape"] = np.array(imgs[0].shape)
    pstore, counter = {}, []
    pair_list = [(a, b) for a in range(len(names)) for b in range(a + 1, len(names))]
    store["pairs"] = np.array(pair_list, dtype=np.int32)
    for p, (a, b) in enumerate(pair_list):
        bf = ref_poly.BFPolygonMatcher(tmp, names[a], names[b], a, b, imgs[a].shape, 1135)
        dms = bf.match(kps[a], kps[b], 0, 0, imgs[a], imgs[b], pstore, a, b, counter)
        store[f"p{p}_dmatches"] = np.array([[m.queryIdx, m.trainIdx] for m in dms], dtype=np.int32).reshape(-1, 2)
        store[f"p{p}_dist"] = np.array([m.distance for m in dms], dtype=np.float32)
        print("c2 pair", (a, b), "dmatches", len(dms))
    for i in range(len(names)):
        q = pstore[i]["polygons_from_ftX"]
        store[f"i{i}_candidates"] = np.array(list(pstore[i]["all_possible_polygons_ftX"]), dtype=np.int32).reshape(-1, 5)
        store[f"i{i}_norm_idx"] = q["keypoint_indices_rotated_list"].astype(np.int32)
        store[f"i{i}_norm_angle"] = q["keypoint_angles_rotated_list"]
        store[f"i{i}_norm_side"] = q["side_lengths_rotated_list"]
        store[f"i{i}_norm_bright"] = q["keypoint_brightnesses_rotated_list"]
    np.savez_compressed(os.path.join(OUT, "eisenberg_c2.npz"), **store)
    print("eisenberg_c2: stars", [len(k) for k in kps], "polygons",
          [len(store[f"i{i}_norm_idx"]) for i in range(len(names))])


def make_reference_matchesinfo_full(names, feats, imgs):
    """All 21 forest images through the unchanged reference matcher, then every entry through the
    reference's real cv2_pickleable container and back (the path the pipeline pickles through,
    stitching_detailed_enhanced.py:1194)."""
    import cv2_pickleable.detail as ref_pickleable  # the unchanged reference package
    tmp = tempfile.mkdtemp(prefix="ssp_golden_full_")
    n = len(names)
    matcher = ref_bf.CustomBruteForceMatcher(
        imgs_features_orb_and_stars=feats, imgs_features_len_orb=[len(f.keypoints) for f in feats],
        conf_thresh=0.2, img_names=list(names), work_scale_imgs=imgs,
        pairwise_matches_output_dir=tmp, disable_prompts=True, output_dir_polygon_matching=tmp,
        predefined_overlaps=[], focal_length_pinhole=1135)
    builtins.input = lambda *a: ""
    infos = matcher.match_n_features()
    store = {"n_images": np.array([n]), "conf_thresh": np.array([0.2]), "cv2_version": cv.__version__}
    for k, mi in enumerate(infos):
        back = ref_pickleable.MatchesInfo(mi).to_cv2()  # raises ValueError if the round trip differs
        matchesinfo_to_arrays(f"m{k}", back, store)
    np.savez_compressed(os.path.join(OUT, "reference_matchesinfo_c1_full.npz"), **store)
    print(f"reference_matchesinfo_c1_full: {len(infos)} entries, "
          f"{sum(1 for mi in infos if mi.src_img_idx >= 0)} kept, all through cv2_pickleable")


# example_06 predefined_overlaps (example_06...py:75-122) restricted to the frames that exist in the
# checkout -- 17-alt2-n (16), 18-alt2-e (17), 19-alt2-s (18), 21-zenith (20) -- and re-indexed 0..3
C2_OVERLAPS = [(0, 1), (1, 2), (0, 3), (1, 3), (2, 3)]


def make_c2_stage():
    names = sorted(n for n in os.listdir(EISENBERG) if n.endswith(".jpg"))
    tmp = tempfile.mkdtemp(prefix="ssp_golden_c2stage_")
    finder = cv.ORB.create(nfeatures=1000, edgeThreshold=10, patchSize=30)  # 'orb-for-starry-sky'
    feats, imgs, n_orb, scale = [], [], [], None
    store = {"names": np.array(names), "focal": np.array([1135.0]), "cv2_version": cv.__version__,
             "overlaps": np.array(C2_OVERLAPS, dtype=np.int32), "conf_thresh": np.array([1.0])}
    for i, name in enumerate(names):
        full = cv.imread(os.path.join(EISENBERG, name))
        if scale is None:
            scale = min(1.0, np.sqrt(1.2 * 1e6 / (full.shape[0] * full.shape[1])))
        img = cv.resize(src=full, dsize=None, fx=scale, fy=scale, interpolation=cv.INTER_AREA)
        # stitching_detailed_enhanced.py:784-834, unchanged order of operations
        orb_ft = cv.detail.computeImageFeatures2(finder, ref_ip.optimize_img_for_feature_detection(img, prompt_dialog=False))
        star_ft = ref_ip.compute_star_features(img, name, tmp, n_brightest_stars=1000)
        merged = cv.detail.computeImageFeatures2(cv.ORB_create(), np.zeros(img.shape, dtype=np.uint8))
        merged.descriptors = cv.vconcat([orb_ft.descriptors, star_ft.descriptors])
        merged.img_idx = -1
        merged.img_size = orb_ft.img_size
        merged.keypoints = orb_ft.keypoints + star_ft.keypoints
        feats.append(merged)
        imgs.append(img)
        n_orb.append(len(orb_ft.keypoints))
        store[f"i{i}_descriptors"] = merged.descriptors.get()
        store[f"i{i}_xy"] = np.array([kp.pt for kp in merged.keypoints], dtype=np.float32).reshape(-1, 2)
        store[f"i{i}_size"] = np.array([kp.size for kp in merged.keypoints], dtype=np.float32)
    store["n_orb"] = np.array(n_orb, dtype=np.int64)
    store["img_shape"] = np.array(imgs[0].shape)
    builtins.input = lambda *a: ""

    # the polygon cache is per matcher (custom_bruteforce_matcher.py:57-59); the three runs hand the
    # same dict on so that the reference's get_star_brightness (148 s per pair) runs once per image.
    # What is cached is a pure function of the image, so each run records what a fresh one would.
    shared_store = {}

    def run(tag, **kw):
        matcher = ref_bf.CustomBruteForceMatcher(
            imgs_features_orb_and_stars=feats, imgs_features_len_orb=n_orb, conf_thresh=1.0,
            img_names=list(names), work_scale_imgs=imgs, pairwise_matches_output_dir=tmp,
            disable_prompts=True, output_dir_polygon_matching=tmp, focal_length_pinhole=1135, **kw)
        matcher.polygon_data_store = shared_store
        raw = {}
        inner = matcher.match_2_features

        def recording(src, dst):
            out = inner(src, dst)
            for side, mi in enumerate(out):
                matchesinfo_to_arrays(f"{tag}_raw_{src}_{dst}_{side}", mi, store)
            raw[(src, dst)] = out
            return out
        matcher.match_2_features = recording
        infos = matcher.match_n_features()
        for k, mi in enumerate(infos):
            matchesinfo_to_arrays(f"{tag}_m{k}", mi, store)
        kinds = sorted(matcher.kind_of_pairwise_matches.items())
        store[f"{tag}_kinds"] = np.array([[s, d, 1 if v == "stars" else 0] for (s, d), v in kinds], dtype=np.int32)
        store[f"{tag}_counter"] = np.array(matcher.fts_calculated_counter, dtype=np.int32)
        print(f"c2 stage [{tag}]: kinds {dict(kinds)}; kept",
              [(mi.src_img_idx, mi.dst_img_idx, mi.confidence, mi.num_inliers) for mi in infos if mi.src_img_idx >= 0])

    run("overlaps", predefined_overlaps=C2_OVERLAPS)
    run("enforced", predefined_overlaps=C2_OVERLAPS, enforced_matches=[(names[0], names[2])],
        enforced_no_matches=[(names[1], names[3])])
    run("allstars", predefined_overlaps=[], enforce_star_polygon_matcher=True)
    np.savez_compressed(os.path.join(OUT, "eisenberg_c2_stage.npz"), **store)
    print("eisenberg_c2_stage: n_orb", n_orb, "stars", [len(f.keypoints) - k for f, k in zip(feats, n_orb)])



if __name__ == "__main__":
    which = sys.argv[1:] or ["orb", "hamming", "matchesinfo", "matchesinfo_full", "polygon", "c2", "c2stage"]
    names = feats = imgs = None
    if "orb" in which or "matchesinfo" in which or "matchesinfo_full" in which:
        names, feats, imgs = make_orb_fixture()
    if "hamming" in which:
        make_hamming_golden()
    if "matchesinfo" in which:
        make_reference_matchesinfo(names, feats, imgs)
    if "matchesinfo_full" in which:
        make_reference_matchesinfo_full(names, feats, imgs)
    if "polygon" in which:
        make_polygon_golden()
    if "c2" in which:
        make_eisenberg_c2()
    if "c2stage" in which:
        make_c2_stage()
