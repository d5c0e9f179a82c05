"""The CustomBruteForceMatcher mirror on the engine against (a) every MatchesInfo field the
UNCHANGED reference produced (reference_matchesinfo_c1.npz, recorded in the build container)
and (b) the same host logic fed by cv.BFMatcher itself, over all 210 pairs of BASELINE config 1."""
import cv2
import numpy as np
import pytest

import ssp_b200
from ssp_b200 import synthetic as synth

pytestmark = pytest.mark.gpu


def features_from_fixture(g, n=None):
    offs = g["offsets"]
    n = len(offs) - 1 if n is None else n
    h, w = int(g["img_shape"][0]), int(g["img_shape"][1])
    orb = cv2.ORB_create()
    feats = []
    for i in range(n):
        ft = cv2.detail.computeImageFeatures2(orb, np.zeros((h, w, 3), np.uint8))
        ft.descriptors = cv2.UMat(np.ascontiguousarray(g["descriptors"][offs[i]:offs[i + 1]]))
        ft.keypoints = tuple(cv2.KeyPoint(float(p[0]), float(p[1]), 30.0) for p in g["points"][offs[i]:offs[i + 1]])
        ft.img_size = (w, h)
        ft.img_idx = -1
        feats.append(ft)
    imgs = [np.zeros((h, w, 3), np.uint8) for _ in range(n)]
    return feats, imgs, [str(s) for s in g["names"][:n]]


def make_matcher(engine, feats, imgs, names, tmp, **kw):
    return ssp_b200.CustomBruteForceMatcher(
        imgs_features_orb_and_stars=feats, imgs_features_len_orb=[len(f.keypoints) for f in feats],
        conf_thresh=0.2, img_names=names, work_scale_imgs=imgs, pairwise_matches_output_dir=str(tmp),
        disable_prompts=True, output_dir_polygon_matching=str(tmp), predefined_overlaps=[],
        focal_length_pinhole=1135, engine=engine, write_debug_images=False, verbose=False, **kw)


def info_fields(mi):
    H = None if mi.H is None or np.size(mi.H) == 0 else np.asarray(mi.H)
    return (mi.src_img_idx, mi.dst_img_idx, mi.num_inliers, float(mi.confidence), H,
            np.asarray(mi.inliers_mask, dtype=np.uint8).ravel(),
            [(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in mi.matches])


def test_matchesinfo_identical_to_reference_recording(engine, golden, tmp_path):
    g = golden("orb_forest_c1.npz")
    ref = golden("reference_matchesinfo_c1.npz")
    n = int(ref["n_images"][0])
    feats, imgs, names = features_from_fixture(g, n)
    infos = make_matcher(engine, feats, imgs, names, tmp_path).match_n_features()
    assert isinstance(infos, tuple) and len(infos) == n * n
    kept = 0
    for k, mi in enumerate(infos):
        src, dst, inl, conf, H, mask, matches = info_fields(mi)
        assert [src, dst, inl] == ref[f"m{k}_ids"].tolist(), k
        assert conf == float(ref[f"m{k}_conf"][0]), k
        np.testing.assert_array_equal(mask, ref[f"m{k}_mask"], err_msg=str(k))
        assert [list(m[:3]) for m in matches] == ref[f"m{k}_matches"].tolist(), k
        np.testing.assert_array_equal(np.array([m[3] for m in matches], dtype=np.float32), ref[f"m{k}_dist"])
        if H is None:
            assert ref[f"m{k}_H"].size == 0
        else:
            np.testing.assert_array_equal(H, ref[f"m{k}_H"])
        kept += src >= 0
    assert kept == 8


def test_quirk_q1_mirror_keeps_unclamped_confidence(engine, golden, tmp_path):
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g, 2)
    fwd, back = make_matcher(engine, feats, imgs, names, tmp_path).match_2_features(0, 1)
    assert fwd.num_inliers > 50 and fwd.confidence == 0.4 and back.confidence > 0.4
    assert back.src_img_idx == 1 and back.dst_img_idx == 0
    assert [(m.queryIdx, m.trainIdx) for m in back.matches] == [(m.trainIdx, m.queryIdx) for m in fwd.matches]
    np.testing.assert_allclose(back.H @ fwd.H, np.eye(3), atol=1e-9)


class _CvMatcherEngineStandIn:
    """Feeds the mirror's host logic from cv.BFMatcher itself (the reference's own call)."""

    def __init__(self):
        self.sets = None

    def upload_descriptor_bank(self, rows):
        self.sets = rows

    def hamming_match_pairs(self, pairs, mode="crosscheck", ratio=0.75, variant="auto"):
        from ssp_b200.engine import HammingResult
        qs, ts, ds, counts = [], [], [], []
        for a, b in pairs:
            ms = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(cv2.UMat(self.sets[a]), cv2.UMat(self.sets[b]))
            qs += [m.queryIdx for m in ms]; ts += [m.trainIdx for m in ms]; ds += [int(m.distance) for m in ms]
            counts.append(len(ms))
        o = np.zeros(len(pairs) + 1, dtype=np.int64)
        np.cumsum(counts, out=o[1:])
        return HammingResult(o, np.array(qs, np.int32), np.array(ts, np.int32), np.array(ds, np.int32))


def test_all_210_pairs_engine_equals_cv_bfmatcher_host_path(engine, golden, tmp_path):
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g)
    ours = make_matcher(engine, feats, imgs, names, tmp_path).match_n_features()
    theirs = make_matcher(_CvMatcherEngineStandIn(), feats, imgs, names, tmp_path).match_n_features()
    assert len(ours) == len(theirs) == 21 * 21
    kept = 0
    for a, b in zip(ours, theirs):
        fa, fb = info_fields(a), info_fields(b)
        assert fa[:4] == fb[:4]
        assert (fa[4] is None) == (fb[4] is None)
        if fa[4] is not None:
            np.testing.assert_array_equal(fa[4], fb[4])
        np.testing.assert_array_equal(fa[5], fb[5])
        assert fa[6] == fb[6]
        kept += a.src_img_idx >= 0
    assert kept >= 60  # the author's run kept 47 pairs (94 directed entries) on opencv 4.6.0


def test_pruning_wrapper_returns_list_that_passes_bundle_adjustment(engine, golden, tmp_path):
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g, 5)
    out = make_matcher(engine, feats, imgs, names, tmp_path).get_valid_pairwise_matches_that_will_pass_bundle_adjuster_wo_crashes()
    assert isinstance(out, list) and len(out) == 25
    assert out[0].src_img_idx == -1 and out[1].src_img_idx == 0 and out[1].dst_img_idx == 1
    cv2.detail.leaveBiggestComponent(feats, out, 0.2)  # the consumer of the return value


def test_cv2_pickleable_style_rebuild_of_engine_matchesinfo(engine, golden, tmp_path):
    """The construction path north_star names: field-by-field rebuild as in
    cv2_pickleable/detail.py:172-211 must accept what the engine path produced."""
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g, 2)
    fwd, _ = make_matcher(engine, feats, imgs, names, tmp_path).match_2_features(0, 1)
    clone = cv2.detail.MatchesInfo()
    clone.H = fwd.H
    clone.confidence = fwd.confidence
    clone.dst_img_idx = fwd.dst_img_idx
    clone.inliers_mask = fwd.inliers_mask
    clone.matches = tuple(cv2.DMatch(_queryIdx=m.queryIdx, _trainIdx=m.trainIdx, _distance=m.distance,
                                     _imgIdx=m.imgIdx) for m in fwd.matches)
    clone.num_inliers = fwd.num_inliers
    clone.src_img_idx = fwd.src_img_idx
    assert np.array_equal(clone.H, fwd.H) and np.array_equal(clone.inliers_mask, fwd.inliers_mask)
    assert all(a.distance == b.distance and a.queryIdx == b.queryIdx and a.trainIdx == b.trainIdx
               for a, b in zip(clone.matches, fwd.matches))
