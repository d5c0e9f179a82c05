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


def assert_info_equals_recording(mi, ref, prefix):
    """Every MatchesInfo field against a recording of the unchanged reference, `H` NOT normalised:
    where the reference left H = None a real entry must read None too; the (-1, -1) dummy is the
    one documented deviation (custom_bruteforce_matcher.py mirror, match_n_features): it reads back
    as an EMPTY (0, 0) float64 array because cv2 4.13's estimator segfaults on a None there."""
    assert [mi.src_img_idx, mi.dst_img_idx, mi.num_inliers] == ref[f"{prefix}_ids"].tolist(), prefix
    assert float(mi.confidence) == float(ref[f"{prefix}_conf"][0]), prefix
    np.testing.assert_array_equal(np.asarray(mi.inliers_mask, dtype=np.uint8).ravel(), ref[f"{prefix}_mask"],
                                  err_msg=prefix)
    assert [[m.queryIdx, m.trainIdx, m.imgIdx] for m in mi.matches] == ref[f"{prefix}_matches"].tolist(), prefix
    np.testing.assert_array_equal(np.array([m.distance for m in mi.matches], dtype=np.float32),
                                  ref[f"{prefix}_dist"], err_msg=prefix)
    if bool(ref[f"{prefix}_H_is_none"][0]):
        if mi.src_img_idx == -1:
            assert isinstance(mi.H, np.ndarray) and mi.H.shape == (0, 0) and mi.H.dtype == np.float64, prefix
        else:
            assert mi.H is None, prefix
    else:
        np.testing.assert_array_equal(np.asarray(mi.H), ref[f"{prefix}_H"], err_msg=prefix)


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


def test_all_441_entries_equal_the_reference_recording_that_went_through_cv2_pickleable(engine, golden, tmp_path):
    """BASELINE config 1 complete: every field of all 441 MatchesInfo equals what the UNCHANGED
    reference matcher returned for the 21 forest frames after each entry went through the
    reference's real cv2_pickleable.detail.MatchesInfo(x).to_cv2() (cv2_pickleable/detail.py:154-211;
    recorded in the build container by tests/golden/make_golden.py matchesinfo_full)."""
    g = golden("orb_forest_c1.npz")
    ref = golden("reference_matchesinfo_c1_full.npz")
    feats, imgs, names = features_from_fixture(g)
    infos = make_matcher(engine, feats, imgs, names, tmp_path).match_n_features()
    assert len(infos) == 441 == int(ref["n_images"][0]) ** 2
    for k, mi in enumerate(infos):
        assert_info_equals_recording(mi, ref, f"m{k}")
    assert sum(mi.src_img_idx >= 0 for mi in infos) == 90


# ---- BASELINE config 2 as a whole stage: ORB -> RANSAC -> fallback decision -> star MatchesInfo ----

def c2_stage_inputs(golden):
    st = golden("eisenberg_c2_stage.npz")
    frames = golden("eisenberg_c2.npz")
    n = len(st["names"])
    h, w = int(st["img_shape"][0]), int(st["img_shape"][1])
    orb = cv2.ORB_create()
    feats, imgs = [], []
    for i in range(n):
        ft = cv2.detail.computeImageFeatures2(orb, np.zeros((h, w, 3), np.uint8))
        ft.descriptors = cv2.UMat(np.ascontiguousarray(st[f"i{i}_descriptors"]))
        ft.keypoints = tuple(cv2.KeyPoint(float(p[0]), float(p[1]), float(sz))
                             for p, sz in zip(st[f"i{i}_xy"], st[f"i{i}_size"]))
        ft.img_size = (w, h)
        ft.img_idx = -1
        feats.append(ft)
        # get_star_brightness only looks at BGR2GRAY of the work image; a grey BGR image converts back exactly
        imgs.append(cv2.cvtColor(frames[f"i{i}_gray"], cv2.COLOR_GRAY2BGR))
    return st, feats, imgs, [str(x) for x in st["names"]], [int(x) for x in st["n_orb"]]


def run_c2_stage(engine, st, feats, imgs, names, n_orb, tmp, tag, **kw):
    m = ssp_b200.CustomBruteForceMatcher(
        imgs_features_orb_and_stars=feats, imgs_features_len_orb=n_orb, conf_thresh=float(st["conf_thresh"][0]),
        img_names=names, work_scale_imgs=imgs, pairwise_matches_output_dir=str(tmp), disable_prompts=True,
        output_dir_polygon_matching=str(tmp), focal_length_pinhole=1135, engine=engine,
        write_debug_images=False, verbose=False, **kw)
    raw = {}
    inner = m.match_2_features

    def recording(src, dst):
        raw[(src, dst)] = inner(src, dst)
        return raw[(src, dst)]
    m.match_2_features = recording
    infos = m.match_n_features()
    n = len(names)
    assert isinstance(infos, tuple) and len(infos) == n * n
    for (src, dst), pair in raw.items():
        for side in (0, 1):
            assert_info_equals_recording(pair[side], st, f"{tag}_raw_{src}_{dst}_{side}")
    for k, mi in enumerate(infos):
        assert_info_equals_recording(mi, st, f"{tag}_m{k}")
    kinds = sorted(m.kind_of_pairwise_matches.items())
    assert [[s, d, 1 if v == "stars" else 0] for (s, d), v in kinds] == st[f"{tag}_kinds"].tolist()
    if tag == "overlaps":
        # only the first recording started from an empty polygon cache (make_golden.py hands the
        # dict on to the other two so that the reference's brightness loop runs once per image)
        assert m.fts_calculated_counter == st[f"{tag}_counter"].tolist()
    return m, raw, infos


def test_c2_stage_star_fallback_equals_reference(engine, golden, tmp_path):
    """example_06 on the frames the reference ships: pairs whose ORB result is weak AND that are in
    predefined_overlaps fall through to the star-polygon matcher (custom_bruteforce_matcher.py:360-552);
    quirk Q4 -- confidence = int(3 * conf_thresh) applied before mirroring -- and the STARS mirror are
    compared field by field with the unchanged reference."""
    st, feats, imgs, names, n_orb = c2_stage_inputs(golden)
    overlaps = [tuple(int(v) for v in p) for p in st["overlaps"]]
    m, raw, infos = run_c2_stage(engine, st, feats, imgs, names, n_orb, tmp_path, "overlaps",
                                 predefined_overlaps=overlaps)
    stars = [k for k, v in m.kind_of_pairwise_matches.items() if v == "stars"]
    assert stars == [(1, 2), (2, 3)]
    fwd, back = raw[(1, 2)]
    assert fwd.confidence == 3 and isinstance(fwd.confidence, float) and back.confidence == 3  # Q4, mirrored after the clamp
    assert all(mm.queryIdx >= n_orb[1] and mm.trainIdx >= n_orb[2] for mm in fwd.matches)      # star keypoints only
    assert [(mm.queryIdx, mm.trainIdx) for mm in back.matches] == [(mm.trainIdx, mm.queryIdx) for mm in fwd.matches]


def test_c2_stage_enforced_matches_and_no_matches_equal_reference(engine, golden, tmp_path):
    st, feats, imgs, names, n_orb = c2_stage_inputs(golden)
    overlaps = [tuple(int(v) for v in p) for p in st["overlaps"]]
    m, raw, infos = run_c2_stage(engine, st, feats, imgs, names, n_orb, tmp_path, "enforced",
                                 predefined_overlaps=overlaps, enforced_matches=[(names[0], names[2])],
                                 enforced_no_matches=[(names[1], names[3])])
    assert raw[(0, 2)][0].confidence == 10 and raw[(0, 2)][1].confidence == 10   # :312-321
    n = len(names)
    assert infos[1 * n + 3].src_img_idx == -1 and infos[3 * n + 1].src_img_idx == -1  # enforced no-match -> dummy
    assert infos[0 * n + 2].src_img_idx == 0 and infos[0 * n + 2].confidence == 10


def test_c2_stage_enforce_star_polygon_matcher_equals_reference(engine, golden, tmp_path):
    st, feats, imgs, names, n_orb = c2_stage_inputs(golden)
    m, raw, infos = run_c2_stage(engine, st, feats, imgs, names, n_orb, tmp_path, "allstars",
                                 predefined_overlaps=[], enforce_star_polygon_matcher=True)
    assert set(m.kind_of_pairwise_matches.values()) == {"stars"} and len(raw) == 6


# ---- one engine context shared by several host objects ------------------------------

def test_bfmatcher_dropin_between_lazy_matcher_calls_does_not_swap_the_bank(engine, golden, tmp_path):
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g, 3)
    want = make_matcher(engine, feats, imgs, names, tmp_path).match_2_features(1, 2)
    m = make_matcher(engine, feats, imgs, names, tmp_path)
    first = m.match_2_features(0, 1)      # uploads the matcher's 3-image bank, lazily
    rows = synth.descriptor_bank(2, 700, seed=5, world=900)
    ssp_b200.BFMatcher(cv2.NORM_HAMMING, crossCheck=True, engine=engine).match(rows[0], rows[1])  # replaces it
    got = m.match_2_features(1, 2)        # must notice and upload again
    assert info_fields(got[0])[:4] == info_fields(want[0])[:4]
    assert info_fields(got[0])[6] == info_fields(want[0])[6] and len(first[0].matches) > 0


def test_two_matchers_with_the_same_image_count_on_one_engine(engine, golden, tmp_path):
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g, 6)
    a = make_matcher(engine, feats[:3], imgs[:3], names[:3], tmp_path)
    b = make_matcher(engine, feats[3:], imgs[3:], names[3:], tmp_path)
    a01 = a.match_2_features(0, 1)
    b01 = b.match_2_features(0, 1)
    a12 = a.match_2_features(1, 2)        # a's bank was replaced by b's in between
    solo = make_matcher(engine, feats[:3], imgs[:3], names[:3], tmp_path).match_2_features(1, 2)
    assert info_fields(a12[0])[6] == info_fields(solo[0])[6]
    assert info_fields(a01[0])[6] != info_fields(b01[0])[6]


def test_default_engine_is_one_context_per_matcher(golden, tmp_path):
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g, 2)
    a = make_matcher(None, feats, imgs, names, tmp_path)
    b = make_matcher(None, feats, imgs, names, tmp_path)
    assert a.engine is not b.engine
    assert len(a.match_2_features(0, 1)[0].matches) == len(b.match_2_features(0, 1)[0].matches) > 0
    a.engine.close()
    b.engine.close()


def test_two_polygon_stores_on_one_engine_do_not_swap_banks(engine, golden, tmp_path):
    """Two polygon_data_store dicts (two matcher instances) that both hold an image 0 share one
    context: the second store's measure of ITS image 0 must not be matched by the first."""
    c2 = golden("eisenberg_c2.npz")
    kps, imgs = [], []
    for i in range(3):
        kps.append(tuple(cv2.KeyPoint(float(p[0]), float(p[1]), float(sz))
                         for p, sz in zip(c2[f"i{i}_xy"], c2[f"i{i}_size"])))
        imgs.append(cv2.cvtColor(c2[f"i{i}_gray"], cv2.COLOR_GRAY2BGR))

    def match(store, counter, a, b, ia, ib):
        bf = ssp_b200.BFPolygonMatcher(str(tmp_path), f"f{a}", f"f{b}", ia, ib, imgs[a].shape, 1135,
                                       engine=engine, write_debug_images=False)
        return [(mm.queryIdx, mm.trainIdx) for mm in bf.match(kps[a], kps[b], 0, 0, imgs[a], imgs[b], store, ia, ib, counter)]
    s1, c1, s2, c2_ = {}, [], {}, []
    want = match(s1, c1, 0, 1, 0, 1)
    match(s2, c2_, 2, 1, 0, 1)            # another store: ITS image 0 is frame 2
    again = match(s1, c1, 0, 1, 0, 1)     # cached in s1, but the engine holds s2's banks now
    assert again == want and c1 == [0, 1]
