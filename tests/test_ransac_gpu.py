"""Row N3: the batched device RANSAC (opt-in stand-in for cv.findHomography(..., cv.RANSAC, 5),
custom_bruteforce_matcher.py:215-246).  Not bit-compatible with OpenCV by nature, so the checks are:
exact recovery of planted homographies, and inlier-count / kept-pair agreement with cv.findHomography
on all 210 pairs of BASELINE config 1."""
import cv2
import numpy as np
import pytest

import ssp_b200
from test_matcher_gpu import features_from_fixture, make_matcher

pytestmark = pytest.mark.gpu


def _planted_pair(rng, n, n_out, H, noise=0.0):
    """Two descriptor sets whose cross-check matches are exactly i <-> i, keypoints related by H for
    the first n - n_out rows and random for the rest."""
    desc = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    src = rng.uniform(20, 1300, size=(n, 2)).astype(np.float32)
    src[:, 1] = rng.uniform(20, 870, size=n).astype(np.float32)
    p = np.concatenate([src.astype(np.float64), np.ones((n, 1))], axis=1) @ H.T
    dst = (p[:, :2] / p[:, 2:3] + rng.normal(0, noise, size=(n, 2))).astype(np.float32)
    dst[n - n_out:] = rng.uniform(20, 1300, size=(n_out, 2)).astype(np.float32)
    return desc, src, dst


def test_planted_homographies_are_recovered_exactly(engine):
    rng = np.random.default_rng(5)
    Hs = [np.array([[1.02, 0.03, 40.0], [-0.02, 0.99, -25.0], [1e-5, -2e-5, 1.0]]),
          np.array([[0.9, -0.2, 300.0], [0.2, 0.9, 10.0], [-3e-5, 1e-5, 1.0]]),
          np.eye(3)]
    sets, kps, truth = [], [], []
    for H, n, n_out in zip(Hs, (400, 90, 12), (160, 30, 3)):
        desc, src, dst = _planted_pair(rng, n, n_out, H)
        flip = desc.copy()
        flip[:, 0] ^= 1                      # distance 1 to its partner, ~128 to everything else
        sets += [desc, flip]
        kps += [src, dst]
        truth.append((H, n, n_out))
    sets.append(rng.integers(0, 256, size=(3, 32), dtype=np.uint8))   # a pair with < 6 matches
    sets.append(rng.integers(0, 256, size=(4, 32), dtype=np.uint8))
    kps += [rng.uniform(0, 100, (3, 2)).astype(np.float32), rng.uniform(0, 100, (4, 2)).astype(np.float32)]
    pairs = [(0, 1), (2, 3), (4, 5), (6, 7)]
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs(pairs, mode="crosscheck")
    H, mask, inl = engine.ransac_homography_pairs(kps, pairs, int(res.offsets[-1]), reproj_threshold=5.0,
                                                  n_hypotheses=2048, seed=0)
    for k, (Ht, n, n_out) in enumerate(truth):
        q, t, d = res.pair(k)
        assert q.tolist() == list(range(n)) and t.tolist() == list(range(n))
        m = mask[res.offsets[k]:res.offsets[k + 1]]
        assert m[:n - n_out].all() and inl[k] == m.sum() and m.sum() <= n - n_out + 2   # a stray outlier may fit
        np.testing.assert_allclose(H[k], Ht / Ht[2, 2], rtol=0, atol=2e-3 * np.abs(Ht).max())
        # float32 keypoints bound the accuracy: corner reprojection within a hundredth of a pixel
        corners = np.array([[0, 0, 1], [1342, 0, 1], [0, 894, 1], [1342, 894, 1]], dtype=np.float64)
        a, b = corners @ H[k].T, corners @ Ht.T
        assert np.abs(a[:, :2] / a[:, 2:] - b[:, :2] / b[:, 2:]).max() < 0.05
    assert inl[3] == -1 and np.isnan(H[3]).all()     # < 6 matches: no model (custom_bruteforce_matcher.py:215)
    # deterministic: same seed, same answer
    H2, mask2, inl2 = engine.ransac_homography_pairs(kps, pairs, int(res.offsets[-1]), seed=0)
    np.testing.assert_array_equal(np.nan_to_num(H), np.nan_to_num(H2))
    np.testing.assert_array_equal(mask, mask2)


def test_device_ransac_agrees_with_cv_findhomography_on_config_1(engine, golden, tmp_path):
    """All 210 pairs of example_01b: the whole stage with ransac='device' against the reference's
    cv.findHomography path on the same matches."""
    g = golden("orb_forest_c1.npz")
    feats, imgs, names = features_from_fixture(g)
    ours = make_matcher(engine, feats, imgs, names, tmp_path, ransac="device").match_n_features()
    theirs = make_matcher(engine, feats, imgs, names, tmp_path).match_n_features()
    kept_o = {(m.src_img_idx, m.dst_img_idx) for m in ours if m.src_img_idx >= 0}
    kept_t = {(m.src_img_idx, m.dst_img_idx) for m in theirs if m.src_img_idx >= 0}
    # pairs kept by one side only sit at the acceptance border (conf > 0.2 and inliers > 5)
    assert len(kept_o ^ kept_t) <= 0.1 * len(kept_t), (sorted(kept_o - kept_t), sorted(kept_t - kept_o))
    rel, reproj, n_ours, n_theirs = [], [], 0, 0
    for a, b in zip(ours, theirs):
        if a.src_img_idx < 0 or b.src_img_idx < 0 or a.src_img_idx > a.dst_img_idx:
            continue
        assert [(m.queryIdx, m.trainIdx) for m in a.matches] == [(m.queryIdx, m.trainIdx) for m in b.matches]
        rel.append(abs(a.num_inliers - b.num_inliers) / max(b.num_inliers, 1))
        n_ours += a.num_inliers
        n_theirs += b.num_inliers
        # OpenCV's inliers under OUR homography
        mask = np.asarray(b.inliers_mask).ravel().astype(bool)
        src = np.array([feats[a.src_img_idx].keypoints[m.queryIdx].pt for m in a.matches])[mask]
        dst = np.array([feats[a.dst_img_idx].keypoints[m.trainIdx].pt for m in a.matches])[mask]
        p = np.concatenate([src, np.ones((len(src), 1))], axis=1) @ np.asarray(a.H).T
        reproj.append(np.median(np.linalg.norm(p[:, :2] / p[:, 2:3] - dst, axis=1)))
    assert len(rel) >= 30
    # the refits that are accepted while they gain inliers find at least what OpenCV's best sample +
    # LM polish finds, pair by pair within a tenth
    assert n_ours >= 0.97 * n_theirs, (n_ours, n_theirs)
    assert np.median(rel) <= 0.10 and np.mean(rel) <= 0.15, (np.median(rel), np.mean(rel), n_ours, n_theirs)
    assert np.median(reproj) < 2.0, np.median(reproj)
