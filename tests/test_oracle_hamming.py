"""Pins oracle/hamming_oracle.c against cv2.BFMatcher: the committed golden vectors (generated
from cv2.BFMatcher by tests/golden/make_golden.py) and the cv2 installed on this machine."""
import cv2
import numpy as np
import pytest

from oracle import hamming_oracle as ho
import ssp_b200.synthetic as synth


def cv_knn2(q, t):
    out = np.full((q.shape[0], 4), -1, dtype=np.int32)
    for i, row in enumerate(cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)):
        for k, m in enumerate(row):
            out[i, 2 * k], out[i, 2 * k + 1] = m.trainIdx, int(m.distance)
    return out


def cv_cross(q, t):
    ms = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(cv2.UMat(q), cv2.UMat(t))
    return np.array([[m.queryIdx, m.trainIdx, int(m.distance)] for m in ms], dtype=np.int32).reshape(-1, 3)


def test_oracle_matches_golden_vectors(golden):
    g = golden("hamming_golden.npz")
    for name in g["case_names"]:
        q, t = g[f"{name}_q"], g[f"{name}_t"]
        np.testing.assert_array_equal(ho.knn2(q, t), g[f"{name}_knn2"], err_msg=str(name))
        mq, mt, md = ho.crosscheck(q, t)
        np.testing.assert_array_equal(np.stack([mq, mt, md], axis=1), g[f"{name}_cross"], err_msg=str(name))


@pytest.mark.parametrize("seed", range(6))
def test_oracle_matches_installed_cv2_on_tie_heavy_inputs(seed):
    rng = np.random.default_rng(100 + seed)
    nq, nt = int(rng.integers(1, 400)), int(rng.integers(2, 400))
    bank = synth.tie_stress_bank(2, max(nq, nt), seed=seed)
    q, t = bank[0][:nq], bank[1][:nt]
    np.testing.assert_array_equal(ho.knn2(q, t), cv_knn2(q, t))
    mq, mt, md = ho.crosscheck(q, t)
    np.testing.assert_array_equal(np.stack([mq, mt, md], axis=1), cv_cross(q, t))


def test_ratio_form_is_the_float_test():
    # m.distance < 0.75 * n.distance evaluated like the reference (float32 distances, python float)
    d1, d2 = np.meshgrid(np.arange(257), np.arange(257), indexing="ij")
    ref = d1.astype(np.float32).astype(np.float64) < 0.75 * d2.astype(np.float32).astype(np.float64)
    np.testing.assert_array_equal(4 * d1 < 3 * d2, ref)


def test_ratio_survivors_on_planted_matches():
    bank = synth.descriptor_bank(2, 600, seed=5, world=800)
    b2 = ho.knn2(bank[0], bank[1])
    mq, mt, md = ho.ratio_survivors(b2)
    keep = [i for i in range(len(b2)) if b2[i, 2] >= 0 and float(b2[i, 1]) < 0.75 * float(b2[i, 3])]
    np.testing.assert_array_equal(mq, np.array(keep, dtype=np.int32))
    assert 50 < len(mq) < 600  # the generator plants real matches
    np.testing.assert_array_equal(mt, b2[keep, 0])


def test_edge_cases():
    q = np.zeros((3, 32), np.uint8)
    assert ho.knn2(q, np.zeros((0, 32), np.uint8)).tolist() == [[-1, -1, -1, -1]] * 3
    one = ho.knn2(q, np.full((1, 32), 255, np.uint8))
    assert one.tolist() == [[0, 256, -1, -1]] * 3
    assert len(ho.crosscheck(q, np.zeros((0, 32), np.uint8))[0]) == 0
    assert len(ho.crosscheck(np.zeros((0, 32), np.uint8), q)[0]) == 0
