"""Pins oracle/polygon_oracle.py against the reference: the known-answer vector of SURVEY.md
§8(c) and the fixtures recorded from the UNCHANGED reference BFPolygonMatcher
(tests/golden/make_golden.py -> polygon_golden.npz)."""
import numpy as np
import pytest

from oracle.polygon_oracle import PolygonOracle, star_brightness

KAT_PTS = [(739.2000122070312, 174.0), (722.0, 247.0), (426.0, 263.0),
           (510.0000305175781, 104.4000015258789), (663.8400268554688, 28.80000114440918)]


def test_known_answer_vector():
    o = PolygonOracle((894, 1342, 3), 1135)
    assert o.angle_distance(KAT_PTS[0], KAT_PTS[1]) == 0.06314409535278809
    assert o.sphere_angle(KAT_PTS[4], KAT_PTS[0], KAT_PTS[1]) == 2.417040566696135
    m = o.measure(KAT_PTS)
    assert m["order"] == [2, 1, 0, 4, 3]
    np.testing.assert_array_equal(m["angles"], [1.0070090368863567, 1.8512833206657646, 2.417040566696135,
                                                1.6426682393834149, 2.538864218589824])
    np.testing.assert_array_equal(m["sides"], [0.25452558770269285, 0.06314409535278809, 0.13342053518746685,
                                               0.14218445156383053, 0.1504438827434782])
    assert sum(m["angles"]) > 3 * np.pi  # positive spherical excess


@pytest.fixture(scope="module")
def g(golden):
    return golden("polygon_golden.npz")


@pytest.mark.parametrize("view", [0, 1])
def test_characteristics_match_reference(g, view):
    o = PolygonOracle(tuple(g["img_shape"]), float(g["focal"][0]))
    cands = g[f"v{view}_candidates"][:2500]
    got = o.characteristics(g[f"v{view}_xy"], g[f"v{view}_brightness"], cands)
    n = len(got["star_idx"])
    assert n > 500
    # reference rows are in candidate (set iteration) order too
    np.testing.assert_array_equal(got["star_idx"], g[f"v{view}_raw_idx"][:n])
    np.testing.assert_allclose(got["angle"], g[f"v{view}_raw_angle"][:n], rtol=0, atol=1e-12)
    np.testing.assert_allclose(got["side"], g[f"v{view}_raw_side"][:n], rtol=0, atol=1e-12)
    np.testing.assert_array_equal(got["brightness"], g[f"v{view}_raw_bright"][:n])
    np.testing.assert_allclose(got["feat"][:, 0:5], g[f"v{view}_norm_angle"][:n], rtol=0, atol=1e-12)
    np.testing.assert_allclose(got["feat"][:, 5:10], g[f"v{view}_norm_side"][:n], rtol=0, atol=1e-12)
    np.testing.assert_array_equal(got["feat"][:, 10:15], g[f"v{view}_norm_bright"][:n])


def test_compare_select_vote_reproduce_reference_dmatches(g):
    feats, idxs = [], []
    for v in (0, 1):
        feats.append(np.concatenate([g[f"v{v}_norm_angle"], g[f"v{v}_norm_side"], g[f"v{v}_norm_bright"]], axis=1))
        idxs.append(g[f"v{v}_norm_idx"])
    m = PolygonOracle.compare(feats[0], feats[1])
    rows, cols, d = PolygonOracle.select(m, 0.1566)
    assert len(rows) == 150  # "Remaining 150 polygon matches" printed by the reference run
    votes = PolygonOracle.vote(rows, cols, d, idxs[0], idxs[1])
    np.testing.assert_array_equal(np.array([(a, b) for a, b, _ in votes], dtype=np.int32), g["dmatches"])
    np.testing.assert_array_equal(np.array([w for _, _, w in votes], dtype=np.float32), g["dmatch_dist"])


def test_numpy_reduction_order_is_sequential():
    # kernel 3 reproduces numpy's float64 rounding; this pins the order numpy uses here
    rng = np.random.default_rng(0)
    a = rng.random((4000, 5)) * rng.choice([1e-8, 1e-3, 1.0, 1e3], size=(4000, 5))
    s = np.square(a)
    seq = (((s[:, 0] + s[:, 1]) + s[:, 2]) + s[:, 3]) + s[:, 4]
    np.testing.assert_array_equal(np.sum(s, axis=1), seq)


def test_sequential_dist2_equals_reference_expression(g):
    f1 = np.concatenate([g["v0_norm_angle"], g["v0_norm_side"], g["v0_norm_bright"]], axis=1)[:300]
    f2 = np.concatenate([g["v1_norm_angle"], g["v1_norm_side"], g["v1_norm_bright"]], axis=1)[:200]
    np.testing.assert_array_equal(np.sqrt(PolygonOracle._dist2_seq(f1, f2)), PolygonOracle.compare(f1, f2))


def test_hull_rejects_interior_and_collinear_points():
    o = PolygonOracle((894, 1342, 3), 1135)
    square = [(100.0, 100.0), (200.0, 100.0), (200.0, 200.0), (100.0, 200.0)]
    assert o.measure(square + [(150.0, 150.0)]) is None         # interior point
    assert o.measure(square + [(150.0, 100.0)]) is None         # collinear on an edge
    assert o.measure(square + [(150.0, 60.0)]) is not None      # strictly convex
    assert o.measure(square + [(103.0, 60.0)][:1] + []) is not None
    assert o.measure([(100.0, 100.0), (103.0, 101.0), (200.0, 200.0), (100.0, 200.0), (150.0, 60.0)]) is None  # < 5 px


def test_brightness_oracle_equals_reference_values(g):
    gray = g["v0_gray"]
    xy = g["v0_xy"]
    for k in (0, 7, 33, len(xy) - 1):
        assert star_brightness(gray, xy[k]) == int(g["v0_brightness"][k])


# ---- BASELINE config 2: the real eisenberg frames through the unchanged reference ----------------

def test_c2_oracle_reproduces_reference_on_real_star_frames(golden):
    """Stars found by the unchanged compute_star_features on the 4 eisenberg JPEGs the reference
    ships; polygon banks and DMatches of the unchanged BFPolygonMatcher for all 6 pairs."""
    c2 = golden("eisenberg_c2.npz")
    o = PolygonOracle(tuple(int(v) for v in c2["img_shape"]), float(c2["focal"][0]))
    feats, idxs = [], []
    for i in range(len(c2["names"])):
        xy = c2[f"i{i}_xy"].astype(np.float64)
        gray = c2[f"i{i}_gray"]
        bright = np.array([star_brightness(gray, p) for p in xy], dtype=np.float64)
        got = o.characteristics(xy, bright, c2[f"i{i}_candidates"])
        np.testing.assert_array_equal(got["star_idx"], c2[f"i{i}_norm_idx"])
        np.testing.assert_allclose(got["feat"][:, 0:5], c2[f"i{i}_norm_angle"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(got["feat"][:, 5:10], c2[f"i{i}_norm_side"], rtol=0, atol=1e-12)
        np.testing.assert_array_equal(got["feat"][:, 10:15], c2[f"i{i}_norm_bright"])
        feats.append(np.concatenate([c2[f"i{i}_norm_angle"], c2[f"i{i}_norm_side"], c2[f"i{i}_norm_bright"]], axis=1))
        idxs.append(c2[f"i{i}_norm_idx"])
    total = 0
    for p, (a, b) in enumerate(c2["pairs"]):
        rows, cols, d = PolygonOracle.select(PolygonOracle.compare(feats[a], feats[b]), 0.1566)
        votes = PolygonOracle.vote(rows, cols, d, idxs[a], idxs[b])
        np.testing.assert_array_equal(np.array([(x, y) for x, y, _ in votes], dtype=np.int32).reshape(-1, 2),
                                      c2[f"p{p}_dmatches"])
        np.testing.assert_array_equal(np.array([w for _, _, w in votes], dtype=np.float32), c2[f"p{p}_dist"])
        total += len(votes)
    assert total > 0
