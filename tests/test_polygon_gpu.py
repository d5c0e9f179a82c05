"""Parity of CUDA kernels 2 and 3 and of the brightness kernel (through the C ABI) with the
reference's recorded outputs (polygon_golden.npz) and with the oracle."""
import cv2
import numpy as np
import pytest

from oracle.polygon_oracle import PolygonOracle, star_brightness
import ssp_b200
from ssp_b200 import synthetic as synth
from ssp_b200.custom_bf_polygon_matcher import disc_offsets

pytestmark = pytest.mark.gpu
ANGLE_TOL = 1e-5  # rad, the bound north_star states; observed differences are ~1e-13


@pytest.fixture(scope="module")
def g(golden):
    return golden("polygon_golden.npz")


def upload_golden_stars(engine, g):
    h, w = int(g["img_shape"][0]), int(g["img_shape"][1])
    engine.upload_star_bank([g["v0_xy"], g["v1_xy"]],
                            [g["v0_brightness"].astype(np.float64), g["v1_brightness"].astype(np.float64)],
                            w, h, float(g["focal"][0]))


def dedup(cands):
    seen, out = set(), []
    for c in cands.tolist():
        k = tuple(sorted(c))
        if k not in seen:
            seen.add(k)
            out.append(c)
    return np.array(out, dtype=np.int32)


def test_star_rays_match_oracle(engine, g):
    upload_golden_stars(engine, g)
    rays = engine.star_rays()
    o = PolygonOracle(tuple(g["img_shape"]), float(g["focal"][0]))
    xy = np.concatenate([g["v0_xy"], g["v1_xy"]])
    want = np.array([o.ray(p) for p in xy])
    np.testing.assert_allclose(rays, want, rtol=0, atol=2e-16)


@pytest.mark.parametrize("view", [0, 1])
def test_measure_matches_reference_recording(engine, g, view):
    upload_golden_stars(engine, g)
    cands = dedup(g[f"v{view}_candidates"])
    n_valid, status = engine.polygon_measure(view, cands, return_status=True)
    got = engine.polygon_fetch(view)
    ref_idx = g[f"v{view}_raw_idx"]
    assert n_valid == len(ref_idx)
    np.testing.assert_array_equal(got["star_idx"], ref_idx)            # canonical vertex order
    assert np.abs(got["angle"] - g[f"v{view}_raw_angle"]).max() < ANGLE_TOL
    assert np.abs(got["side"] - g[f"v{view}_raw_side"]).max() < ANGLE_TOL
    np.testing.assert_allclose(got["angle"], g[f"v{view}_raw_angle"], rtol=0, atol=1e-10)
    np.testing.assert_allclose(got["side"], g[f"v{view}_raw_side"], rtol=0, atol=1e-12)
    np.testing.assert_array_equal(got["brightness"], g[f"v{view}_raw_bright"])
    np.testing.assert_allclose(got["feat"][:, 0:5], g[f"v{view}_norm_angle"], rtol=0, atol=1e-10)
    np.testing.assert_allclose(got["feat"][:, 5:10], g[f"v{view}_norm_side"], rtol=0, atol=1e-10)
    np.testing.assert_array_equal(got["feat"][:, 10:15], g[f"v{view}_norm_bright"])
    assert set(np.unique(status)) <= {0, 1, 2}


def test_measure_status_codes_and_degenerate_inputs(engine):
    xy = np.array([[100, 100], [200, 100], [200, 200], [100, 200], [150, 150], [150, 100], [150, 60],
                   [103, 101], [300, 300]], dtype=np.float64)
    engine.upload_star_bank([xy], [np.full(len(xy), 100.0)], 1342, 894, 1135.0)
    cands = np.array([[0, 1, 2, 3, 4],    # interior point
                      [0, 1, 2, 3, 5],    # collinear on an edge
                      [0, 1, 2, 3, 6],    # strictly convex
                      [0, 7, 2, 3, 6],    # two stars closer than 5 px
                      [6, 3, 2, 1, 0],    # same star set again, other order: same polygon
                      [0, 1, 2, 3, 99]], dtype=np.int32)  # bad index
    n_valid, status = engine.polygon_measure(0, cands, return_status=True)
    assert status.tolist() == [2, 2, 0, 1, 0, 4]
    got = engine.polygon_fetch(0)
    assert n_valid == 2
    np.testing.assert_array_equal(got["star_idx"][0], got["star_idx"][1])
    o = PolygonOracle((894, 1342, 3), 1135.0)
    m = o.measure([tuple(xy[i]) for i in cands[2]])
    np.testing.assert_array_equal(got["star_idx"][0], cands[2][m["order"]])
    np.testing.assert_allclose(got["angle"][0], m["angles"], rtol=0, atol=1e-12)
    assert engine.polygon_measure(0, np.zeros((0, 5), np.int32)) == 0


def test_measure_against_oracle_on_integer_pixel_field(engine):
    xy_l, b_l, _ = synth.star_fields(1, 400, seed=5)
    engine.upload_star_bank(xy_l, b_l, 1342, 894, 1135.0)
    cands = synth.random_candidates(len(xy_l[0]), 4000, seed=6)
    # keep polygons local like the reference's regions do: sort stars by x and take windows
    order = np.argsort(xy_l[0][:, 0] + 0.3 * xy_l[0][:, 1])
    rng = np.random.default_rng(7)
    for i in range(len(cands)):
        s = rng.integers(0, len(order) - 12)
        cands[i] = order[s + rng.choice(12, size=5, replace=False)]
    cands = dedup(cands)
    n_valid, status = engine.polygon_measure(0, cands, return_status=True)
    got = engine.polygon_fetch(0)
    o = PolygonOracle((894, 1342, 3), 1135.0)
    want = o.characteristics(xy_l[0], b_l[0], cands)
    np.testing.assert_array_equal(status == 0, want["valid"])
    np.testing.assert_array_equal(got["star_idx"], want["star_idx"])
    np.testing.assert_allclose(got["angle"], want["angle"], rtol=0, atol=1e-9)
    np.testing.assert_allclose(got["side"], want["side"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(got["feat"], want["feat"], rtol=0, atol=1e-9)
    assert 200 < n_valid < len(cands)


def features_from_golden(g, v):
    return np.concatenate([g[f"v{v}_norm_angle"], g[f"v{v}_norm_side"], g[f"v{v}_norm_bright"]], axis=1)


def test_compare_is_bit_exact_with_numpy(engine, g):
    f0, f1 = features_from_golden(g, 0), features_from_golden(g, 1)
    engine.polygon_upload_features(0, f0, g["v0_norm_idx"])
    engine.polygon_upload_features(1, f1, g["v1_norm_idx"])
    res = engine.polygon_match_pairs([(0, 1), (1, 0)], return_rowmin=True)
    for p, (a, b) in enumerate([(f0, f1), (f1, f0)]):
        m = PolygonOracle.compare(a, b)
        lo, hi = res.row_offsets[p], res.row_offsets[p + 1]
        np.testing.assert_array_equal(res.argmin[lo:hi], np.argmin(m, axis=1))
        np.testing.assert_array_equal(res.mindist[lo:hi], m.min(axis=1))  # bit for bit
        rows, cols, d = PolygonOracle.select(m, 0.1566)
        r, c, dd = res.pair(p)
        np.testing.assert_array_equal(r, rows)
        np.testing.assert_array_equal(c, cols)
        np.testing.assert_array_equal(dd, d)
    assert res.offsets[1] == 150  # the reference run printed "Remaining 150 polygon matches"


def test_compare_ties_keep_first_column(engine):
    f = synth.polygon_feature_bank(2, 300, seed=1)
    f[1][100] = f[1][7]      # duplicate columns: argmin must keep the first
    f[1][250] = f[1][7]
    f[0][5] = f[1][7]
    engine.polygon_upload_features(0, f[0])
    engine.polygon_upload_features(1, f[1])
    res = engine.polygon_match_pairs([(0, 1)], threshold=10.0, return_rowmin=True)
    m = PolygonOracle.compare(f[0], f[1])
    np.testing.assert_array_equal(res.argmin, np.argmin(m, axis=1))
    np.testing.assert_array_equal(res.mindist, m.min(axis=1))
    assert res.argmin[5] == 7 and res.mindist[5] == 0.0


def test_compare_optional_modes(engine):
    f = synth.polygon_feature_bank(2, 500, seed=2)
    engine.polygon_upload_features(0, f[0])
    engine.polygon_upload_features(1, f[1])
    m = PolygonOracle.compare(f[0], f[1])
    fast = engine.polygon_match_pairs([(0, 1)], threshold=10.0, fp32=True, return_rowmin=True)
    np.testing.assert_allclose(fast.mindist, m.min(axis=1), rtol=1e-5, atol=1e-6)
    shifts = engine.polygon_match_pairs([(0, 1)], threshold=10.0, min_over_shifts=True, return_rowmin=True)
    ms = PolygonOracle.compare_min_over_shifts(f[0], f[1])
    np.testing.assert_array_equal(shifts.mindist, ms.min(axis=1))
    assert (shifts.mindist <= m.min(axis=1)).all()  # a superset of alignments can only lower distances


def test_empty_polygon_banks(engine):
    f = synth.polygon_feature_bank(1, 10, seed=3)
    engine.polygon_upload_features(0, f[0])
    engine.polygon_upload_features(1, np.zeros((0, 15)))
    res = engine.polygon_match_pairs([(0, 1), (1, 0)])
    assert res.offsets.tolist() == [0, 0, 0]


def test_brightness_kernel_matches_reference_values(engine, g):
    for v in (0, 1):
        gray, xy = g[f"v{v}_gray"], g[f"v{v}_xy"]
        centers = np.stack([xy[:, 0].astype(np.int64), xy[:, 1].astype(np.int64)], axis=1)
        got = engine.star_brightness(gray, centers, disc_offsets(10), 10)
        np.testing.assert_array_equal(got, g[f"v{v}_brightness"])


def test_brightness_kernel_at_borders_against_oracle(engine):
    rng = np.random.default_rng(4)
    gray = rng.integers(0, 256, size=(120, 160), dtype=np.uint8)
    pts = [(0, 0), (159, 119), (3, 117), (80, 60), (155, 2), (9, 10), (10, 9), (150, 110)]
    got = engine.star_brightness(gray, np.array(pts), disc_offsets(10), 10)
    want = [star_brightness(gray, p) for p in pts]
    assert got.tolist() == want


def test_full_match_reproduces_reference_dmatches(engine, g, tmp_path):
    """BFPolygonMatcher.match end to end (brightness -> candidates -> measure -> compare -> vote)
    against the DMatches the unchanged reference produced on the same two views."""
    h, w = int(g["img_shape"][0]), int(g["img_shape"][1])
    kps, imgs = [], []
    for v in (0, 1):
        xy, br = g[f"v{v}_xy"], None
        kps.append(tuple(cv2.KeyPoint(float(p[0]), float(p[1]), 4.0) for p in xy.astype(np.float32)))
        gray = g[f"v{v}_gray"]
        imgs.append(cv2.cvtColor(gray, cv2.COLOR_GRAY2BGR))
    bf = ssp_b200.BFPolygonMatcher(str(tmp_path), "view0", "view1", 0, 1, imgs[0].shape, 1135, engine=engine,
                                   write_debug_images=False)
    store, counter = {}, []
    dms = bf.match(kps[0], kps[1], 0, 0, imgs[0], imgs[1], store, 0, 1, counter)
    assert counter == [0, 1]
    for v in (0, 1):
        p = store[v]["polygons_from_ftX"]
        # same candidate set (sklearn clustering + regions) -> same polygons in the same order
        np.testing.assert_array_equal(p["keypoint_indices_rotated_list"], g[f"v{v}_norm_idx"])
        np.testing.assert_allclose(p["keypoint_angles_rotated_list"], g[f"v{v}_norm_angle"], rtol=0, atol=1e-10)
    got = np.array([[m.queryIdx, m.trainIdx] for m in dms], dtype=np.int32)
    np.testing.assert_array_equal(got, g["dmatches"])
    np.testing.assert_allclose(np.array([m.distance for m in dms], dtype=np.float32), g["dmatch_dist"],
                               rtol=0, atol=1e-6)
    assert len(bf.last_polygon_matches[0]) == 150


def test_c2_real_star_frames_all_pairs_reproduce_reference(engine, golden, tmp_path):
    """BASELINE config 2: BFPolygonMatcher.match (brightness kernel -> host candidates -> kernel 2
    -> kernel 3 -> voting) on the 4 eisenberg frames the reference ships, all 6 pairs with the
    polygon cache shared across pairs; DMatches and polygon banks equal the unchanged reference's."""
    c2 = golden("eisenberg_c2.npz")
    n = len(c2["names"])
    kps, imgs = [], []
    for i in range(n):
        kps.append(tuple(cv2.KeyPoint(float(p[0]), float(p[1]), float(s))
                         for p, s in zip(c2[f"i{i}_xy"], c2[f"i{i}_size"])))
        imgs.append(cv2.cvtColor(c2[f"i{i}_gray"], cv2.COLOR_GRAY2BGR))
    store, counter = {}, []
    total = 0
    for p, (a, b) in enumerate(c2["pairs"]):
        a, b = int(a), int(b)
        bf = ssp_b200.BFPolygonMatcher(str(tmp_path), str(c2["names"][a]), str(c2["names"][b]), a, b,
                                       imgs[a].shape, 1135, engine=engine, write_debug_images=False)
        dms = bf.match(kps[a], kps[b], 0, 0, imgs[a], imgs[b], store, a, b, counter)
        got = np.array([[m.queryIdx, m.trainIdx] for m in dms], dtype=np.int32).reshape(-1, 2)
        np.testing.assert_array_equal(got, c2[f"p{p}_dmatches"], err_msg=f"pair {(a, b)}")
        np.testing.assert_allclose(np.array([m.distance for m in dms], dtype=np.float32), c2[f"p{p}_dist"],
                                   rtol=0, atol=1e-6)
        total += len(dms)
    assert counter == list(range(n)) and total > 0
    for i in range(n):
        q = store[i]["polygons_from_ftX"]
        np.testing.assert_array_equal(q["keypoint_indices_rotated_list"], c2[f"i{i}_norm_idx"])
        assert np.abs(q["keypoint_angles_rotated_list"] * np.pi - c2[f"i{i}_norm_angle"] * np.pi).max() < 1e-5
        np.testing.assert_allclose(q["side_lengths_rotated_list"], c2[f"i{i}_norm_side"], rtol=0, atol=1e-10)
        np.testing.assert_array_equal(q["keypoint_brightnesses_rotated_list"], c2[f"i{i}_norm_bright"])


def test_no_wait_measure_of_a_batch_equals_the_waiting_form(engine):
    """ssp_polygon_measure without status / n_valid does not block; counts (and banks) delivered by
    the next consumer equal those of the waiting calls, also when images are re-measured."""
    import torch
    xy_l, b_l, _ = synth.star_fields(3, 300, seed=9)
    engine.upload_star_bank(xy_l, b_l, 1342, 894, 1135.0)
    cands = [synth.random_candidates(len(xy_l[i]), 1500 + 500 * i, seed=20 + i) for i in range(3)]
    want = []
    for i in range(3):
        n = engine.polygon_measure(i, cands[i])
        want.append((n, engine.polygon_fetch(i)))
    ref = engine.polygon_match_pairs([(0, 1), (1, 2), (0, 2)], threshold=10.0, return_rowmin=True)
    dev = [torch.from_numpy(c).cuda() for c in cands]
    for rep in range(2):
        for i in (2, 0, 1):
            assert engine.polygon_measure_resident(i, dev[i].data_ptr(), len(cands[i]), wait=False) is None
        for i in range(3):
            assert engine.polygon_count(i) == want[i][0]
            got = engine.polygon_fetch(i)
            for k in ("star_idx", "feat", "angle", "side", "brightness"):
                np.testing.assert_array_equal(got[k], want[i][1][k])
    for i in range(3):
        engine.polygon_measure_resident(i, dev[i].data_ptr(), len(cands[i]), wait=False)
    res = engine.polygon_match_pairs([(0, 1), (1, 2), (0, 2)], threshold=10.0, return_rowmin=True)
    np.testing.assert_array_equal(res.argmin, ref.argmin)
    np.testing.assert_array_equal(res.mindist, ref.mindist)
    np.testing.assert_array_equal(res.offsets, ref.offsets)


# ---- row N2: candidate generation on the device ------------------------------------------------------

def _c2_frames(c2):
    kps, imgs = [], []
    for i in range(len(c2["names"])):
        kps.append(tuple(cv2.KeyPoint(float(p[0]), float(p[1]), float(s))
                         for p, s in zip(c2[f"i{i}_xy"], c2[f"i{i}_size"])))
        imgs.append(cv2.cvtColor(c2[f"i{i}_gray"], cv2.COLOR_GRAY2BGR))
    return kps, imgs


def test_star_regions_kernel_equals_the_reference_region_filters(engine, golden):
    """cbpm:328-375: 12 overlapping regions, strict float64 comparisons -- including stars placed
    exactly ON the region borders."""
    c2 = golden("eisenberg_c2.npz")
    kps, imgs = _c2_frames(c2)
    h, w = imgs[0].shape[:2]
    xy = np.array([kp.pt for kp in kps[0]], dtype=np.float64)
    border = np.array([[np.float32(0.33 * w), np.float32(0.70 * h)], [0.5 * w, 0.5 * h], [np.float32(0.67 * w), np.float32(0.3 * h)],
                       [0.5 * w - 0.25 * w, 10.0], [0.5 * w + 0.25 * w, h - 10.0], [w * 0.33, h * 0.33]], dtype=np.float64)
    xy = np.concatenate([xy, border.astype(np.float32).astype(np.float64)])
    engine.upload_star_bank([xy], [np.ones(len(xy))], w, h, 1135.0)
    masks = engine.star_regions(0)
    bf = ssp_b200.BFPolygonMatcher("/tmp", "a", "b", 0, 1, imgs[0].shape, 1135, engine=engine, write_debug_images=False)
    all_kps = tuple(cv2.KeyPoint(float(np.float32(p[0])), float(np.float32(p[1])), 4.0) for p in xy)
    want = bf._regions(all_kps, w, h)
    for r, name in enumerate(want):
        assert np.nonzero(masks & (1 << r))[0].tolist() == want[name], name


def test_cluster_brightest_kernel_keeps_the_last_of_equally_bright_stars(engine):
    rng = np.random.default_rng(3)
    n = 400
    xy = rng.uniform(20, 800, size=(n, 2))
    bright = rng.integers(2500, 2551, size=n).astype(np.float64)   # saturated stars: ties are the norm
    engine.upload_star_bank([xy], [bright], 1342, 894, 1135.0)
    members, labels, off = [], [], [0]
    for g in range(5):
        m = rng.permutation(n)[:60 + 10 * g].tolist()
        members += m
        labels += rng.integers(0, 15, size=len(m)).tolist()
        off.append(len(members))
    got = engine.cluster_brightest(0, members, labels, off, 15)
    for g in range(5):
        mem, lab = members[off[g]:off[g + 1]], labels[off[g]:off[g + 1]]
        for label in range(15):
            idx = [i for i in range(len(mem)) if lab[i] == label]
            # cbpm:241: ascending stable sort by brightness, reversed, first element
            want = sorted(((i, bright[mem[i]]) for i in idx), key=lambda x: x[1])[::-1][0][0] if idx else None
            assert got[g, label] == (mem[want] if idx else -1)


def test_device_candidate_generation_yields_the_reference_candidate_set(engine, golden, tmp_path):
    """Row N2 on the 4 real eisenberg frames: the candidate SET per image equals the one the unchanged
    reference enumerated (recorded), hence the same polygon set; matches and DMatches agree as sets."""
    c2 = golden("eisenberg_c2.npz")
    kps, imgs = _c2_frames(c2)
    n = len(kps)
    store, counter = {}, []
    dm_sets = []
    for p, (a, b) in enumerate(c2["pairs"]):
        a, b = int(a), int(b)
        bf = ssp_b200.BFPolygonMatcher(str(tmp_path), f"f{a}", f"f{b}", a, b, imgs[a].shape, 1135, engine=engine,
                                       write_debug_images=False, candidate_generation="device")
        dms = bf.match(kps[a], kps[b], 0, 0, imgs[a], imgs[b], store, a, b, counter)
        assert {(m.queryIdx, m.trainIdx) for m in dms} == {tuple(x) for x in c2[f"p{p}_dmatches"].tolist()}, (a, b)
    for i in range(n):
        want = {tuple(sorted(c)) for c in c2[f"i{i}_candidates"].tolist()}
        got = {tuple(sorted(c)) for c in store[i]["all_possible_polygons_ftX"]}
        assert got == want and len(store[i]["all_possible_polygons_ftX"]) == len(want), i
        ref_polys = {tuple(r) for r in c2[f"i{i}_norm_idx"].tolist()}
        assert {tuple(r) for r in store[i]["polygons_from_ftX"]["keypoint_indices_rotated_list"].tolist()} == ref_polys


# ---- kernel 3: the compacted-pruning variant is the same function -----------------------------------

@pytest.mark.parametrize("rows,cols", [(1, 1), (33, 257), (700, 4100), (3000, 3000)])
def test_queue_compare_kernel_equals_per_lane_pruning_and_numpy(engine, monkeypatch, rows, cols):
    """The default kernel (single-precision screen, FP64 evaluation of what passes) and
    SSP_B200_POLY_COMPARE=queue (survivors of the first property group are queued per warp and
    evaluated as dense batches) must return the per-lane kernel's argmin / mindist bit for bit, also
    with duplicate columns (ties keep the first) and ragged tile ends."""
    feat = synth.polygon_feature_bank(2, max(rows, cols), seed=41)
    f1, f2 = feat[0][:rows].copy(), feat[1][:cols].copy()
    if cols > 40:
        f2[cols // 2] = f2[7]            # duplicate columns: equal distances, the earlier one must win
        f2[cols - 1] = f2[3]
    if rows > 5 and cols > 9:
        f1[5] = f2[9]                    # an exact match: distance 0
    engine.polygon_upload_features(0, f1)
    engine.polygon_upload_features(1, f2)
    out = {}
    for kind in ("legacy", "queue", "tensor", "screen"):
        monkeypatch.setenv("SSP_B200_POLY_COMPARE", kind)
        out[kind] = engine.polygon_match_pairs([(0, 1)], threshold=10.0, return_rowmin=True)
        st = engine.polygon_last_stats()
        assert (st["executed_flops"] is None) == (kind == "legacy")
        ts = st["tensor_screen"]
        assert ts["used"] == (kind == "tensor") and ts["fallback_flags"] == 0
        if kind == "tensor":
            # every row has at least one candidate; the screen stays far inside its error bound
            assert rows <= ts["candidates"] <= 24 * rows + 8192 and ts["worst_error_coef"] < ts["bound_coef"] / 4
    m = PolygonOracle.compare(f1, f2)
    for kind in ("queue", "screen", "tensor"):
        np.testing.assert_array_equal(out[kind].argmin, out["legacy"].argmin)
        np.testing.assert_array_equal(out[kind].mindist, out["legacy"].mindist)
        np.testing.assert_array_equal(out[kind].argmin, np.argmin(m, axis=1))
        np.testing.assert_array_equal(out[kind].mindist, m.min(axis=1))
    # the screen evaluates every cell once in single precision and only what passes in FP64
    assert st["executed_flops_fp32"] == 44 * rows * cols and 45 * rows <= st["executed_flops_fp64"] <= 45 * rows * cols


@pytest.mark.parametrize("case", ["offset_1000", "differences_below_float_resolution", "scale_1e20",
                                  "scale_1e100", "first_group_sum_at_the_threshold",
                                  "every_column_improves_every_row", "nan_and_inf_features",
                                  "all_zero_features"])
def test_compare_float_screens_never_drop_a_column_they_must_keep(engine, monkeypatch, case):
    """The default kernel screens every cell in packed single precision (all 15 properties), the queue
    kernel its first property group, each with a margin that makes the float sum conservative
    (csrc/polygon.cu).  Inputs chosen against that margin: features far from [0, 1], differences a
    float cannot resolve, magnitudes beyond the float range, sums within a few double ulps of the row's
    threshold, a bank in which every cell passes (queue pressure), and NaN / inf features."""
    rng = np.random.default_rng(77)
    rows, cols = 300, 700
    f1 = rng.uniform(0.0, 1.0, size=(rows, 15))
    f2 = rng.uniform(0.0, 1.0, size=(cols, 15))
    if case == "offset_1000":
        f1 += 1000.0
        f2 += 1000.0
    elif case == "differences_below_float_resolution":
        base = rng.uniform(0.2, 0.8, size=15)
        f1 = base + rng.uniform(-1e-9, 1e-9, size=(rows, 15))
        f2 = base + rng.uniform(-1e-9, 1e-9, size=(cols, 15))
    elif case == "scale_1e20":
        f1 *= 1e20
        f2 *= 1e20
    elif case == "scale_1e100":
        f1 *= 1e100
        f2 *= 1e100
    elif case == "every_column_improves_every_row":
        base = rng.uniform(0.2, 0.8, size=15)
        f1 = base + rng.uniform(-1e-4, 1e-4, size=(rows, 15))
        f2 = np.repeat(base[None, :], cols, axis=0)
        f2[:, 3] += (cols - np.arange(cols)) * 1e-3       # distances fall monotonically with the column
    elif case == "all_zero_features":
        f1[:] = 0.0                                        # every column ties at distance 0: column 0 wins
        f2[:] = 0.0
    elif case == "nan_and_inf_features":
        f2[5, 2] = np.nan
        f2[17, 11] = np.inf
        f1[3, 7] = np.nan
        f1[9, 0] = -np.inf
    else:
        # every column differs from row r only in ONE feature of ONE group by the same amount up to
        # an ulp: all three group orders see sums that straddle the running threshold
        f2 = np.repeat(f1[:1], cols, axis=0)
        step = 0.125 * (1.0 + np.arange(cols) % 7 * 2.0 ** -52)
        for c in range(cols):
            f2[c, (c * 5) % 15] += step[c]
        f1 = np.repeat(f1[:1], rows, axis=0)
    engine.polygon_upload_features(0, f1)
    engine.polygon_upload_features(1, f2)
    out = {}
    for kind in ("legacy", "queue", "screen", "tensor"):
        monkeypatch.setenv("SSP_B200_POLY_COMPARE", kind)
        out[kind] = engine.polygon_match_pairs([(0, 1)], threshold=1e300, return_rowmin=True)
    ts = engine.polygon_last_stats()["tensor_screen"]
    # features beyond the float range (also +-inf) make the tensor-core screen hand the call over
    want_flags = 2 if case in ("scale_1e20", "scale_1e100", "nan_and_inf_features") else None
    if want_flags is not None:
        assert not ts["used"] and ts["fallback_flags"] & want_flags
    else:
        assert ts["used"] or ts["fallback_flags"] == 4, ts      # 4: candidate pool overflow (every cell a candidate)
        assert ts["worst_error_coef"] is None or ts["worst_error_coef"] < ts["bound_coef"] / 4 or not ts["used"]
    for kind in ("queue", "screen", "tensor"):
        np.testing.assert_array_equal(out[kind].argmin, out["legacy"].argmin)
        np.testing.assert_array_equal(out[kind].mindist, out["legacy"].mindist)
    if case != "nan_and_inf_features":     # numpy's argmin stops at a NaN; the kernels skip it like the per-lane one
        m = PolygonOracle.compare(f1, f2)
        np.testing.assert_array_equal(out["screen"].argmin, np.argmin(m, axis=1))
        np.testing.assert_array_equal(out["screen"].mindist, m.min(axis=1))


def test_batch_measure_of_all_images_equals_the_per_image_form(engine):
    """ssp_polygon_measure_batch: several images, ONE measure + ONE compaction launch; banks, counts and
    the matches that follow equal those of the per-image calls -- also with an image without
    candidates and with invalid candidates in the mix."""
    import torch
    xy_l, b_l, _ = synth.star_fields(4, 300, seed=12)
    engine.upload_star_bank(xy_l, b_l, 1342, 894, 1135.0)
    cands = [synth.random_candidates(len(xy_l[i]), 900 + 333 * i, seed=50 + i) for i in range(4)]
    cands[2] = cands[2][:0]
    want = []
    for i in range(4):
        n = engine.polygon_measure(i, cands[i])
        want.append((n, engine.polygon_fetch(i)))
    ref = engine.polygon_match_pairs([(0, 1), (1, 3), (0, 3), (2, 3)], threshold=10.0, return_rowmin=True)
    offs = np.zeros(5, dtype=np.int64)
    np.cumsum([len(c) for c in cands], out=offs[1:])
    table = torch.from_numpy(np.concatenate(cands).astype(np.int32)).cuda()
    order = [3, 0, 2, 1]                         # any order of images, offsets follow the order
    perm_offs = np.zeros(5, dtype=np.int64)
    np.cumsum([len(cands[i]) for i in order], out=perm_offs[1:])
    perm_table = torch.from_numpy(np.concatenate([cands[i] for i in order]).astype(np.int32)).cuda()
    for ids, tab, o in (([0, 1, 2, 3], table, offs), (order, perm_table, perm_offs)):
        engine.polygon_measure_batch_resident(ids, tab.data_ptr(), o)
        for i in range(4):
            assert engine.polygon_count(i) == want[i][0]
            got = engine.polygon_fetch(i)
            for k in ("star_idx", "angle", "side", "brightness", "feat"):
                np.testing.assert_array_equal(got[k], want[i][1][k], err_msg=f"image {i} {k}")
        res = engine.polygon_match_pairs([(0, 1), (1, 3), (0, 3), (2, 3)], threshold=10.0, return_rowmin=True)
        np.testing.assert_array_equal(res.argmin, ref.argmin)
        np.testing.assert_array_equal(res.mindist, ref.mindist)


def test_compute_star_features_mirror_reproduces_the_reference_star_finder(engine, golden, tmp_path):
    """image_processors.compute_star_features (Canny -> contours -> brightness -> the n brightest -> ORB)
    with the per-star brightness loop replaced by ONE launch of the brightness kernel: keypoints, their
    order and the descriptors equal what the UNCHANGED reference produced on the 4 eisenberg frames
    (eisenberg_c2.npz: keypoints; eisenberg_c2_stage.npz: the star rows of the merged descriptors)."""
    from ssp_b200.image_processors import compute_star_features, get_star_brightness
    c2 = golden("eisenberg_c2.npz")
    st = golden("eisenberg_c2_stage.npz")
    for i in range(len(c2["names"])):
        img = cv2.cvtColor(c2[f"i{i}_gray"], cv2.COLOR_GRAY2BGR)
        feats = compute_star_features(img, str(c2["names"][i]), str(tmp_path), n_brightest_stars=1000, engine=engine,
                                      write_debug_images=False)
        xy = np.array([kp.pt for kp in feats.keypoints], dtype=np.float32).reshape(-1, 2)
        np.testing.assert_array_equal(xy, c2[f"i{i}_xy"], err_msg=f"frame {i} keypoint positions / order")
        np.testing.assert_array_equal(np.array([kp.size for kp in feats.keypoints], dtype=np.float32), c2[f"i{i}_size"])
        n_orb = int(st["n_orb"][i])
        np.testing.assert_array_equal(feats.descriptors.get(), st[f"i{i}_descriptors"][n_orb:])
        assert feats.img_size == (img.shape[1], img.shape[0])
    kp = feats.keypoints[0]
    assert get_star_brightness(kp, img, engine) == star_brightness(c2[f"i{len(c2['names']) - 1}_gray"], kp.pt)
