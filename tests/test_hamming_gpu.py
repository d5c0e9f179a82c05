"""Parity of CUDA kernel 1 (through the C ABI) with the oracle and with cv2.BFMatcher: bit-exact
distances, tie-broken indices and survivor sets."""
import cv2
import numpy as np
import pytest

from oracle import hamming_oracle as ho
import ssp_b200
from ssp_b200 import synthetic as synth

pytestmark = pytest.mark.gpu
VARIANTS = ["popc", "popc_csa", "tensor"]


def oracle_pairs(sets, pairs, mode):
    out = []
    for a, b in pairs:
        if mode == "crosscheck":
            out.append(ho.crosscheck(sets[a], sets[b]))
        else:
            out.append(ho.ratio_survivors(ho.knn2(sets[a], sets[b])))
    return out


def assert_result_equals(res, expected):
    for p, (mq, mt, md) in enumerate(expected):
        q, t, d = res.pair(p)
        np.testing.assert_array_equal(q, mq, err_msg=f"pair {p} queryIdx")
        np.testing.assert_array_equal(t, mt, err_msg=f"pair {p} trainIdx")
        np.testing.assert_array_equal(d, md, err_msg=f"pair {p} distance")


@pytest.mark.parametrize("variant", VARIANTS)
def test_golden_vectors_from_cv2(engine, golden, variant):
    g = golden("hamming_golden.npz")
    for name in g["case_names"]:
        q, t = g[f"{name}_q"], g[f"{name}_t"]
        engine.upload_descriptor_bank([q, t])
        res = engine.hamming_match_pairs([(0, 1)], mode="knn2_ratio", variant=variant, return_best2=True)
        np.testing.assert_array_equal(res.best2, g[f"{name}_knn2"], err_msg=f"{name} knn2")
        res = engine.hamming_match_pairs([(0, 1)], mode="crosscheck", variant=variant)
        got = np.stack(res.pair(0), axis=1)
        np.testing.assert_array_equal(got, g[f"{name}_cross"], err_msg=f"{name} crosscheck")


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("mode", ["crosscheck", "knn2_ratio"])
def test_ragged_multi_pair_batch_against_oracle(engine, variant, mode):
    sizes = [0, 1, 2, 255, 256, 257, 1023, 1024, 1025, 1500]
    tie = synth.tie_stress_bank(len(sizes), max(sizes), seed=7)
    sets = [tie[i][:n] for i, n in enumerate(sizes)]
    pairs = [(i, j) for i in range(len(sizes)) for j in range(len(sizes)) if i != j]
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs(pairs, mode=mode, variant=variant)
    assert_result_equals(res, oracle_pairs(sets, pairs, mode))


@pytest.mark.parametrize("mode", ["crosscheck", "knn2_ratio"])
def test_tie_stress_over_many_train_tiles(engine, mode):
    """Ties across MANY 256-row train tiles: the tensor variant's two warpgroups exchange
    thresholds (with a tie margin that is dropped once a value is two tiles old), so equal
    distances in neighbouring tiles and in the two halves of one tile all have to resolve to the
    lowest train index."""
    sizes = [8191, 7937, 8000]
    tie = synth.tie_stress_bank(len(sizes), max(sizes), seed=11, dup_frac=0.3)
    sets = [tie[i][:n] for i, n in enumerate(sizes)]
    pairs = [(i, j) for i in range(3) for j in range(3) if i != j]
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs(pairs, mode=mode, variant="tensor", return_best2=(mode == "knn2_ratio"))
    assert_result_equals(res, oracle_pairs(sets, pairs, mode))
    if mode == "knn2_ratio":
        for p, (a, b) in enumerate(pairs):
            lo, hi = res.best2_offsets[p], res.best2_offsets[p + 1]
            np.testing.assert_array_equal(res.best2[lo:hi], ho.knn2(sets[a], sets[b]))


def test_crosscheck_sparse_reverse_pass_many_items_per_block(engine):
    """The tensor variant runs the reverse direction of the cross-check over the CLAIMED train
    rows only (sparse query tiles, gathered row by row).  42 ordered pairs of 6 000 rows give
    every thread block several work items in a row, tail tiles with empty slots included; twice,
    because the hazards of this path are timing-dependent."""
    sets = synth.descriptor_bank(7, 6000, seed=5, world=20_000)
    pairs = [(i, j) for i in range(7) for j in range(7) if i != j]
    engine.upload_descriptor_bank(sets)
    expected = oracle_pairs(sets, pairs, "crosscheck")
    for _ in range(2):
        assert_result_equals(engine.hamming_match_pairs(pairs, mode="crosscheck"), expected)
    assert_result_equals(engine.hamming_match_pairs(pairs, mode="crosscheck", variant="popc_csa"), expected)


@pytest.mark.parametrize("variant", VARIANTS)
def test_best2_records_against_oracle_with_planted_matches(engine, variant):
    sets = synth.descriptor_bank(3, 3000, seed=4, world=4000)
    engine.upload_descriptor_bank(sets)
    pairs = [(0, 1), (1, 2), (2, 0)]
    res = engine.hamming_match_pairs(pairs, mode="knn2_ratio", variant=variant, return_best2=True)
    for p, (a, b) in enumerate(pairs):
        lo, hi = res.best2_offsets[p], res.best2_offsets[p + 1]
        np.testing.assert_array_equal(res.best2[lo:hi], ho.knn2(sets[a], sets[b]))
    assert res.offsets[-1] > 500  # planted matches survive the ratio test


def test_against_installed_cv2_bfmatcher(engine):
    sets = synth.descriptor_bank(2, 2000, seed=9, world=2500)
    bf = ssp_b200.BFMatcher(cv2.NORM_HAMMING, crossCheck=True, engine=engine)
    ours = bf.match(cv2.UMat(sets[0]), cv2.UMat(sets[1]))
    theirs = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(cv2.UMat(sets[0]), cv2.UMat(sets[1]))
    assert [(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in ours] == \
           [(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in theirs]
    knn_ours = ssp_b200.BFMatcher(cv2.NORM_HAMMING, engine=engine).knnMatch(sets[0], sets[1], k=2)
    knn_theirs = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(sets[0], sets[1], k=2)
    assert [[(m.queryIdx, m.trainIdx, m.distance) for m in row] for row in knn_ours] == \
           [[(m.queryIdx, m.trainIdx, m.distance) for m in row] for row in knn_theirs]


def test_real_orb_fixture_all_210_pairs(engine, golden):
    g = golden("orb_forest_c1.npz")
    offs = g["offsets"]
    sets = [g["descriptors"][offs[i]:offs[i + 1]] for i in range(len(offs) - 1)]
    pairs = synth.all_pairs(len(sets))
    assert len(pairs) == 210
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs(pairs, mode="crosscheck")
    assert_result_equals(res, oracle_pairs(sets, pairs, "crosscheck"))
    counts = np.diff(res.offsets)
    assert 100 < counts.min() and counts.max() < 400  # the author's runs: 149-294 per pair


def test_full_size_pair_50k_against_oracle_and_properties(engine):
    """One C3-sized pair (50 000 x 50 000) against the C oracle, plus size-independent
    properties on a second pair: mirror symmetry of cross-check and distance recomputation."""
    sets = synth.descriptor_bank(3, 50_000, seed=0)
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs([(0, 1), (1, 0), (1, 2)], mode="crosscheck")
    mq, mt, md = ho.crosscheck(sets[0], sets[1])
    q, t, d = res.pair(0)
    np.testing.assert_array_equal(q, mq)
    np.testing.assert_array_equal(t, mt)
    np.testing.assert_array_equal(d, md)
    # (1,0) is the mirror of (0,1): same couples with roles swapped
    q2, t2, d2 = res.pair(1)
    order = np.argsort(t2, kind="stable")
    np.testing.assert_array_equal(t2[order], q)
    np.testing.assert_array_equal(q2[order], t)
    np.testing.assert_array_equal(d2[order], d)
    # reported distances are the true Hamming distances of the reported couples
    q3, t3, d3 = res.pair(2)
    x = np.unpackbits(sets[1][q3] ^ sets[2][t3], axis=1).sum(axis=1)
    np.testing.assert_array_equal(x, d3)
    assert len(q) > 10_000


def test_full_size_knn2_ratio_sampled_rows(engine):
    sets = synth.descriptor_bank(2, 50_000, seed=0)
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs([(0, 1)], mode="knn2_ratio", return_best2=True)
    rows = np.random.default_rng(1).choice(50_000, size=256, replace=False)
    np.testing.assert_array_equal(res.best2[rows], ho.knn2(sets[0][rows], sets[1]))
    mq, mt, md = ho.ratio_survivors(res.best2)
    np.testing.assert_array_equal(res.query_idx, mq)
    np.testing.assert_array_equal(res.train_idx, mt)
    np.testing.assert_array_equal(res.distance, md)


def test_errors_are_loud(engine):
    engine.upload_descriptor_bank([np.zeros((4, 32), np.uint8)])
    with pytest.raises(ssp_b200.EngineError, match="out of range"):
        engine.hamming_match_pairs([(0, 3)])
    assert engine.hamming_match_pairs([]).offsets.tolist() == [0]
    # OpenCV's own limit on rows per train set (IMGIDX_SHIFT = 18) is an error, not a wrong answer
    engine.upload_descriptor_bank([np.zeros((8, 32), np.uint8), np.zeros((1 << 18, 32), np.uint8)])
    with pytest.raises(ssp_b200.EngineError, match="2\\^18"):
        engine.hamming_match_pairs([(0, 1)], mode="knn2_ratio")
    with pytest.raises(ssp_b200.EngineError, match="2\\^18"):
        engine.hamming_match_pairs([(1, 0)], mode="crosscheck")
    assert engine.hamming_match_pairs([(1, 0)], mode="knn2_ratio").offsets.tolist() == [0, 0]  # 8 equal rows: no ratio survivors
    fresh = ssp_b200.Engine(0)
    with pytest.raises(ssp_b200.EngineError, match="no descriptor bank"):
        fresh.hamming_match_pairs([(0, 0)])
    fresh.close()


def test_pinned_results_and_fetch_into_equal_plain_fetch(engine):
    """The page-locked result views (ssp_host_alloc) and the fetch-by-address path used by the
    multi-GPU gather return exactly what the plain numpy fetch returns, also after the buffers
    were reused for a smaller and then a larger result."""
    sets = synth.descriptor_bank(4, 900, seed=11, world=1200)
    engine.upload_descriptor_bank(sets)
    for pairs in ([(0, 1), (1, 2), (2, 3)], [(0, 2)], synth.all_pairs(4)):
        plain = engine.hamming_match_pairs(pairs, mode="knn2_ratio")
        pinned = engine.hamming_match_pairs(pairs, mode="knn2_ratio", pinned=True)
        np.testing.assert_array_equal(plain.offsets, pinned.offsets)
        for a, b in zip((plain.query_idx, plain.train_idx, plain.distance),
                        (pinned.query_idx, pinned.train_idx, pinned.distance)):
            np.testing.assert_array_equal(a, b)
        lazy = engine.hamming_match_pairs(pairs, mode="knn2_ratio", fetch=False)
        n = int(lazy.offsets[-1])
        bufs = [engine.pinned_array(f"test{k}", n, np.int32) for k in range(3)]
        engine.hamming_fetch_into(*(b.ctypes.data for b in bufs), n, device=False)
        for a, b in zip((plain.query_idx, plain.train_idx, plain.distance), bufs):
            np.testing.assert_array_equal(a, b)
    assert plain.offsets[-1] > 0


# ---- round 2: the three forms of the tensor variant's cross-check, larger multi-item cases ----------

CROSS_FORMS = ["fused", "sparse", "dense"]


@pytest.fixture
def cross_form(monkeypatch):
    def use(form):
        monkeypatch.setenv("SSP_B200_CROSSCHECK", form)   # read per call by libssp_b200.so
    return use


@pytest.mark.parametrize("form", CROSS_FORMS)
def test_crosscheck_forms_on_ragged_tie_stress(engine, cross_form, form):
    """Every form against the oracle on tie-heavy ragged sets, including pairs in both orders where
    one image has no rows at all (forward launch empty, reverse work present: ADVICE r1)."""
    cross_form(form)
    sizes = [0, 1, 7, 128, 129, 255, 256, 257, 700, 1500, 2100]
    tie = synth.tie_stress_bank(len(sizes), max(sizes), seed=17)
    sets = [tie[i][:n] for i, n in enumerate(sizes)]
    pairs = [(i, j) for i in range(len(sizes)) for j in range(len(sizes)) if i != j]
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs(pairs, mode="crosscheck", variant="tensor")
    assert_result_equals(res, oracle_pairs(sets, pairs, "crosscheck"))
    only_empty_queries = [(0, 9), (9, 0), (0, 10), (10, 0)]
    res = engine.hamming_match_pairs(only_empty_queries, mode="crosscheck", variant="tensor")
    assert res.offsets.tolist() == [0, 0, 0, 0, 0]


@pytest.mark.parametrize("form", CROSS_FORMS)
def test_crosscheck_forms_tie_stress_over_many_tiles_and_items(engine, cross_form, form):
    """Ties across many train tiles AND several work items per CTA: 5 images x 9 000 tie-heavy rows,
    all 20 ordered pairs = 1 420 query tiles over 148 CTAs.  In the fused form the train sweep of
    every CTA is rotated, so equal distances arrive in every index order."""
    cross_form(form)
    sets = synth.tie_stress_bank(5, 9000, seed=23)
    pairs = [(i, j) for i in range(5) for j in range(5) if i != j]
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs(pairs, mode="crosscheck", variant="tensor")
    assert_result_equals(res, oracle_pairs(sets, pairs, "crosscheck"))


@pytest.mark.parametrize("mode", ["crosscheck", "knn2_ratio"])
def test_c5_sized_items_20k_several_items_per_cta(engine, mode):
    """BASELINE configs[4] shape: 20 000 x 20 000 items (79 train tiles, where warps spend most of an
    item settling their thresholds), 4 images -> 6 pairs -> 942 query tiles, 6+ items per CTA."""
    sets = synth.descriptor_bank(4, 20_000, seed=1, world=60_000)
    pairs = synth.all_pairs(4)
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs(pairs, mode=mode, return_best2=(mode == "knn2_ratio"))
    assert_result_equals(res, oracle_pairs(sets, pairs, mode))
    if mode == "knn2_ratio":
        for p, (a, b) in enumerate(pairs[:2]):
            lo, hi = res.best2_offsets[p], res.best2_offsets[p + 1]
            np.testing.assert_array_equal(res.best2[lo:hi], ho.knn2(sets[a], sets[b]))
    assert res.offsets[-1] > 1000


@pytest.mark.parametrize("mode", ["crosscheck", "knn2_ratio"])
def test_full_size_tie_stress_pair_50k(engine, mode):
    """50 000 x 50 000 rows drawn from a 2-bit-per-byte alphabet with exact duplicates: thousands of
    equal distances per row and per column, at the headline size."""
    sets = synth.tie_stress_bank(2, 50_000, seed=29)
    engine.upload_descriptor_bank(sets)
    res = engine.hamming_match_pairs([(0, 1)], mode=mode, return_best2=(mode == "knn2_ratio"))
    assert_result_equals(res, oracle_pairs(sets, [(0, 1)], mode))
    if mode == "knn2_ratio":
        np.testing.assert_array_equal(res.best2, ho.knn2(sets[0], sets[1]))


def test_ratio_is_pythons_float64_comparison_for_any_ratio(engine):
    """`m.distance < ratio * n.distance` (custom_bruteforce_matcher.py:164) for ratios that have no
    small rational form; 0.7000001 keeps (7, 10), which a rounded 7/10 would reject."""
    sets = synth.descriptor_bank(2, 3000, seed=31, world=3500)
    engine.upload_descriptor_bank(sets)
    for ratio in (0.75, 0.7000001, 0.1, 1.0 / 3.0, 0.9999999999, 0.8):
        res = engine.hamming_match_pairs([(0, 1)], mode="knn2_ratio", ratio=ratio, return_best2=True)
        b = res.best2
        keep = [i for i in range(len(b)) if b[i, 2] >= 0 and float(np.float32(b[i, 1])) < ratio * float(np.float32(b[i, 3]))]
        np.testing.assert_array_equal(res.query_idx, np.array(keep, dtype=np.int32), err_msg=str(ratio))
    # the integer entry point is the same test with ratio = num / den
    res = engine.hamming_match_pairs([(0, 1)], mode="knn2_ratio", ratio=(7, 10), return_best2=True)
    keep = [i for i in range(len(res.best2)) if res.best2[i, 2] >= 0 and float(res.best2[i, 1]) < 0.7 * float(res.best2[i, 3])]
    np.testing.assert_array_equal(res.query_idx, np.array(keep, dtype=np.int32))


@pytest.mark.parametrize("default_stream", [False, True])
def test_engine_work_is_ordered_with_the_callers_torch_stream(default_stream):
    """The bank is produced on the caller's stream by an asynchronous copy that sits behind 100 ms of
    other work; the engine, told to use that stream, must see the copied bytes.  torch names its
    default stream 0, which the C ABI reads as "own stream": Engine.set_stream maps it to
    cudaStreamLegacy (round 1 ran kernel 1 on the own stream without anybody noticing)."""
    import torch
    sets = synth.descriptor_bank(2, 3000, seed=77, world=4000)
    bank, offs = ssp_b200.Engine.pack_descriptors(sets)
    pinned = torch.from_numpy(bank).pin_memory()
    eng = ssp_b200.Engine(0)
    stream = torch.cuda.default_stream() if default_stream else torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        dev = torch.zeros_like(pinned, device="cuda")
        torch.cuda.synchronize()
        torch.cuda._sleep(200_000_000)                 # ~0.1 s of work in front of the copy
        dev.copy_(pinned, non_blocking=True)
        eng.bind_descriptor_bank(dev.data_ptr(), offs, keepalive=dev)
        res = eng.hamming_match_pairs([(0, 1)], mode="crosscheck")
    mq, mt, md = ho.crosscheck(sets[0], sets[1])
    np.testing.assert_array_equal(res.query_idx, mq)
    np.testing.assert_array_equal(res.train_idx, mt)
    assert len(mq) > 500
    eng.close()
