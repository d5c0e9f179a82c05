"""Host-side logic that needs no GPU: ratio terms, packing, sharding, generators, ordering."""
import os

import numpy as np
import pytest

import ssp_b200
from ssp_b200 import synthetic as synth
from ssp_b200.engine import Engine, _ratio_value
from ssp_b200.distributed import shard_pairs
from ssp_b200.custom_bf_polygon_matcher import disc_offsets


def test_ratio_value_is_the_float_the_reference_multiplies_with():
    assert _ratio_value(0.75) == 0.75
    assert _ratio_value((7, 10)) == 0.7
    assert _ratio_value(0.7000001) == 0.7000001   # no rational approximation any more
    with pytest.raises(ValueError):
        _ratio_value(0.0)
    with pytest.raises(ValueError):
        _ratio_value(float("nan"))


def test_pack_descriptors_handles_ragged_and_empty():
    a = np.arange(64, dtype=np.uint8).reshape(2, 32)
    bank, offs = Engine.pack_descriptors([a, np.zeros((0, 32), np.uint8), a[:1]])
    assert offs.tolist() == [0, 2, 2, 3]
    np.testing.assert_array_equal(bank[2], a[0])
    with pytest.raises(ValueError):
        Engine.pack_descriptors([np.zeros((2, 16), np.uint8)])


def test_shard_pairs_round_robin_covers_everything_once():
    pairs = synth.all_pairs(7)
    seen = []
    for r in range(4):
        local, pos = shard_pairs(pairs, r, 4)
        assert local == [pairs[i] for i in pos]
        seen += pos
    assert sorted(seen) == list(range(len(pairs)))


def test_descriptor_bank_plants_matches():
    bank = synth.descriptor_bank(3, 400, seed=0, world=600)
    assert all(b.shape == (400, 32) and b.dtype == np.uint8 for b in bank)
    x = np.unpackbits(bank[0][:, None, :] ^ bank[1][None, :, :], axis=2).sum(axis=2)
    best = x.min(axis=1)
    assert (best < 50).sum() > 100 and (best > 90).sum() > 50  # matched vs unmatched rows


def test_star_fields_shapes():
    xy, b, ids = synth.star_fields(4, 300, seed=2)
    for a, c, i in zip(xy, b, ids):
        assert a.shape[1] == 2 and len(a) == len(c) == len(i) and 150 < len(a) < 450
        assert a[:, 0].min() >= 0 and a[:, 0].max() < 1342 and a[:, 1].max() < 894


def test_disc_offsets_is_opencv_radius_10_disc():
    d = disc_offsets(10)
    assert d.shape == (317, 2) or 300 < d.shape[0] < 350
    assert (np.abs(d).max(axis=0) == 10).all()
    assert (d[:, 0] ** 2 + d[:, 1] ** 2 <= 110).all()


def test_strict_order_and_dummy_layout():
    cls = ssp_b200.CustomBruteForceMatcher
    fake = cls.__new__(cls)
    fake.imgs_features_orb_and_stars = [None] * 3
    assert fake.strict_order_for_pairwise_matches_list() == [
        (-1, -1), (0, 1), (0, 2), (1, 0), (-1, -1), (1, 2), (2, 0), (2, 1), (-1, -1)]


def test_bench_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` needs no GPU: one JSON line on stdout with the contract's keys."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--workload", "c1",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["impl"] == "reference" and d["value"] > 0 and d["steps"] == 2 and d["warmup"] == 1
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    assert "workload" in d["config"]
