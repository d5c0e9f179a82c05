"""world_size-2/3 gloo runs of the pair-sharded paths on CPU: packed broadcast -> contiguous shards ->
exact-length gather must reproduce the single-process result, for the Hamming path and for the
star-polygon path (broadcast star bank, measure split by image + all-gather, gather matches).
The arithmetic is injected (the oracles stand in for the engine here because there is no GPU;
the production path passes the engine)."""
import ctypes
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle_match_fn(bank_holder, mode):
    from oracle import hamming_oracle as ho
    from ssp_b200.engine import HammingResult

    def run(local_pairs):
        bank, offs = bank_holder["bank"], bank_holder["offs"]
        qs, ts, ds, counts = [], [], [], []
        for a, b in local_pairs:
            q, t = bank[offs[a]:offs[a + 1]], bank[offs[b]:offs[b + 1]]
            if mode == "crosscheck":
                mq, mt, md = ho.crosscheck(q, t)
            else:
                mq, mt, md = ho.ratio_survivors(ho.knn2(q, t))
            qs.append(mq); ts.append(mt); ds.append(md); counts.append(len(mq))
        o = np.zeros(len(local_pairs) + 1, dtype=np.int64)
        np.cumsum(counts, out=o[1:])
        cat = lambda xs: np.concatenate(xs).astype(np.int32) if xs else np.zeros(0, np.int32)
        return HammingResult(o, cat(qs), cat(ts), cat(ds))
    return run


def _init(rank, world, port):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    return dist


def _worker(rank, world, port, mode, out_path):
    dist = _init(rank, world, port)
    from ssp_b200 import synthetic as synth
    from ssp_b200.engine import Engine
    from ssp_b200.distributed import match_pairs_sharded
    pairs = synth.all_pairs(5)
    bank = offs = None
    if rank == 0:
        sets = synth.descriptor_bank(5, 120, seed=3, world=200)
        sets[3] = sets[3][:77]  # ragged
        bank, offs = Engine.pack_descriptors(sets)
    holder = {}

    def bind(bank_t, off):
        holder["bank"], holder["offs"] = bank_t.numpy(), off

    res = match_pairs_sharded(_oracle_match_fn(holder, mode), bank, offs, pairs, bind_fn=bind)
    if rank == 0:
        np.savez(out_path, offsets=res.offsets, q=res.query_idx, t=res.train_idx, d=res.distance)
    else:
        assert res is None
    dist.destroy_process_group()


def _read(ptr, n, dtype):
    dt = np.dtype(dtype)
    if n == 0:
        return np.zeros(0, dt)
    return np.frombuffer((ctypes.c_char * (n * dt.itemsize)).from_address(ptr), dtype=dt, count=n).copy()


def _write(ptr, arr):
    a = np.ascontiguousarray(arr)
    if a.nbytes and ptr:
        ctypes.memmove(ptr, a.ctypes.data, a.nbytes)


class _AddressEngine:
    """Stands in for Engine.hamming_fetch_into on CPU: copies a HammingResult to raw addresses."""

    def __init__(self, res):
        self.res = res

    def hamming_fetch_into(self, q_ptr, t_ptr, d_ptr, capacity, device=True):
        n = int(self.res.offsets[-1])
        assert capacity >= n and not device    # CPU tensors under gloo
        for ptr, arr in ((q_ptr, self.res.query_idx), (t_ptr, self.res.train_idx), (d_ptr, self.res.distance)):
            _write(ptr, np.ascontiguousarray(arr, dtype=np.int32))


def _hamming_inputs():
    from ssp_b200 import synthetic as synth
    from ssp_b200.engine import Engine
    sets = synth.descriptor_bank(6, 90, seed=5, world=150)
    sets[2] = sets[2][:41]
    sets[4] = sets[4][:0]     # an image without descriptors: pairs with zero survivors
    bank, offs = Engine.pack_descriptors(sets)
    return synth.all_pairs(6), bank, offs


def _worker_resident(rank, world, port, out_path):
    dist = _init(rank, world, port)
    from ssp_b200.distributed import (broadcast_descriptor_bank, gather_matches_resident, shard_pairs_contiguous,
                                      _broadcaster)
    pairs, bank, offs = _hamming_inputs()
    collectives = []
    _broadcaster("descriptors", 0).CHUNK = 4096     # several pieces even for this small bank
    for step in range(3):     # the steady state is one pass over the buffer
        bank_t, off = broadcast_descriptor_bank(bank if rank == 0 else None, offs if rank == 0 else None, 0)
        collectives.append(_broadcaster("descriptors", 0).collectives)
        np.testing.assert_array_equal(off, offs)
        np.testing.assert_array_equal(bank_t.numpy(), bank)
    pieces = -(-_broadcaster("descriptors", 0).capacity // 4096)
    assert pieces > 2 and collectives[1] - collectives[0] == pieces and collectives[2] - collectives[1] == pieces
    cost = [(offs[a + 1] - offs[a]) * (offs[b + 1] - offs[b]) for a, b in pairs]
    local_pairs, lo, hi = shard_pairs_contiguous(pairs, rank, world, cost)
    local = _oracle_match_fn({"bank": bank_t.numpy(), "offs": off}, "knn2_ratio")(local_pairs)
    res = gather_matches_resident(_AddressEngine(local), local.offsets, dst=0)
    if rank == 0:
        np.savez(out_path, offsets=res.offsets, q=res.query_idx, t=res.train_idx, d=res.distance)
    else:
        assert res is None
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_packed_broadcast_and_exact_length_gather_restore_global_pair_order(tmp_path, world):
    out = str(tmp_path / "res.npz")
    port = 31500 + (os.getpid() % 2000) + world
    mp.spawn(_worker_resident, args=(world, port, out), nprocs=world, join=True)
    got = np.load(out)
    pairs, bank, offs = _hamming_inputs()
    single = _oracle_match_fn({"bank": bank, "offs": offs}, "knn2_ratio")(pairs)
    np.testing.assert_array_equal(got["offsets"], single.offsets)
    np.testing.assert_array_equal(got["q"], single.query_idx)
    np.testing.assert_array_equal(got["t"], single.train_idx)
    np.testing.assert_array_equal(got["d"], single.distance)
    assert single.offsets[-1] > 0


@pytest.mark.parametrize("mode", ["crosscheck", "knn2_ratio"])
def test_two_rank_gloo_equals_single_process(tmp_path, mode):
    from ssp_b200 import synthetic as synth
    from ssp_b200.engine import Engine
    out = str(tmp_path / "res.npz")
    port = 29500 + (os.getpid() % 2000) + (0 if mode == "crosscheck" else 1)
    mp.spawn(_worker, args=(2, port, mode, out), nprocs=2, join=True)
    got = np.load(out)
    sets = synth.descriptor_bank(5, 120, seed=3, world=200)
    sets[3] = sets[3][:77]
    bank, offs = Engine.pack_descriptors(sets)
    single = _oracle_match_fn({"bank": bank, "offs": offs}, mode)(synth.all_pairs(5))
    np.testing.assert_array_equal(got["offsets"], single.offsets)
    np.testing.assert_array_equal(got["q"], single.query_idx)
    np.testing.assert_array_equal(got["t"], single.train_idx)
    np.testing.assert_array_equal(got["d"], single.distance)
    assert single.offsets[-1] > 0


def test_shard_bounds_are_contiguous_cover_everything_and_balance_cost():
    from ssp_b200.distributed import shard_bounds, shard_pairs_contiguous
    assert shard_bounds(10, 4) == [0, 2, 5, 7, 10]
    cost = [100] + [1] * 99
    b = shard_bounds(100, 4, cost)
    assert b[0] == 0 and b[-1] == 100 and all(x <= y for x, y in zip(b, b[1:]))
    pairs = [(i, i + 1) for i in range(7)]
    seen = []
    for r in range(3):
        local, lo, hi = shard_pairs_contiguous(pairs, r, 3)
        assert local == pairs[lo:hi]
        seen += local
    assert seen == pairs
    assert shard_bounds(0, 3) == [0, 0, 0, 0]


# ---- star-polygon path: broadcast star bank, measure split by image, all-gather banks, gather matches ----
class _PolygonOracleEngine:
    """The engine's polygon entry points on CPU memory, arithmetic from oracle/polygon_oracle.py."""

    def __init__(self):
        from oracle.polygon_oracle import PolygonOracle
        self.P = PolygonOracle
        self.oracle = None
        self.bank = {}
        self.last = None

    def upload_star_bank_resident(self, xy_ptr, br_ptr, star_offsets, img_w, img_h, focal_px, owner=None):
        n = int(star_offsets[-1])
        self.xy = _read(xy_ptr, 2 * n, np.float64).reshape(n, 2)
        self.br = _read(br_ptr, n, np.float64)
        self.off = np.asarray(star_offsets)
        self.oracle = self.P((img_h, img_w, 3), focal_px)

    def polygon_measure_resident(self, image, device_ptr, n_candidates, min_side_px=5.0, wait=True):
        cands = _read(device_ptr, 5 * n_candidates, np.int32).reshape(-1, 5)
        lo, hi = int(self.off[image]), int(self.off[image + 1])
        got = self.oracle.characteristics(self.xy[lo:hi], self.br[lo:hi], cands, min_side_px)
        self.bank[image] = (np.ascontiguousarray(got["feat"], dtype=np.float64).reshape(-1, 15),
                            np.ascontiguousarray(got["star_idx"], dtype=np.int32).reshape(-1, 5))

    def polygon_measure_batch_resident(self, images, device_ptr, cand_offsets, min_side_px=5.0):
        for k, image in enumerate(images):
            lo, n = int(cand_offsets[k]), int(cand_offsets[k + 1] - cand_offsets[k])
            self.polygon_measure_resident(image, device_ptr + lo * 20 if n else 0, n, min_side_px)

    def polygon_count(self, image):
        return len(self.bank[image][0])

    def polygon_fetch_resident(self, image, star_idx_ptr, feat_ptr, capacity):
        feat, idx = self.bank[image]
        assert capacity >= len(feat)
        _write(star_idx_ptr, idx)
        _write(feat_ptr, feat)

    def polygon_upload_features_resident(self, image, feat_ptr, star_idx_ptr, n_polygons, owner=None):
        self.bank[image] = (_read(feat_ptr, 15 * n_polygons, np.float64).reshape(-1, 15),
                            _read(star_idx_ptr, 5 * n_polygons, np.int32).reshape(-1, 5))

    def polygon_match_pairs(self, pairs, threshold=0.1566, fetch=True):
        from ssp_b200.engine import PolygonMatchResult
        r_, c_, d_, counts = [], [], [], []
        for a, b in pairs:
            rows, cols, vals = self.P.select(self.P.compare(self.bank[a][0], self.bank[b][0]), threshold)
            r_.append(rows); c_.append(cols); d_.append(vals); counts.append(len(rows))
        o = np.zeros(len(pairs) + 1, dtype=np.int64)
        np.cumsum(counts, out=o[1:])
        cat = lambda xs, dt: np.concatenate(xs).astype(dt) if xs else np.zeros(0, dt)
        self.last = PolygonMatchResult(o, cat(r_, np.int32), cat(c_, np.int32), cat(d_, np.float64))
        return self.last

    def polygon_fetch_into(self, p1_ptr, p2_ptr, dist_ptr, capacity, device=True):
        assert capacity >= int(self.last.offsets[-1]) and not device
        _write(p1_ptr, self.last.poly1)
        _write(p2_ptr, self.last.poly2)
        _write(dist_ptr, self.last.dist)


def _polygon_inputs():
    from ssp_b200 import synthetic as synth
    xy, br, _ = synth.star_fields(5, 160, seed=4)
    cands = [synth.random_candidates(len(xy[i]), 250 + 40 * i, seed=30 + i) for i in range(5)]
    cands[3] = cands[3][:0]      # an image without candidates: no polygons, no matches
    return xy, br, cands, synth.all_pairs(5)


def _worker_polygons(rank, world, port, out_path):
    dist = _init(rank, world, port)
    from ssp_b200.distributed import (broadcast_star_bank, gather_polygon_matches_resident, measure_polygons_sharded,
                                      shard_pairs_contiguous)
    xy, br, cands, pairs = _polygon_inputs()
    src = (xy, br, cands) if rank == 0 else (None, None, None)
    so, co, xy_t, br_t, cand_t = broadcast_star_bank(*src, src=0)
    eng = _PolygonOracleEngine()
    # world 2: the exchange (split by image + all-gather); world 3: every rank measures everything
    counts = measure_polygons_sharded(eng, so, co, xy_t, br_t, cand_t, 1342, 894, 1135.0,
                                      replicate_below=0 if world == 2 else 1 << 20)
    assert [eng.polygon_count(i) for i in range(5)] == counts.tolist()   # every rank holds every bank now
    cost = [int(counts[a]) * int(counts[b]) for a, b in pairs]
    local_pairs, lo, hi = shard_pairs_contiguous(pairs, rank, world, cost)
    local = eng.polygon_match_pairs(local_pairs, threshold=0.6, fetch=False)
    res = gather_polygon_matches_resident(eng, local.offsets, dst=0)
    if rank == 0:
        np.savez(out_path, offsets=res.offsets, p1=res.poly1, p2=res.poly2, d=res.dist, counts=counts)
    else:
        assert res is None
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_polygon_path_sharded_equals_single_process(tmp_path, world):
    out = str(tmp_path / "poly.npz")
    port = 33500 + (os.getpid() % 2000) + world
    mp.spawn(_worker_polygons, args=(world, port, out), nprocs=world, join=True)
    got = np.load(out)
    xy, br, cands, pairs = _polygon_inputs()
    eng = _PolygonOracleEngine()
    so = np.zeros(6, dtype=np.int64)
    np.cumsum([len(a) for a in xy], out=so[1:])
    allxy = np.ascontiguousarray(np.concatenate(xy), dtype=np.float64)
    allbr = np.ascontiguousarray(np.concatenate(br), dtype=np.float64)
    eng.upload_star_bank_resident(allxy.ctypes.data, allbr.ctypes.data, so, 1342, 894, 1135.0)
    for i in range(5):
        c = np.ascontiguousarray(cands[i], dtype=np.int32)
        eng.polygon_measure_resident(i, c.ctypes.data if len(c) else 0, len(c))
    single = eng.polygon_match_pairs(pairs, threshold=0.6)
    assert got["counts"].tolist() == [eng.polygon_count(i) for i in range(5)] and got["counts"][3] == 0
    np.testing.assert_array_equal(got["offsets"], single.offsets)
    np.testing.assert_array_equal(got["p1"], single.poly1)
    np.testing.assert_array_equal(got["p2"], single.poly2)
    np.testing.assert_array_equal(got["d"], single.dist)
    assert single.offsets[-1] > 0
