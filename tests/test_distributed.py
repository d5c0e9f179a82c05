"""world_size-2 gloo run of the pair-sharded path on CPU: broadcast bank -> shard -> gather must
reproduce the single-process result.  The matcher function is injected (the oracle stands in
for the engine here because there is no GPU; the production path passes the engine)."""
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


def _worker(rank, world, port, mode, out_path):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    from ssp_b200 import synthetic as synth
    from ssp_b200.engine import Engine
    from ssp_b200.distributed import match_pairs_sharded
    dist.init_process_group("gloo", rank=rank, world_size=world)
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


class _AddressEngine:
    """Stands in for Engine.hamming_fetch_into on CPU: copies a HammingResult to raw addresses."""

    def __init__(self, res):
        self.res = res

    def hamming_fetch_into(self, q_ptr, t_ptr, d_ptr, capacity, device=True):
        import ctypes
        n = int(self.res.offsets[-1])
        assert capacity >= n
        for ptr, arr in ((q_ptr, self.res.query_idx), (t_ptr, self.res.train_idx), (d_ptr, self.res.distance)):
            a = np.ascontiguousarray(arr, dtype=np.int32)
            ctypes.memmove(ptr, a.ctypes.data, n * 4)


def _worker_resident(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    from ssp_b200 import synthetic as synth
    from ssp_b200.engine import Engine
    from ssp_b200.distributed import shard_pairs, gather_matches_resident
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pairs = synth.all_pairs(6)
    sets = synth.descriptor_bank(6, 90, seed=5, world=150)
    sets[2] = sets[2][:41]
    bank, offs = Engine.pack_descriptors(sets)
    local_pairs, pos = shard_pairs(pairs, rank, world)
    local = _oracle_match_fn({"bank": bank, "offs": offs}, "knn2_ratio")(local_pairs)
    res = gather_matches_resident(_AddressEngine(local), local.offsets, pos, len(pairs), dst=0)
    if rank == 0:
        np.savez(out_path, offsets=res.offsets, q=res.query_idx, t=res.train_idx, d=res.distance)
    else:
        assert res is None
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_resident_gather_restores_global_pair_order(tmp_path, world):
    from ssp_b200 import synthetic as synth
    from ssp_b200.engine import Engine
    out = str(tmp_path / "res.npz")
    port = 31500 + (os.getpid() % 2000) + world
    mp.spawn(_worker_resident, args=(world, port, out), nprocs=world, join=True)
    got = np.load(out)
    sets = synth.descriptor_bank(6, 90, seed=5, world=150)
    sets[2] = sets[2][:41]
    bank, offs = Engine.pack_descriptors(sets)
    single = _oracle_match_fn({"bank": bank, "offs": offs}, "knn2_ratio")(synth.all_pairs(6))
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
