"""Pair-sharded multi-GPU matching.

Image pairs are independent given read-only banks (custom_bruteforce_matcher.py:597, :615-620),
so the pair list is dealt round-robin to the ranks of one 8 x B200 box, one process per GPU.
The only exchanges are the two real ones the path has:

  * the descriptor bank is broadcast from rank 0 (NCCL over NVLink; 33.6 MB at C3, 128 MB at C5)
  * per-pair survivors are gathered back to rank 0 (variable length: counts first, then one
    padded gather)

There is no collective inside the matching itself.  `torch.distributed` is plumbing only;
with the gloo backend the same code runs on CPU tensors, which is how the sharding / gather
logic is tested without GPUs (tests/test_distributed.py injects the matcher function).
"""
from typing import Callable, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from .engine import HammingResult


def shard_pairs(pairs: Sequence, rank: int, world_size: int):
    """Round-robin shard of the pair list; returns (local pairs, their global positions)."""
    pos = list(range(rank, len(pairs), world_size))
    return [tuple(pairs[i]) for i in pos], pos


def _device_for_backend():
    if dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def broadcast_descriptor_bank(bank: Optional[np.ndarray], offsets: Optional[np.ndarray], src: int = 0):
    """Rank `src` passes the packed bank (uint8 (N, 32)) and row offsets; every rank gets a tensor
    on its own device (cuda for NCCL) plus the offsets as numpy."""
    dev = _device_for_backend()
    rank = dist.get_rank()
    meta = torch.zeros(2, dtype=torch.int64, device=dev)
    if rank == src:
        meta[0], meta[1] = bank.shape[0], len(offsets)
    dist.broadcast(meta, src)
    n_rows, n_off = int(meta[0]), int(meta[1])
    off_t = torch.zeros(n_off, dtype=torch.int64, device=dev)
    if rank == src:
        off_t.copy_(torch.from_numpy(np.ascontiguousarray(offsets, dtype=np.int64)))
        bank_t = torch.from_numpy(np.ascontiguousarray(bank)).to(dev, non_blocking=False)
    else:
        bank_t = torch.empty((n_rows, 32), dtype=torch.uint8, device=dev)
    dist.broadcast(off_t, src)
    dist.broadcast(bank_t, src)
    return bank_t, off_t.cpu().numpy()


def gather_matches(local: HammingResult, local_positions: Sequence[int], n_pairs: int, dst: int = 0):
    """Gathers the per-rank survivors to rank `dst` and restores global pair order.
    Returns a HammingResult on `dst`, None elsewhere."""
    dev = _device_for_backend()
    rank, world = dist.get_rank(), dist.get_world_size()
    # (1) per-pair counts, scattered into the global pair order and summed over ranks
    counts = torch.zeros(n_pairs, dtype=torch.int64, device=dev)
    if len(local_positions):
        local_counts = np.diff(local.offsets).astype(np.int64)
        counts[torch.as_tensor(list(local_positions), dtype=torch.int64, device=dev)] = \
            torch.from_numpy(local_counts).to(dev)
    dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    counts_np = counts.cpu().numpy()
    # (2) one padded gather of the packed (query, train, distance) triples
    per_rank = np.array([counts_np[r::world].sum() for r in range(world)], dtype=np.int64)
    width = int(per_rank.max()) if world else 0
    payload = torch.zeros((3, max(width, 1)), dtype=torch.int32, device=dev)
    n_local = int(local.offsets[-1])
    if n_local:
        packed = np.stack([local.query_idx, local.train_idx, local.distance]).astype(np.int32)
        payload[:, :n_local] = torch.from_numpy(packed).to(dev)
    bufs = [torch.empty_like(payload) for _ in range(world)] if rank == dst else None
    dist.gather(payload, bufs, dst=dst)
    if rank != dst:
        return None
    offsets = np.zeros(n_pairs + 1, dtype=np.int64)
    np.cumsum(counts_np, out=offsets[1:])
    total = int(offsets[-1])
    q = np.empty(total, dtype=np.int32)
    t = np.empty(total, dtype=np.int32)
    d = np.empty(total, dtype=np.int32)
    for r in range(world):
        data = bufs[r].cpu().numpy()
        cursor = 0
        for p in range(r, n_pairs, world):
            c = int(counts_np[p])
            lo = int(offsets[p])
            q[lo:lo + c] = data[0, cursor:cursor + c]
            t[lo:lo + c] = data[1, cursor:cursor + c]
            d[lo:lo + c] = data[2, cursor:cursor + c]
            cursor += c
    return HammingResult(offsets, q, t, d)


_PINNED = {}


def _pinned_i32(tag: str, n: int) -> torch.Tensor:
    """Grow-only pinned staging buffer (cudaHostAlloc per step would cost more than the copy)."""
    buf = _PINNED.get(tag)
    if buf is None or buf.numel() < n:
        buf = torch.empty(max(n, 1), dtype=torch.int32)
        if torch.cuda.is_available():
            buf = buf.pin_memory()
        _PINNED[tag] = buf
    return buf[:n]


def gather_matches_resident(engine, local_offsets: np.ndarray, local_positions: Sequence[int],
                            n_pairs: int, dst: int = 0):
    """`gather_matches` without the host round trip (NCCL backend): the rank's survivors of the
    last `hamming_match_pairs(..., fetch=False)` go device -> NCCL gather -> one permutation on
    rank `dst`'s GPU (round-robin shards back into global pair order) -> one copy into pinned
    host memory.  Returns a HammingResult on `dst`, None elsewhere; its arrays are views of a
    reused staging buffer, valid until the next call."""
    dev = _device_for_backend()
    rank, world = dist.get_rank(), dist.get_world_size()
    counts = torch.zeros(n_pairs, dtype=torch.int64, device=dev)
    if len(local_positions):
        counts[rank::world] = torch.from_numpy(np.diff(local_offsets).astype(np.int64)).to(dev)
    dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    per_rank = torch.stack([counts[r::world].sum() for r in range(world)])
    width = max(int(per_rank.max().item()), 1)
    payload = torch.empty((3, width), dtype=torch.int32, device=dev)
    n_local = int(local_offsets[-1])
    if n_local:
        engine.hamming_fetch_into(payload[0].data_ptr(), payload[1].data_ptr(), payload[2].data_ptr(),
                                  width, device=True)
    bufs = [torch.empty_like(payload) for _ in range(world)] if rank == dst else None
    dist.gather(payload, bufs, dst=dst)
    if rank != dst:
        return None
    offsets = torch.zeros(n_pairs + 1, dtype=torch.int64, device=dev)
    torch.cumsum(counts, 0, out=offsets[1:])
    total = int(offsets[-1].item())
    # where pair p's block starts inside the concatenated per-rank payloads
    base = torch.empty(n_pairs, dtype=torch.int64, device=dev)
    for r in range(world):
        c = counts[r::world]
        base[r::world] = r * width + torch.cumsum(c, 0) - c
    src = torch.repeat_interleave(base - offsets[:-1], counts, output_size=total) + \
        torch.arange(total, dtype=torch.int64, device=dev)
    flat = torch.cat(bufs, dim=1)                       # (3, world * width)
    ordered = flat.index_select(1, src)                 # (3, total) in global pair order
    host = _pinned_i32("gather", 3 * total).view(3, total)
    host.copy_(ordered, non_blocking=False)
    h = host.numpy()
    return HammingResult(offsets.cpu().numpy(), h[0], h[1], h[2])


def match_pairs_sharded(match_fn: Callable, bank: Optional[np.ndarray], offsets: Optional[np.ndarray],
                        pairs: Sequence, bind_fn: Optional[Callable] = None, src: int = 0):
    """End-to-end sharded run: broadcast bank -> every rank matches pairs[rank::world] -> gather.

    match_fn(local_pairs) -> HammingResult for the rank's shard (the engine's
    hamming_match_pairs in production); bind_fn(bank_tensor, offsets) hands the broadcast bank
    to the rank's engine (Engine.bind_descriptor_bank with the tensor's data_ptr)."""
    bank_t, off = broadcast_descriptor_bank(bank, offsets, src)
    if bind_fn is not None:
        bind_fn(bank_t, off)
    local_pairs, pos = shard_pairs(pairs, dist.get_rank(), dist.get_world_size())
    local = match_fn(local_pairs)
    return gather_matches(local, pos, len(pairs), dst=src)
