"""Pair-sharded multi-GPU matching.

Image pairs are independent given read-only banks (custom_bruteforce_matcher.py:597, :615-620),
so the pair list is cut into one CONTIGUOUS, cost-balanced block per rank of one 8 x B200 box, one
process per GPU.  The only exchanges are the real ones the path has:

  * descriptor banks, star banks and candidate tables are broadcast from rank 0 as ONE packed
    buffer (header + small tables + payload; NCCL over NVLink);
  * the per-image polygon banks, measured by the rank that owns the image, are all-gathered;
  * per-pair results (Hamming survivors, polygon matches) go back to rank 0 with exact-length
    point-to-point transfers straight into their place in rank 0's final arrays -- contiguous
    shards make every rank's results one contiguous slice of the global pair order.

There is no collective inside the matching itself.  `torch.distributed` is plumbing only; with
the gloo backend the same code runs on CPU tensors, which is how this logic is tested without
GPUs (tests/test_distributed.py injects stand-in engines).
"""
from typing import Callable, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from .engine import HammingResult, PolygonMatchResult


# ---- sharding -------------------------------------------------------------------------------------
def shard_pairs(pairs: Sequence, rank: int, world_size: int):
    """Round-robin shard of the pair list; returns (local pairs, their global positions)."""
    pos = list(range(rank, len(pairs), world_size))
    return [tuple(pairs[i]) for i in pos], pos


def shard_bounds(n_items: int, world_size: int, cost: Optional[Sequence[float]] = None) -> List[int]:
    """Cuts [0, n_items) into world_size contiguous blocks of (nearly) equal total cost; returns the
    world_size + 1 boundaries.  cost[i] = work of item i (Nq * Nt for a Hamming pair, P1 * P2 for
    a polygon pair); None = all equal."""
    if cost is None:
        return [(n_items * r) // world_size for r in range(world_size + 1)]
    c = np.cumsum(np.asarray(cost, dtype=np.float64))
    total = float(c[-1]) if len(c) else 0.0
    bounds = [0]
    for r in range(1, world_size):
        bounds.append(max(bounds[-1], int(np.searchsorted(c, total * r / world_size, side="left"))))
    bounds.append(n_items)
    return bounds


def shard_pairs_contiguous(pairs: Sequence, rank: int, world_size: int, cost: Optional[Sequence[float]] = None):
    """This rank's contiguous block of the pair list: (local pairs, lo, hi)."""
    b = shard_bounds(len(pairs), world_size, cost)
    lo, hi = b[rank], b[rank + 1]
    return [tuple(p) for p in pairs[lo:hi]], lo, hi


def _device_for_backend():
    if dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


# ---- one packed broadcast -------------------------------------------------------------------------
_DTYPES = [np.dtype(np.uint8), np.dtype(np.int32), np.dtype(np.int64), np.dtype(np.float64), np.dtype(np.float32)]
_MAGIC = 0x53535042  # "SSPB"
_HEADER_SLOTS = 64    # arrays per message
_HEADER_BYTES = 8 * (4 + 6 * _HEADER_SLOTS)


class PackedBroadcaster:
    """Broadcasts a list of numpy arrays from `src` as ONE buffer per call.

    Layout: int64 header {magic, total bytes, n arrays, 0} + per array {dtype code, ndim, d0, d1,
    byte offset, host flag}, then the arrays, each 256-byte aligned.  Receivers do not know the
    size in advance, so every rank keeps a capacity: a call broadcasts `capacity` bytes; if the
    header says the message is longer, all ranks grow to that size and broadcast once more.  In
    the steady state (same sizes step after step, or shrinking) it is one pass over the buffer -- one
    collective per 8 MiB piece, all in flight, pipelined with rank src's host -> device copies -- and
    one small device -> host read of the header and of the arrays flagged `to_host` (offset tables)."""

    def __init__(self, src: int = 0, capacity: int = 1 << 16):
        self.src = src
        self.capacity = max(int(capacity), _HEADER_BYTES)
        self.dev = _device_for_backend()
        self.buf = torch.empty(self.capacity, dtype=torch.uint8, device=self.dev)
        self._stage = None   # pinned staging for the header and the small host-side tables
        self.collectives = 0

    CHUNK = 8 << 20   # a large message goes out in pieces so that rank src's H2D of piece k + 1 overlaps the
                      # NVLink broadcast of piece k (33.6 MB bank at C3: 0.6 ms of PCIe + 0.3 ms of NVLink otherwise)

    def _grow(self, nbytes):
        self.capacity = (int(nbytes) + 4095) // 4096 * 4096     # tight: every call broadcasts the whole capacity
        self.buf = torch.empty(self.capacity, dtype=torch.uint8, device=self.dev)

    def _layout(self, arrays, to_host):
        head = np.zeros(4 + 6 * _HEADER_SLOTS, dtype=np.int64)
        cursor = _HEADER_BYTES
        for k, a in enumerate(arrays):
            cursor = (cursor + 255) // 256 * 256
            shape = list(a.shape) + [1, 1]
            head[4 + 6 * k: 10 + 6 * k] = [_DTYPES.index(a.dtype), a.ndim, shape[0], shape[1], cursor,
                                           1 if k in to_host else 0]
            cursor += a.nbytes
        head[0:4] = [_MAGIC, cursor, len(arrays), 0]
        return head, int(cursor)

    def _broadcast_buffer(self, fill=None):
        """Broadcasts the whole capacity in CHUNK pieces, all in flight; `fill(lo, hi)` (rank src) puts
        the bytes [lo, hi) of the message into the buffer right before their piece goes out."""
        works = []
        for lo in range(0, self.capacity, self.CHUNK):
            hi = min(lo + self.CHUNK, self.capacity)
            if fill is not None:
                fill(lo, hi)
            works.append(dist.broadcast(self.buf[lo:hi], self.src, async_op=True))
            self.collectives += 1
        for w in works:
            w.wait()

    def broadcast(self, arrays: Optional[Sequence[np.ndarray]], to_host: Sequence[int] = ()):
        """`arrays` on rank src (1- or 2-D, C-contiguous; pass page-locked memory for large ones),
        None elsewhere.  Returns a list with, per array, a numpy array (indices in `to_host`) or a
        tensor view of the receive buffer on this rank's device (valid until the next call)."""
        rank = dist.get_rank()
        if rank == self.src:
            arrays = [np.ascontiguousarray(a) for a in arrays]
            assert len(arrays) <= _HEADER_SLOTS and all(a.ndim in (1, 2) for a in arrays)
            head, total = self._layout(arrays, set(to_host))
            if total > self.capacity:
                # receivers cannot know yet: send the header alone in the old capacity, then grow
                self.buf[:_HEADER_BYTES].copy_(torch.from_numpy(head.view(np.uint8)))
                self._broadcast_buffer()
                self._grow(total)
            spans = [(0, torch.from_numpy(head.view(np.uint8)))]
            for k, a in enumerate(arrays):
                if a.nbytes:
                    spans.append((int(head[4 + 6 * k + 4]), torch.from_numpy(a.reshape(-1).view(np.uint8))))

            def fill(lo, hi):
                for off, src_bytes in spans:
                    a_lo, a_hi = max(lo, off), min(hi, off + src_bytes.numel())
                    if a_lo < a_hi:
                        self.buf[a_lo:a_hi].copy_(src_bytes[a_lo - off:a_hi - off], non_blocking=True)
            self._broadcast_buffer(fill)
        else:
            self._broadcast_buffer()
            head = self.buf[:_HEADER_BYTES].cpu().numpy().view(np.int64)
            assert head[0] == _MAGIC, "packed broadcast: bad header"
            if int(head[1]) > self.capacity:
                self._grow(int(head[1]))
                self._broadcast_buffer()
                head = self.buf[:_HEADER_BYTES].cpu().numpy().view(np.int64)
        out = []
        host_spans = []
        for k in range(int(head[2])):
            code, ndim, d0, d1, off, host = (int(v) for v in head[4 + 6 * k: 10 + 6 * k])
            dt = _DTYPES[code]
            shape = (d0,) if ndim == 1 else (d0, d1)
            nbytes = int(np.prod(shape)) * dt.itemsize
            view = self.buf[off:off + nbytes]
            if host:
                if rank == self.src:
                    out.append(arrays[k])
                else:
                    host_spans.append((len(out), view, dt, shape))
                    out.append(None)
            else:
                out.append(view.view(_TORCH_DTYPES[code]).view(shape))
        for idx, view, dt, shape in host_spans:
            out[idx] = view.cpu().numpy().view(dt).reshape(shape)
        return out


_TORCH_DTYPES = [torch.uint8, torch.int32, torch.int64, torch.float64, torch.float32]
_BROADCASTERS = {}


def _broadcaster(tag: str, src: int) -> PackedBroadcaster:
    key = (tag, src, dist.get_backend())
    b = _BROADCASTERS.get(key)
    if b is None or b.dev != _device_for_backend():
        b = _BROADCASTERS[key] = PackedBroadcaster(src)
    return b


def broadcast_descriptor_bank(bank: Optional[np.ndarray], offsets: Optional[np.ndarray], src: int = 0):
    """Rank `src` passes the packed bank (uint8 (N, 32)) and row offsets; every rank gets a tensor
    on its own device (cuda for NCCL; a view of a reused receive buffer) plus the offsets as
    numpy.  One collective per call in the steady state (PackedBroadcaster)."""
    arrays = None
    if dist.get_rank() == src:
        arrays = [np.ascontiguousarray(offsets, dtype=np.int64), np.ascontiguousarray(bank, dtype=np.uint8).reshape(-1, 32)]
    off, bank_t = _broadcaster("descriptors", src).broadcast(arrays, to_host=(0,))
    return bank_t, np.asarray(off, dtype=np.int64)


# ---- exact-length gather of per-pair results -------------------------------------------------------
_PINNED = {}
_DEVBUF = {}


def _pinned(tag: str, n: int, dtype: torch.dtype) -> torch.Tensor:
    """Grow-only pinned staging buffer (cudaHostAlloc per step would cost more than the copy)."""
    buf = _PINNED.get((tag, dtype))
    if buf is None or buf.numel() < n:
        buf = torch.empty(max(n + n // 8, 1), dtype=dtype)
        if torch.cuda.is_available():
            buf = buf.pin_memory()
        _PINNED[(tag, dtype)] = buf
    return buf[:n]


def _devbuf(tag: str, n: int, dtype: torch.dtype, dev) -> torch.Tensor:
    buf = _DEVBUF.get((tag, dtype, str(dev)))
    if buf is None or buf.numel() < n:
        buf = torch.empty(max(n + n // 8, 1), dtype=dtype, device=dev)
        _DEVBUF[(tag, dtype, str(dev))] = buf
    return buf[:n]


def gather_segments(tag: str, fetch_into: Callable, dtypes: Sequence[torch.dtype], local_counts: np.ndarray,
                    dst: int = 0, max_block: Optional[int] = None):
    """Every rank holds, for its CONTIGUOUS block of pairs, `len(dtypes)` result arrays with
    local_counts[p] entries for its p-th pair.  Brings them to rank `dst` in global pair order:

      1. one all-gather of the per-pair counts (ranks own blocks of different length: padded);
      2. rank dst sizes its final device arrays; every other rank sends its arrays with their exact
         length straight into its slice of them (one batch of point-to-point transfers; dst's own
         block is written there by `fetch_into`);
      3. one device -> pinned-host copy per array on dst.

    fetch_into(list of device tensors of length sum(local_counts)) must fill them (the engine's
    *_fetch_* entry points with SSP_MEM_DEVICE).  Returns (offsets int64 (n_pairs + 1), [numpy
    arrays]) on dst -- views of reused staging buffers, valid until the next call -- else None."""
    dev = _device_for_backend()
    rank, world = dist.get_rank(), dist.get_world_size()
    local_counts = np.asarray(local_counts, dtype=np.int64)
    n_local = int(local_counts.sum())
    # (1) counts: [n_pairs_of_rank, counts..., 0 padding]; block lengths differ by the cost balance.
    # max_block = the longest block of any rank, when the caller knows it (every rank can compute the
    # shard bounds itself): saves one collective and one host synchronisation per step
    if max_block is None:
        width = torch.tensor([len(local_counts)], dtype=torch.int64, device=dev)
        dist.all_reduce(width, op=dist.ReduceOp.MAX)
        w = int(width.item())
    else:
        w = int(max_block)
        assert w >= len(local_counts)
    mine = torch.zeros(w + 1, dtype=torch.int64)
    mine[0] = len(local_counts)
    mine[1:1 + len(local_counts)] = torch.from_numpy(local_counts)
    table = torch.empty(world * (w + 1), dtype=torch.int64, device=dev)   # flat: gloo accepts nothing else
    dist.all_gather_into_tensor(table, mine.to(dev))
    table = table.cpu().numpy().reshape(world, w + 1)
    per_rank = [table[r, 1:1 + int(table[r, 0])] for r in range(world)]
    totals = np.array([int(c.sum()) for c in per_rank], dtype=np.int64)
    base = np.concatenate([[0], np.cumsum(totals)])
    # (2) exact-length transfers into dst's final arrays
    if rank == dst:
        total = int(base[-1])
        final = [_devbuf(f"{tag}{k}", total, dt, dev) for k, dt in enumerate(dtypes)]
        ops = []
        for r in range(world):
            if r == dst or totals[r] == 0:
                continue
            for k in range(len(dtypes)):
                ops.append(dist.P2POp(dist.irecv, final[k][int(base[r]):int(base[r + 1])], r))
        if n_local:
            fetch_into([f[int(base[dst]):int(base[dst + 1])] for f in final])
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        # (3) final arrays -> pinned host
        host = []
        for k, dt in enumerate(dtypes):
            h = _pinned(f"{tag}{k}", total, dt)
            h.copy_(final[k], non_blocking=True)     # all arrays in flight, one synchronisation below
            host.append(h.numpy())
        if dev.type == "cuda":
            torch.cuda.current_stream(dev).synchronize()
        offsets = np.zeros(sum(len(c) for c in per_rank) + 1, dtype=np.int64)
        np.cumsum(np.concatenate(per_rank) if per_rank else np.zeros(0, np.int64), out=offsets[1:])
        return offsets, host
    if n_local:
        send = [_devbuf(f"{tag}s{k}", n_local, dt, dev) for k, dt in enumerate(dtypes)]
        fetch_into(send)
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, s, dst) for s in send]):
            req.wait()
    return None


def gather_matches_resident(engine, local_offsets: np.ndarray, dst: int = 0, max_block: Optional[int] = None):
    """Hamming survivors of the last `hamming_match_pairs(local block, fetch=False)` of every rank
    -> HammingResult in global pair order on rank `dst` (None elsewhere), without a host round
    trip on the sending side."""
    def fetch(bufs):
        engine.hamming_fetch_into(bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[2].data_ptr(), bufs[0].numel(),
                                  device=bufs[0].is_cuda)
    out = gather_segments("ham", fetch, (torch.int32, torch.int32, torch.int32), np.diff(local_offsets), dst, max_block)
    if out is None:
        return None
    offsets, (q, t, d) = out
    return HammingResult(offsets, q, t, d)


def gather_polygon_matches_resident(engine, local_offsets: np.ndarray, dst: int = 0, max_block: Optional[int] = None):
    """Polygon matches (poly1, poly2, dist) of every rank's block -> PolygonMatchResult on `dst`."""
    def fetch(bufs):
        engine.polygon_fetch_into(bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[2].data_ptr(), bufs[0].numel(),
                                  device=bufs[0].is_cuda)
    out = gather_segments("poly", fetch, (torch.int32, torch.int32, torch.float64), np.diff(local_offsets), dst, max_block)
    if out is None:
        return None
    offsets, (p1, p2, d) = out
    return PolygonMatchResult(offsets, p1, p2, d)


# ---- star / polygon path on several GPUs ------------------------------------------------------------
def broadcast_star_bank(xy: Optional[Sequence[np.ndarray]], brightness: Optional[Sequence[np.ndarray]],
                        candidates: Optional[Sequence[np.ndarray]], src: int = 0):
    """Rank `src` passes per-image star coordinates (n_i, 2), brightness (n_i,) and candidate
    5-tuples (c_i, 5); every rank gets (star_offsets, cand_offsets) as numpy and (xy float64 (N, 2),
    brightness float64 (N,), candidates int32 (C, 5)) as tensors on its device.  One collective."""
    arrays = None
    if dist.get_rank() == src:
        so = np.zeros(len(xy) + 1, dtype=np.int64)
        co = np.zeros(len(xy) + 1, dtype=np.int64)
        np.cumsum([len(a) for a in xy], out=so[1:])
        np.cumsum([len(c) for c in candidates], out=co[1:])
        cat = lambda parts, dt, cols: (np.concatenate([np.asarray(p, dtype=dt).reshape(-1, cols) for p in parts])
                                       if len(parts) else np.zeros((0, cols), dt))
        arrays = [so, co, cat(xy, np.float64, 2), cat(brightness, np.float64, 1).reshape(-1),
                  cat(candidates, np.int32, 5)]
    so, co, xy_t, br_t, cand_t = _broadcaster("stars", src).broadcast(arrays, to_host=(0, 1))
    return np.asarray(so), np.asarray(co), xy_t, br_t, cand_t


def measure_polygons_sharded(engine, star_off, cand_off, xy_t, br_t, cand_t, img_w, img_h, focal,
                             min_side_px: float = 5.0, replicate_below: int = 1 << 20):
    """Kernel 2 split by image (the reference's per-image cache, custom_bruteforce_matcher.py:57-59 /
    cbpm:744-748: measure once per image, compare per pair): every rank measures ONE contiguous block
    of images (balanced by candidate count) in one launch pair, then the polygon banks are all-gathered
    so that every rank can compare any pair.  Returns the polygon count of every image.

    With fewer than `replicate_below` candidates in total (a C4 step has 0.2 M: 40 us of kernel 2) every
    rank simply measures ALL images in its one launch pair: the exchange -- three collectives, two host
    synchronisations and a device copy per foreign image -- would cost more than it saves."""
    dev = _device_for_backend()
    rank, world = dist.get_rank(), dist.get_world_size()
    n_img = len(star_off) - 1
    engine.upload_star_bank_resident(xy_t.data_ptr(), br_t.data_ptr(), star_off, img_w, img_h, focal)
    if int(cand_off[-1]) < replicate_below:
        engine.polygon_measure_batch_resident(list(range(n_img)), cand_t.data_ptr(), np.asarray(cand_off), min_side_px)
        return np.array([engine.polygon_count(i) for i in range(n_img)], dtype=np.int64)
    bounds = shard_bounds(n_img, world, cost=np.diff(cand_off).astype(np.float64) + 1.0)
    lo, hi = bounds[rank], bounds[rank + 1]
    mine = list(range(lo, hi))
    if mine:
        engine.polygon_measure_batch_resident(mine, cand_t[int(cand_off[lo]):].data_ptr(),
                                              np.asarray(cand_off[lo:hi + 1]) - int(cand_off[lo]), min_side_px)
    counts = torch.zeros(n_img, dtype=torch.int64, device=dev)
    if mine:
        counts[lo:hi] = torch.tensor([engine.polygon_count(i) for i in mine], dtype=torch.int64).to(dev)
    dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    counts = counts.cpu().numpy()
    # every rank's banks back to back, padded to the longest rank, one all-gather per array
    rows_of = [int(counts[bounds[r]:bounds[r + 1]].sum()) for r in range(world)]
    width = max(max(rows_of), 1)
    feat = _devbuf("pg_feat", width * 15, torch.float64, dev).view(width, 15)
    idx = _devbuf("pg_idx", width * 5, torch.int32, dev).view(width, 5)
    cursor = 0
    for i in mine:
        n = int(counts[i])
        if n:
            engine.polygon_fetch_resident(i, idx[cursor:].data_ptr(), feat[cursor:].data_ptr(), n)
        cursor += n
    all_feat = _devbuf("pg_all_feat", world * width * 15, torch.float64, dev).view(world, width, 15)
    all_idx = _devbuf("pg_all_idx", world * width * 5, torch.int32, dev).view(world, width, 5)
    dist.all_gather_into_tensor(all_feat.view(-1), feat.view(-1))
    dist.all_gather_into_tensor(all_idx.view(-1), idx.view(-1))
    for r in range(world):
        if r == rank:
            continue
        cursor = 0
        for i in range(bounds[r], bounds[r + 1]):
            n = int(counts[i])
            engine.polygon_upload_features_resident(i, all_feat[r, cursor:].data_ptr(), all_idx[r, cursor:].data_ptr(), n)
            cursor += n
    return counts


# ---- host-staged forms (gloo / CPU tests of the same logic, and callers without device buffers) -----
def gather_matches(local: HammingResult, local_positions: Sequence[int], n_pairs: int, dst: int = 0):
    """Gathers the per-rank survivors of a ROUND-ROBIN shard to rank `dst` through host memory and
    restores global pair order.  Returns a HammingResult on `dst`, None elsewhere."""
    dev = _device_for_backend()
    rank, world = dist.get_rank(), dist.get_world_size()
    counts = torch.zeros(n_pairs, dtype=torch.int64, device=dev)
    if len(local_positions):
        local_counts = np.diff(local.offsets).astype(np.int64)
        counts[torch.as_tensor(list(local_positions), dtype=torch.int64, device=dev)] = \
            torch.from_numpy(local_counts).to(dev)
    dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    counts_np = counts.cpu().numpy()
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
