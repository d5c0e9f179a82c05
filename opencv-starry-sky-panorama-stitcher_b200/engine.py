"""Thin Python face of the C ABI: numpy in, numpy out, device work in libssp_b200.so.

PyTorch is optional plumbing here (device buffers and the stream the kernels run on); every
computation happens in the hand-written CUDA kernels behind `include/ssp_b200.h`.
"""
import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import _native as N


class EngineError(RuntimeError):
    """Any failure reported by libssp_b200.so other than the reference's own ValueError case."""


@dataclass
class HammingResult:
    """Survivors of `hamming_match_pairs`: pair p owns [offsets[p], offsets[p+1])."""
    offsets: np.ndarray      # int64 (n_pairs + 1)
    query_idx: np.ndarray    # int32 (n_matches)  ascending inside each pair
    train_idx: np.ndarray    # int32 (n_matches)
    distance: np.ndarray     # int32 (n_matches)  Hamming distance (DMatch.distance as integer)
    best2: Optional[np.ndarray] = None       # int32 (rows, 4) {idx1, d1, idx2, d2}
    best2_offsets: Optional[np.ndarray] = None  # int64 (n_pairs + 1) row ranges into best2

    def pair(self, p):
        lo, hi = int(self.offsets[p]), int(self.offsets[p + 1])
        return self.query_idx[lo:hi], self.train_idx[lo:hi], self.distance[lo:hi]


@dataclass
class PolygonMatchResult:
    """Polygon matches of `polygon_match_pairs` (cbpm:1145-1152 `valid_polygon_matches_*`)."""
    offsets: np.ndarray   # int64 (n_pairs + 1)
    poly1: np.ndarray     # int32 row index in the first image's polygon bank
    poly2: np.ndarray     # int32 argmin row in the second image's bank
    dist: np.ndarray      # float64 polygon distance
    argmin: Optional[np.ndarray] = None       # int32 per row of every pair (before threshold)
    mindist: Optional[np.ndarray] = None      # float64
    row_offsets: Optional[np.ndarray] = None  # int64 (n_pairs + 1)

    def pair(self, p):
        lo, hi = int(self.offsets[p]), int(self.offsets[p + 1])
        return self.poly1[lo:hi], self.poly2[lo:hi], self.dist[lo:hi]


_MODES = {"knn2_ratio": N.MODE_KNN2_RATIO, "crosscheck": N.MODE_CROSSCHECK}
_VARIANTS = {"auto": N.HAMMING_AUTO, "popc": N.HAMMING_POPC, "popc_csa": N.HAMMING_POPC_CSA,
             "tensor": N.HAMMING_TENSOR}


def _ratio_value(ratio):
    """Lowe ratio as the float64 the reference multiplies with (`m.distance < 0.75 * n.distance`,
    custom_bruteforce_matcher.py:164); a (num, den) pair means num / den.  The device evaluates
    float64(d1) < ratio * float64(d2), which IS Python's comparison for any ratio."""
    if isinstance(ratio, (tuple, list)):
        num, den = int(ratio[0]), int(ratio[1])
        if num <= 0 or den <= 0:
            raise ValueError("ratio must be positive")
        return num / den
    r = float(ratio)
    if not r > 0.0:
        raise ValueError("ratio must be positive")
    return r


def _ptr(a):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


def _as_pairs(pairs):
    arr = np.ascontiguousarray(np.asarray(pairs, dtype=np.int32).reshape(-1, 2))
    return arr


class Engine:
    """One engine context per matcher instance and CUDA device."""

    def __init__(self, device: int = 0, stream: Optional[int] = None):
        self._lib = N.lib()
        self._ctx = ctypes.c_void_p()
        rc = self._lib.ssp_create(ctypes.byref(self._ctx), int(device))
        if rc != N.OK:
            msg = self._lib.ssp_last_error(None).decode()
            self._ctx = None
            raise EngineError(f"ssp_create failed ({rc}): {msg}")
        self.device = int(device)
        self._desc_offsets = None
        self._star_offsets = None
        self._bound = None  # keeps a bound torch tensor alive
        # Who owns what on this context.  Host mirrors that cache "my bank is uploaded" compare these
        # with what they remember before every call and upload again on a mismatch, so that two
        # matchers (or a matcher and the BFMatcher drop-in) can share one context safely.
        self.bank_generation = 0   # bumped by every descriptor-bank upload / bind
        self.star_owner = None     # whoever uploaded the current star bank
        self.poly_owner = {}       # image id -> whoever measured / uploaded that polygon bank
        self._pinned = {}   # tag -> (address, bytes): engine-owned page-locked result buffers
        if stream is not None:
            self.set_stream(stream)

    # -- lifetime -------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None):
            for addr, _ in getattr(self, "_pinned", {}).values():
                self._lib.ssp_host_free(self._ctx, ctypes.c_void_p(addr))
            self._pinned = {}
            self._lib.ssp_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc == N.OK:
            return
        msg = self._lib.ssp_last_error(self._ctx).decode()
        if rc == N.ERR_DOMAIN:
            # the reference's math.acos raises ValueError("math domain error") (cbpm:134)
            raise ValueError(f"math domain error: {msg}")
        raise EngineError(f"libssp_b200 error {rc}: {msg}")

    def pinned_array(self, tag: str, n: int, dtype) -> np.ndarray:
        """numpy view of `n` items of an engine-owned page-locked buffer (grow-only, one per
        `tag`); the view is valid until the next request for the same tag or `close()`."""
        dt = np.dtype(dtype)
        need = max(int(n) * dt.itemsize, 1)
        addr, size = self._pinned.get(tag, (0, 0))
        if size < need:
            if addr:
                self._check(self._lib.ssp_host_free(self._ctx, ctypes.c_void_p(addr)))
                self._pinned.pop(tag)
            p = ctypes.c_void_p()
            size = need + need // 4
            self._check(self._lib.ssp_host_alloc(self._ctx, size, ctypes.byref(p)))
            addr = p.value
            self._pinned[tag] = (addr, size)
        raw = (ctypes.c_char * need).from_address(addr)
        return np.frombuffer(raw, dtype=dt, count=int(n))

    CUDA_STREAM_LEGACY = 0x1   # cudaStreamLegacy: the handle that NAMES the default stream

    def set_stream(self, cuda_stream_handle: Optional[int]):
        """Run subsequent work on this cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream).
        None = the context's own (non-blocking) stream.  torch reports its DEFAULT stream as handle 0,
        which the C ABI reads as "own stream" too: 0 is therefore passed on as cudaStreamLegacy, so
        that the engine's kernels really are ordered with the caller's torch / NCCL work (round 1 and
        the first round-2 bench silently ran on the own stream: an NCCL broadcast enqueued before
        kernel 1 then sat behind the persistent kernel that holds every SM)."""
        if cuda_stream_handle is None:
            handle = 0
        else:
            handle = int(cuda_stream_handle) or self.CUDA_STREAM_LEGACY
        self._check(self._lib.ssp_set_stream(self._ctx, ctypes.c_void_p(handle)))

    def synchronize(self):
        self._check(self._lib.ssp_synchronize(self._ctx))

    def device_info(self):
        sm, khz, mem = ctypes.c_int(), ctypes.c_int(), ctypes.c_int64()
        self._check(self._lib.ssp_device_info(self._ctx, ctypes.byref(sm), ctypes.byref(khz),
                                              ctypes.byref(mem)))
        return {"sm_count": sm.value, "sm_clock_khz": khz.value, "hbm_bytes": mem.value}

    # -- descriptor bank ------------------------------------------------------------------
    @staticmethod
    def pack_descriptors(descriptors: Sequence[np.ndarray]):
        """list of uint8 (n_i, 32) -> (uint8 (sum n_i, 32) C-contiguous, int64 offsets)."""
        offs = np.zeros(len(descriptors) + 1, dtype=np.int64)
        for i, d in enumerate(descriptors):
            d = np.asarray(d)
            if d.size and (d.dtype != np.uint8 or d.ndim != 2 or d.shape[1] != 32):
                raise ValueError(f"descriptor set {i}: expected uint8 (n, 32), got {d.dtype} {d.shape}")
            offs[i + 1] = offs[i] + (d.shape[0] if d.size else 0)
        bank = np.empty((int(offs[-1]), 32), dtype=np.uint8)
        for i, d in enumerate(descriptors):
            if offs[i + 1] > offs[i]:
                bank[offs[i]:offs[i + 1]] = d
        return bank, offs

    def upload_descriptor_bank(self, descriptors, row_offsets=None):
        """`descriptors`: list of uint8 (n_i, 32) arrays, or one packed (N, 32) array together
        with `row_offsets`.  Copies host -> device."""
        if row_offsets is None:
            bank, offs = self.pack_descriptors(descriptors)
        else:
            bank = np.ascontiguousarray(descriptors, dtype=np.uint8).reshape(-1, 32)
            offs = np.ascontiguousarray(row_offsets, dtype=np.int64)
            if offs[-1] != bank.shape[0]:
                raise ValueError("row_offsets[-1] must equal the number of descriptor rows")
        self._check(self._lib.ssp_upload_descriptor_bank(self._ctx, _ptr(bank), _ptr(offs),
                                                         len(offs) - 1, N.MEM_HOST))
        self.synchronize()  # `bank` may be a temporary
        self._desc_offsets = offs.copy()
        self._bound = None
        self.bank_generation += 1
        return self.bank_generation

    def bind_descriptor_bank(self, device_ptr: int, row_offsets, keepalive=None):
        """Zero-copy: the bank already lives on this device (e.g. a torch uint8 tensor);
        `keepalive` is held so the buffer outlives the binding."""
        offs = np.ascontiguousarray(row_offsets, dtype=np.int64)
        self._check(self._lib.ssp_bind_descriptor_bank(self._ctx, ctypes.c_void_p(device_ptr),
                                                       _ptr(offs), len(offs) - 1))
        self._desc_offsets = offs.copy()
        self._bound = keepalive
        self.bank_generation += 1
        return self.bank_generation

    # -- kernel 1 -------------------------------------------------------------------------
    def hamming_match_pairs(self, pairs, mode="crosscheck", ratio=0.75, variant="auto",
                            return_best2=False, fetch=True, pinned=False) -> HammingResult:
        """All-pairs Hamming matching of image pairs (query image, train image).

        mode="crosscheck"  == cv.BFMatcher(cv.NORM_HAMMING, crossCheck=True).match per pair
                              (custom_bruteforce_matcher.py:193-199)
        mode="knn2_ratio"  == knnMatch(k=2) + `m.distance < ratio * n.distance`
                              (custom_bruteforce_matcher.py:160-172)
        pinned=True returns views of engine-owned page-locked buffers (valid until the next
        call) instead of fresh numpy arrays: the device -> host copy then runs at link rate.
        """
        if self._desc_offsets is None:
            raise EngineError("no descriptor bank uploaded")
        p = _as_pairs(pairs)
        offs = np.zeros(p.shape[0] + 1, dtype=np.int64)
        self._check(self._lib.ssp_hamming_match_pairs_f64(
            self._ctx, _ptr(p), p.shape[0], _MODES[mode], _ratio_value(ratio), _VARIANTS[variant], _ptr(offs)))
        n = int(offs[-1])
        if pinned and fetch:
            q, t, d = (self.pinned_array(k, n, np.int32) for k in ("hq", "ht", "hd"))
        else:
            q = np.empty(n, dtype=np.int32)
            t = np.empty(n, dtype=np.int32)
            d = np.empty(n, dtype=np.int32)
        if fetch:
            self._check(self._lib.ssp_hamming_fetch_matches(self._ctx, _ptr(q), _ptr(t), _ptr(d), n,
                                                            N.MEM_HOST))
        res = HammingResult(offs, q, t, d)
        if return_best2:
            nq = (self._desc_offsets[1:] - self._desc_offsets[:-1])[p[:, 0]] if p.size else \
                np.zeros(0, dtype=np.int64)
            b_off = np.zeros(p.shape[0] + 1, dtype=np.int64)
            np.cumsum(nq, out=b_off[1:])
            b2 = np.empty((int(b_off[-1]), 4), dtype=np.int32)
            self._check(self._lib.ssp_hamming_fetch_best2(self._ctx, _ptr(b2), b2.shape[0], N.MEM_HOST))
            res.best2, res.best2_offsets = b2, b_off
        return res

    def hamming_fetch_into(self, q_ptr: int, t_ptr: int, d_ptr: int, capacity: int, device: bool = True):
        """Copies the survivors of the last `hamming_match_pairs(..., fetch=False)` into three
        int32 buffers given by address (device memory, or pinned / pageable host memory)."""
        self._check(self._lib.ssp_hamming_fetch_matches(
            self._ctx, ctypes.c_void_p(q_ptr), ctypes.c_void_p(t_ptr), ctypes.c_void_p(d_ptr),
            int(capacity), N.MEM_DEVICE if device else N.MEM_HOST))

    def hamming_last_stats(self):
        n, ms, cmp_ = ctypes.c_int(), ctypes.c_float(), ctypes.c_int64()
        self._check(self._lib.ssp_hamming_last_stats(self._ctx, ctypes.byref(n), ctypes.byref(ms),
                                                     ctypes.byref(cmp_)))
        return {"launches": n.value, "main_kernel_ms": ms.value, "comparisons": cmp_.value}

    # -- kernel 2 -------------------------------------------------------------------------
    def upload_star_bank(self, xy: Sequence[np.ndarray], brightness: Sequence[np.ndarray],
                         img_w: int, img_h: int, focal_px: float, owner=None):
        """`xy[i]`: (n_i, 2) pixel coordinates of image i's stars (keypoints[n_orb:], cbpm:710);
        `brightness[i]`: (n_i,) get_star_brightness values.  img_w/img_h = work image size."""
        offs = np.zeros(len(xy) + 1, dtype=np.int64)
        for i, a in enumerate(xy):
            offs[i + 1] = offs[i] + len(a)
        allxy = np.zeros((int(offs[-1]), 2), dtype=np.float64)
        allb = np.zeros(int(offs[-1]), dtype=np.float64)
        for i, (a, b) in enumerate(zip(xy, brightness)):
            if len(a) != len(b):
                raise ValueError(f"image {i}: {len(a)} stars but {len(b)} brightness values")
            if len(a):
                allxy[offs[i]:offs[i + 1]] = np.asarray(a, dtype=np.float64).reshape(-1, 2)
                allb[offs[i]:offs[i + 1]] = np.asarray(b, dtype=np.float64)
        self._check(self._lib.ssp_upload_star_bank(self._ctx, _ptr(allxy), _ptr(allb), _ptr(offs),
                                                   len(offs) - 1, int(img_w), int(img_h),
                                                   float(focal_px), N.MEM_HOST))
        self.synchronize()
        self._star_offsets = offs
        self.star_owner = owner

    def upload_star_bank_resident(self, xy_ptr: int, brightness_ptr: int, star_offsets, img_w: int, img_h: int,
                                  focal_px: float, owner=None):
        """`upload_star_bank` for float64 arrays already on this device (e.g. the receive buffer of
        an NCCL broadcast): xy (N, 2) and brightness (N,), images back to back per `star_offsets`."""
        offs = np.ascontiguousarray(star_offsets, dtype=np.int64)
        self._check(self._lib.ssp_upload_star_bank(self._ctx, ctypes.c_void_p(xy_ptr), ctypes.c_void_p(brightness_ptr),
                                                   _ptr(offs), len(offs) - 1, int(img_w), int(img_h),
                                                   float(focal_px), N.MEM_DEVICE))
        self._star_offsets = offs.copy()
        self.star_owner = owner

    def star_rays(self):
        n = int(self._star_offsets[-1])
        rays = np.empty((n, 3), dtype=np.float64)
        self._check(self._lib.ssp_fetch_star_rays(self._ctx, _ptr(rays), n, N.MEM_HOST))
        return rays

    def polygon_measure(self, image: int, candidates, min_side_px: float = 5.0,
                        return_status: bool = False, owner=None):
        """Measures candidate 5-star polygons of one image (cbpm:500-600 + :722-740) into the
        engine's polygon bank.  Returns n_valid (and the per-candidate status)."""
        cand = np.ascontiguousarray(np.asarray(candidates, dtype=np.int32).reshape(-1, 5))
        status = np.empty(cand.shape[0], dtype=np.uint8) if return_status else None
        n_valid = ctypes.c_int64()
        self.poly_owner[int(image)] = owner
        self._check(self._lib.ssp_polygon_measure(self._ctx, int(image), _ptr(cand), cand.shape[0],
                                                  float(min_side_px), N.MEM_HOST, _ptr(status),
                                                  ctypes.byref(n_valid)))
        return (n_valid.value, status) if return_status else n_valid.value

    def polygon_measure_resident(self, image: int, device_ptr: int, n_candidates: int,
                                 min_side_px: float = 5.0, wait: bool = True):
        """Same as `polygon_measure` for an int32 (n, 5) candidate table already on this device.
        wait=False does not wait for the device (measure a batch of images back to back); the
        count is then available from `polygon_count` and returned as None here."""
        n_valid = ctypes.c_int64()
        self.poly_owner[int(image)] = None
        self._check(self._lib.ssp_polygon_measure(self._ctx, int(image), ctypes.c_void_p(device_ptr),
                                                  int(n_candidates), float(min_side_px), N.MEM_DEVICE,
                                                  None, ctypes.byref(n_valid) if wait else None))
        return n_valid.value if wait else None

    def polygon_measure_batch_resident(self, images: Sequence[int], device_ptr: int, cand_offsets,
                                       min_side_px: float = 5.0):
        """Measures several images in one launch pair: images[k] owns rows [cand_offsets[k],
        cand_offsets[k+1]) of the int32 (n, 5) candidate table at `device_ptr`.  Does not wait; counts
        come from `polygon_count`."""
        ids = np.ascontiguousarray(images, dtype=np.int32)
        offs = np.ascontiguousarray(cand_offsets, dtype=np.int64)
        for i in ids.tolist():
            self.poly_owner[int(i)] = None
        self._check(self._lib.ssp_polygon_measure_batch(self._ctx, _ptr(ids), len(ids), ctypes.c_void_p(device_ptr),
                                                        _ptr(offs), float(min_side_px), N.MEM_DEVICE))

    def polygon_count(self, image: int) -> int:
        n = ctypes.c_int64()
        self._check(self._lib.ssp_polygon_count(self._ctx, int(image), ctypes.byref(n)))
        return n.value

    def polygon_fetch(self, image: int, raw: bool = True):
        """Polygon bank of one image as numpy arrays in the reference's layout."""
        n = self.polygon_count(image)
        out = {"star_idx": np.empty((n, 5), dtype=np.int32), "feat": np.empty((n, 15), dtype=np.float64)}
        if raw:
            for k in ("angle", "side", "brightness"):
                out[k] = np.empty((n, 5), dtype=np.float64)
        self._check(self._lib.ssp_polygon_fetch(
            self._ctx, int(image), _ptr(out["star_idx"]), _ptr(out.get("angle")), _ptr(out.get("side")),
            _ptr(out.get("brightness")), _ptr(out["feat"]), n, N.MEM_HOST))
        return out

    def polygon_fetch_resident(self, image: int, star_idx_ptr: int, feat_ptr: int, capacity: int):
        """Copies the stored polygons of one image device -> device: int32 (P, 5) star indices and
        float64 (P, 15) normalised features (either address may be 0)."""
        self._check(self._lib.ssp_polygon_fetch(
            self._ctx, int(image), ctypes.c_void_p(star_idx_ptr or None), None, None, None,
            ctypes.c_void_p(feat_ptr or None), int(capacity), N.MEM_DEVICE))

    def polygon_upload_features_resident(self, image: int, feat_ptr: int, star_idx_ptr: int, n_polygons: int,
                                         owner=None):
        """`polygon_upload_features` from device memory (the all-gathered banks of other ranks)."""
        self.poly_owner[int(image)] = owner
        self._check(self._lib.ssp_polygon_upload_features(
            self._ctx, int(image), ctypes.c_void_p(feat_ptr), ctypes.c_void_p(star_idx_ptr or None),
            int(n_polygons), N.MEM_DEVICE))

    def polygon_upload_features(self, image: int, feat, star_idx=None, owner=None):
        self.poly_owner[int(image)] = owner
        f = np.ascontiguousarray(np.asarray(feat, dtype=np.float64).reshape(-1, 15))
        s = None
        if star_idx is not None:
            s = np.ascontiguousarray(np.asarray(star_idx, dtype=np.int32).reshape(-1, 5))
            if s.shape[0] != f.shape[0]:
                raise ValueError("star_idx and feat row counts differ")
        self._check(self._lib.ssp_polygon_upload_features(self._ctx, int(image), _ptr(f), _ptr(s),
                                                          f.shape[0], N.MEM_HOST))

    # -- kernel 3 -------------------------------------------------------------------------
    def polygon_match_pairs(self, pairs, threshold: float = 0.1566, min_over_shifts: bool = False,
                            fp32: bool = False, return_rowmin: bool = False,
                            fetch: bool = True) -> PolygonMatchResult:
        """Per polygon of the first image the nearest polygon of the second (cbpm:992-1067), kept
        iff distance < threshold (cbpm:1145-1148).  Defaults are the reference's parity mode."""
        p = _as_pairs(pairs)
        offs = np.zeros(p.shape[0] + 1, dtype=np.int64)
        self._check(self._lib.ssp_polygon_match_pairs(
            self._ctx, _ptr(p), p.shape[0], float(threshold),
            N.SHIFT_MIN_OVER_SHIFTS if min_over_shifts else N.SHIFT_CANONICAL,
            N.PRECISION_FP32 if fp32 else N.PRECISION_FP64, _ptr(offs)))
        n = int(offs[-1])
        m1 = np.empty(n, dtype=np.int32)
        m2 = np.empty(n, dtype=np.int32)
        md = np.empty(n, dtype=np.float64)
        if fetch:
            self._check(self._lib.ssp_polygon_fetch_matches(self._ctx, _ptr(m1), _ptr(m2), _ptr(md), n,
                                                            N.MEM_HOST))
        res = PolygonMatchResult(offs, m1, m2, md)
        if return_rowmin:
            rows = np.array([self.polygon_count(int(a)) for a in p[:, 0]], dtype=np.int64)
            r_off = np.zeros(p.shape[0] + 1, dtype=np.int64)
            np.cumsum(rows, out=r_off[1:])
            am = np.empty(int(r_off[-1]), dtype=np.int32)
            mn = np.empty(int(r_off[-1]), dtype=np.float64)
            self._check(self._lib.ssp_polygon_fetch_rowmin(self._ctx, _ptr(am), _ptr(mn), am.shape[0],
                                                           N.MEM_HOST))
            res.argmin, res.mindist, res.row_offsets = am, mn, r_off
        return res

    def polygon_fetch_into(self, p1_ptr: int, p2_ptr: int, dist_ptr: int, capacity: int, device: bool = True):
        """Copies the matches of the last `polygon_match_pairs(..., fetch=False)` into int32 / int32 /
        float64 buffers given by address (device memory, or host memory)."""
        self._check(self._lib.ssp_polygon_fetch_matches(
            self._ctx, ctypes.c_void_p(p1_ptr), ctypes.c_void_p(p2_ptr), ctypes.c_void_p(dist_ptr),
            int(capacity), N.MEM_DEVICE if device else N.MEM_HOST))

    def polygon_last_stats(self):
        n, ms, ev = ctypes.c_int(), ctypes.c_float(), ctypes.c_int64()
        self._check(self._lib.ssp_polygon_last_stats(self._ctx, ctypes.byref(n), ctypes.byref(ms),
                                                     ctypes.byref(ev)))
        fl = ctypes.c_int64()
        self._check(self._lib.ssp_polygon_last_executed_flops(self._ctx, ctypes.byref(fl)))
        f64, f32 = ctypes.c_int64(), ctypes.c_int64()
        self._check(self._lib.ssp_polygon_last_executed_flops_by_precision(self._ctx, ctypes.byref(f64), ctypes.byref(f32)))
        used, fb, cand = ctypes.c_int(), ctypes.c_int(), ctypes.c_int64()
        worst, bound = ctypes.c_double(), ctypes.c_double()
        self._check(self._lib.ssp_polygon_last_screen_stats(self._ctx, ctypes.byref(used), ctypes.byref(fb), ctypes.byref(cand),
                                                            ctypes.byref(worst), ctypes.byref(bound)))
        return {"launches": n.value, "main_kernel_ms": ms.value, "evaluations": ev.value,
                "tensor_screen": {"used": bool(used.value), "fallback_flags": fb.value,
                                  "candidates": cand.value if cand.value >= 0 else None,
                                  "worst_error_coef": worst.value if worst.value >= 0 else None,
                                  "bound_coef": bound.value},
                "executed_flops": fl.value if fl.value >= 0 else None,
                "executed_flops_fp64": f64.value if f64.value >= 0 else None,
                "executed_flops_fp32": f32.value if f32.value >= 0 else None}

    # -- row N3 ---------------------------------------------------------------------------
    def ransac_homography_pairs(self, keypoint_xy: Sequence[np.ndarray], pairs, n_matches: int,
                                reproj_threshold: float = 5.0, n_hypotheses: int = 2048, seed: int = 0,
                                min_matches: int = 6):
        """Batched RANSAC homographies for the pairs of the LAST `hamming_match_pairs` call (same pair
        list), point pairs gathered on the device from its survivors.  keypoint_xy[i]: (n_i, 2) pixel
        coordinates of image i's keypoints.  Returns (H float64 (n_pairs, 3, 3) with NaN where no
        model exists, inlier mask uint8 (n_matches,), n_inliers int32 (n_pairs,), -1 = fewer than
        min_matches matches).  Opt-in stand-in for cv.findHomography(..., cv.RANSAC, 5)
        (custom_bruteforce_matcher.py:215-246); same model, not OpenCV's random numbers."""
        p = _as_pairs(pairs)
        offs = np.zeros(len(keypoint_xy) + 1, dtype=np.int64)
        np.cumsum([len(a) for a in keypoint_xy], out=offs[1:])
        xy = np.ascontiguousarray(np.concatenate([np.asarray(a, dtype=np.float32).reshape(-1, 2) for a in keypoint_xy])
                                  if offs[-1] else np.zeros((0, 2), np.float32))
        H = np.empty((p.shape[0], 9), dtype=np.float64)
        mask = np.empty(max(int(n_matches), 1), dtype=np.uint8)
        inl = np.empty(p.shape[0], dtype=np.int32)
        self._check(self._lib.ssp_ransac_homography_pairs(
            self._ctx, _ptr(xy), _ptr(offs), len(offs) - 1, _ptr(p), p.shape[0], float(reproj_threshold),
            int(n_hypotheses), int(seed) & 0xFFFFFFFF, int(min_matches), _ptr(H), _ptr(mask), mask.shape[0], _ptr(inl)))
        return H.reshape(-1, 3, 3), mask[:int(n_matches)], inl

    # -- row N2 ---------------------------------------------------------------------------
    def star_regions(self, image: int) -> np.ndarray:
        """uint16 bit mask per star of `image`: membership in the 12 border regions of cbpm:328-375."""
        n = int(self._star_offsets[image + 1] - self._star_offsets[image])
        out = np.zeros(n, dtype=np.uint16)
        self._check(self._lib.ssp_star_regions(self._ctx, int(image), _ptr(out), n))
        return out

    def cluster_brightest(self, image: int, members, labels, region_offsets, max_labels: int) -> np.ndarray:
        """Brightest star per (region, cluster label), last of equals (cbpm:230-245); int32
        (n_regions, max_labels), -1 for unused labels."""
        m = np.ascontiguousarray(members, dtype=np.int32)
        lab = np.ascontiguousarray(labels, dtype=np.int32)
        off = np.ascontiguousarray(region_offsets, dtype=np.int32)
        out = np.empty((len(off) - 1, int(max_labels)), dtype=np.int32)
        self._check(self._lib.ssp_cluster_brightest(self._ctx, int(image), _ptr(m), _ptr(lab), _ptr(off),
                                                    len(off) - 1, int(max_labels), _ptr(out)))
        return out

    def polygon_candidates(self, region_members: Sequence[Sequence[int]], max_candidates: int = 15000):
        """set(itertools.combinations(region list, 5) for every region) (cbpm:407-413); every region's
        stars in LIST order.  Returns (n_tuples = len() of the reference's set, n_unique star sets,
        device address of the int32 (n_unique, 5) table or None when n_tuples is 0 or above the maximum)."""
        counts = np.array([len(r) for r in region_members], dtype=np.int32)
        flat = np.ascontiguousarray(np.concatenate([np.asarray(r, dtype=np.int32) for r in region_members])
                                    if counts.sum() else np.zeros(1, np.int32), dtype=np.int32)
        n_t, n_u = ctypes.c_int64(), ctypes.c_int64()
        ptr = ctypes.c_void_p()
        self._check(self._lib.ssp_polygon_candidates(self._ctx, _ptr(flat), _ptr(counts), len(counts), int(max_candidates),
                                                     ctypes.byref(n_t), ctypes.byref(n_u), ctypes.byref(ptr)))
        return n_t.value, n_u.value, ptr.value

    def copy_to_host(self, device_ptr: int, shape, dtype) -> np.ndarray:
        """numpy copy of an engine-owned device table (e.g. the candidate table)."""
        out = np.empty(shape, dtype=dtype)
        self._check(self._lib.ssp_copy_to_host(self._ctx, _ptr(out), ctypes.c_void_p(device_ptr), out.nbytes))
        return out

    # -- row N1 ---------------------------------------------------------------------------
    def star_brightness(self, gray: np.ndarray, centers_xy, disc_dxdy, top_k: int = 10):
        """get_star_brightness (image_processors.py:222-254) for many stars of one grey image."""
        g = np.ascontiguousarray(gray, dtype=np.uint8)
        if g.ndim != 2:
            raise ValueError("gray must be (h, w) uint8")
        c = np.ascontiguousarray(np.asarray(centers_xy, dtype=np.int32).reshape(-1, 2))
        d = np.ascontiguousarray(np.asarray(disc_dxdy, dtype=np.int32).reshape(-1, 2))
        out = np.empty(c.shape[0], dtype=np.int32)
        self._check(self._lib.ssp_star_brightness(self._ctx, _ptr(g), g.shape[0], g.shape[1], _ptr(c),
                                                  c.shape[0], _ptr(d), d.shape[0], int(top_k),
                                                  _ptr(out), N.MEM_HOST))
        return out

    def star_brightness_static(self, gray: np.ndarray, centers_xy, disc_dxdy, top_k: int = 10):
        """`star_brightness` through `ssp_star_brightness_static`: no allocation on the device, the staging
        lives in a fixed device array (at most 2 Mi pixels and 16 384 stars per call; calls on one GPU must
        not overlap).  No counterpart in the reference."""
        g = np.ascontiguousarray(gray, dtype=np.uint8)
        if g.ndim != 2:
            raise ValueError("gray must be (h, w) uint8")
        c = np.ascontiguousarray(np.asarray(centers_xy, dtype=np.int32).reshape(-1, 2))
        d = np.ascontiguousarray(np.asarray(disc_dxdy, dtype=np.int32).reshape(-1, 2))
        out = np.empty(c.shape[0], dtype=np.int32)
        self._check(self._lib.ssp_star_brightness_static(self._ctx, _ptr(g), g.shape[0], g.shape[1], _ptr(c),
                                                         c.shape[0], _ptr(d), d.shape[0], int(top_k),
                                                         _ptr(out), N.MEM_HOST))
        return out
