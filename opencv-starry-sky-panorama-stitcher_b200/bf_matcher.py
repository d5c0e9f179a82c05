"""Drop-in for `cv.BFMatcher(cv.NORM_HAMMING[, crossCheck=True])` backed by CUDA kernel 1.

Replaces the two call sites of the reference:
  custom_bruteforce_matcher.py:160-162   bf.knnMatch(q, t, k=2)
  custom_bruteforce_matcher.py:193-199   bf.match(cv.UMat(q), cv.UMat(t))   (crossCheck=True)
Results are `cv2.DMatch` tuples in OpenCV's order (ascending queryIdx, imgIdx 0, distance =
integer Hamming distance as float) so the unchanged host code cannot tell the difference.
"""
import cv2
import numpy as np

from .engine import Engine

_shared_engine = None


def shared_engine(device: int = 0) -> Engine:
    """Process-wide engine used by the drop-in classes when none is passed explicitly."""
    global _shared_engine
    if _shared_engine is None:
        _shared_engine = Engine(device)
    return _shared_engine


def as_descriptor_array(d):
    """cv2.UMat | ndarray | None -> uint8 (n, 32) C-contiguous (custom_bruteforce_matcher.py:197)."""
    if d is None:
        return np.zeros((0, 32), dtype=np.uint8)
    if isinstance(d, cv2.UMat):
        d = d.get()
    d = np.asarray(d)
    if d.size == 0:
        return np.zeros((0, 32), dtype=np.uint8)
    if d.dtype != np.uint8 or d.ndim != 2 or d.shape[1] != 32:
        raise ValueError(f"expected 256-bit ORB descriptors uint8 (n, 32), got {d.dtype} {d.shape}")
    return np.ascontiguousarray(d)


def dmatches_from_arrays(query_idx, train_idx, distance):
    """SoA engine output -> tuple of cv2.DMatch exactly as cv.BFMatcher builds them."""
    return tuple(cv2.DMatch(int(q), int(t), 0, float(d))
                 for q, t, d in zip(query_idx.tolist(), train_idx.tolist(), distance.tolist()))


class BFMatcher:
    def __init__(self, normType=cv2.NORM_HAMMING, crossCheck=False, engine: Engine = None,
                 variant: str = "auto"):
        if normType not in (cv2.NORM_HAMMING,):
            raise NotImplementedError("the B200 engine implements NORM_HAMMING on 256-bit descriptors only")
        self.crossCheck = bool(crossCheck)
        self.engine = engine
        self.variant = variant

    @classmethod
    def create(cls, normType=cv2.NORM_HAMMING, crossCheck=False, **kw):
        return cls(normType, crossCheck, **kw)

    def _engine(self):
        return self.engine if self.engine is not None else shared_engine()

    def _run(self, q, t, mode, ratio=0.75, best2=False):
        eng = self._engine()
        eng.upload_descriptor_bank([as_descriptor_array(q), as_descriptor_array(t)])
        return eng.hamming_match_pairs([(0, 1)], mode=mode, ratio=ratio, variant=self.variant,
                                       return_best2=best2)

    def match(self, queryDescriptors, trainDescriptors):
        if self.crossCheck:
            res = self._run(queryDescriptors, trainDescriptors, "crosscheck")
            return dmatches_from_arrays(*res.pair(0))
        # plain nearest neighbour: one DMatch per query row that has a neighbour
        res = self._run(queryDescriptors, trainDescriptors, "knn2_ratio", best2=True)
        b = res.best2
        rows = np.nonzero(b[:, 0] >= 0)[0]
        return dmatches_from_arrays(rows, b[rows, 0], b[rows, 1])

    def knnMatch(self, queryDescriptors, trainDescriptors, k=2):
        if self.crossCheck or k not in (1, 2):
            raise NotImplementedError("knnMatch is implemented for k in (1, 2) without crossCheck")
        res = self._run(queryDescriptors, trainDescriptors, "knn2_ratio", best2=True)
        out = []
        for qi, (i1, d1, i2, d2) in enumerate(res.best2.tolist()):
            row = []
            if i1 >= 0:
                row.append(cv2.DMatch(qi, i1, 0, float(d1)))
            if k == 2 and i2 >= 0:
                row.append(cv2.DMatch(qi, i2, 0, float(d2)))
            out.append(tuple(row))
        return tuple(out)
