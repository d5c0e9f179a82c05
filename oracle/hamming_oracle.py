"""oracle/hamming_oracle.py — TEST INFRASTRUCTURE, not product code.

ctypes face of oracle/hamming_oracle.c (the CPU restatement of cv::batchDistance's Hamming
kNN-2 / crossCheck semantics; see that file's header for the reference call sites
custom_bruteforce_matcher.py:160-172 and :193-199 and for how parity is pinned).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "build", "libhamming_oracle.so")
_lib = None


def build():
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_hamming_crosscheck.restype = ctypes.c_int
        _lib.oracle_ratio_survivors.restype = ctypes.c_int
    return _lib


def _rows(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if a.size == 0:
        return a.reshape(0, 32)
    assert a.ndim == 2 and a.shape[1] == 32, a.shape
    return a


def _p(a):
    return ctypes.c_void_p(a.ctypes.data)


def knn2(q, t):
    """int32 (nq, 4) {idx1, d1, idx2, d2}: first two rows of a stable sort of each distance row."""
    q, t = _rows(q), _rows(t)
    out = np.full((q.shape[0], 4), -1, dtype=np.int32)
    _load().oracle_hamming_knn2(_p(q), ctypes.c_int(q.shape[0]), _p(t), ctypes.c_int(t.shape[0]), _p(out))
    return out


def ratio_survivors(best2, num=3, den=4):
    """(query_idx, train_idx, distance) of rows with den*d1 < num*d2 (Lowe ratio num/den)."""
    best2 = np.ascontiguousarray(best2, dtype=np.int32)
    n = best2.shape[0]
    mq, mt, md = (np.empty(n, dtype=np.int32) for _ in range(3))
    k = _load().oracle_ratio_survivors(_p(best2), ctypes.c_int(n), ctypes.c_int(num), ctypes.c_int(den),
                                       _p(mq), _p(mt), _p(md))
    return mq[:k].copy(), mt[:k].copy(), md[:k].copy()


def crosscheck(q, t):
    """(query_idx, train_idx, distance) of BFMatcher(NORM_HAMMING, crossCheck=True).match(q, t)."""
    q, t = _rows(q), _rows(t)
    nq, nt = q.shape[0], t.shape[0]
    work = np.empty(2 * (nq + nt) + 4, dtype=np.int32)
    mq, mt, md = (np.empty(max(nq, 1), dtype=np.int32) for _ in range(3))
    k = _load().oracle_hamming_crosscheck(_p(q), ctypes.c_int(nq), _p(t), ctypes.c_int(nt), _p(work),
                                          _p(mq), _p(mt), _p(md))
    return mq[:k].copy(), mt[:k].copy(), md[:k].copy()
