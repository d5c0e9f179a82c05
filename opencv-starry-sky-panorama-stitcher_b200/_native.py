"""ctypes binding of libssp_b200.so (the C ABI declared in include/ssp_b200.h).

There is no CPU fallback: importing this module fails loudly when the shared library has not
been built, and creating an engine fails when no B200 is visible.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libssp_b200.so")

# constants mirrored from include/ssp_b200.h
ABI_VERSION = 1
OK = 0
ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_CAPACITY, ERR_DOMAIN = -1, -2, -3, -4, -5
MEM_HOST, MEM_DEVICE = 0, 1
MODE_KNN2_RATIO, MODE_CROSSCHECK = 0, 1
HAMMING_AUTO, HAMMING_POPC, HAMMING_POPC_CSA, HAMMING_TENSOR = 0, 1, 2, 3
SHIFT_CANONICAL, SHIFT_MIN_OVER_SHIFTS = 0, 1
PRECISION_FP64, PRECISION_FP32 = 0, 1
POLY_VALID, POLY_SHORT_SIDE, POLY_NOT_CONVEX, POLY_DOMAIN, POLY_BAD_INDEX = 0, 1, 2, 3, 4

_c = ctypes
_ctx_p = _c.c_void_p
_vp = _c.c_void_p
_i32p = _c.c_void_p
_i64p = _c.c_void_p
_f64p = _c.c_void_p
_u8p = _c.c_void_p

# name -> (restype, argtypes); every symbol include/ssp_b200.h declares
PROTOTYPES = {
    "ssp_create": (_c.c_int, [_c.POINTER(_ctx_p), _c.c_int]),
    "ssp_destroy": (_c.c_int, [_ctx_p]),
    "ssp_last_error": (_c.c_char_p, [_ctx_p]),
    "ssp_abi_version": (_c.c_int, []),
    "ssp_set_stream": (_c.c_int, [_ctx_p, _vp]),
    "ssp_synchronize": (_c.c_int, [_ctx_p]),
    "ssp_device_info": (_c.c_int, [_ctx_p, _c.POINTER(_c.c_int), _c.POINTER(_c.c_int),
                                   _c.POINTER(_c.c_int64)]),
    "ssp_host_alloc": (_c.c_int, [_ctx_p, _c.c_int64, _c.POINTER(_c.c_void_p)]),
    "ssp_host_free": (_c.c_int, [_ctx_p, _vp]),
    "ssp_copy_to_host": (_c.c_int, [_ctx_p, _vp, _vp, _c.c_int64]),
    "ssp_upload_descriptor_bank": (_c.c_int, [_ctx_p, _u8p, _i64p, _c.c_int, _c.c_int]),
    "ssp_bind_descriptor_bank": (_c.c_int, [_ctx_p, _u8p, _i64p, _c.c_int]),
    "ssp_hamming_match_pairs": (_c.c_int, [_ctx_p, _i32p, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                           _c.c_int, _i64p]),
    "ssp_hamming_match_pairs_f64": (_c.c_int, [_ctx_p, _i32p, _c.c_int, _c.c_int, _c.c_double,
                                               _c.c_int, _i64p]),
    "ssp_hamming_fetch_matches": (_c.c_int, [_ctx_p, _i32p, _i32p, _i32p, _c.c_int64, _c.c_int]),
    "ssp_hamming_fetch_best2": (_c.c_int, [_ctx_p, _i32p, _c.c_int64, _c.c_int]),
    "ssp_hamming_last_stats": (_c.c_int, [_ctx_p, _c.POINTER(_c.c_int), _c.POINTER(_c.c_float),
                                          _c.POINTER(_c.c_int64)]),
    "ssp_upload_star_bank": (_c.c_int, [_ctx_p, _f64p, _f64p, _i64p, _c.c_int, _c.c_int, _c.c_int,
                                        _c.c_double, _c.c_int]),
    "ssp_fetch_star_rays": (_c.c_int, [_ctx_p, _f64p, _c.c_int64, _c.c_int]),
    "ssp_polygon_measure": (_c.c_int, [_ctx_p, _c.c_int, _i32p, _c.c_int64, _c.c_double, _c.c_int,
                                       _u8p, _c.POINTER(_c.c_int64)]),
    "ssp_polygon_measure_batch": (_c.c_int, [_ctx_p, _i32p, _c.c_int, _i32p, _i64p, _c.c_double, _c.c_int]),
    "ssp_polygon_fetch": (_c.c_int, [_ctx_p, _c.c_int, _i32p, _f64p, _f64p, _f64p, _f64p,
                                     _c.c_int64, _c.c_int]),
    "ssp_polygon_count": (_c.c_int, [_ctx_p, _c.c_int, _c.POINTER(_c.c_int64)]),
    "ssp_polygon_upload_features": (_c.c_int, [_ctx_p, _c.c_int, _f64p, _i32p, _c.c_int64,
                                               _c.c_int]),
    "ssp_polygon_match_pairs": (_c.c_int, [_ctx_p, _i32p, _c.c_int, _c.c_double, _c.c_int,
                                           _c.c_int, _i64p]),
    "ssp_polygon_fetch_matches": (_c.c_int, [_ctx_p, _i32p, _i32p, _f64p, _c.c_int64, _c.c_int]),
    "ssp_polygon_fetch_rowmin": (_c.c_int, [_ctx_p, _i32p, _f64p, _c.c_int64, _c.c_int]),
    "ssp_polygon_last_stats": (_c.c_int, [_ctx_p, _c.POINTER(_c.c_int), _c.POINTER(_c.c_float),
                                          _c.POINTER(_c.c_int64)]),
    "ssp_polygon_last_executed_flops": (_c.c_int, [_ctx_p, _c.POINTER(_c.c_int64)]),
    "ssp_polygon_last_executed_flops_by_precision": (_c.c_int, [_ctx_p, _c.POINTER(_c.c_int64), _c.POINTER(_c.c_int64)]),
    "ssp_polygon_last_screen_stats": (_c.c_int, [_ctx_p, _c.POINTER(_c.c_int), _c.POINTER(_c.c_int), _c.POINTER(_c.c_int64),
                                                 _c.POINTER(_c.c_double), _c.POINTER(_c.c_double)]),
    "ssp_ransac_homography_pairs": (_c.c_int, [_ctx_p, _vp, _i64p, _c.c_int, _i32p, _c.c_int, _c.c_double, _c.c_int,
                                               _c.c_uint32, _c.c_int, _f64p, _u8p, _c.c_int64, _i32p]),
    "ssp_star_regions": (_c.c_int, [_ctx_p, _c.c_int, _vp, _c.c_int64]),
    "ssp_cluster_brightest": (_c.c_int, [_ctx_p, _c.c_int, _i32p, _i32p, _i32p, _c.c_int, _c.c_int, _i32p]),
    "ssp_polygon_candidates": (_c.c_int, [_ctx_p, _i32p, _i32p, _c.c_int, _c.c_int64, _c.POINTER(_c.c_int64),
                                          _c.POINTER(_c.c_int64), _c.POINTER(_c.c_void_p)]),
    "ssp_star_brightness": (_c.c_int, [_ctx_p, _u8p, _c.c_int, _c.c_int, _i32p, _c.c_int64, _i32p,
                                       _c.c_int, _c.c_int, _i32p, _c.c_int]),
    "ssp_star_brightness_static": (_c.c_int, [_ctx_p, _u8p, _c.c_int, _c.c_int, _i32p, _c.c_int64, _i32p,
                                              _c.c_int, _c.c_int, _i32p, _c.c_int]),
}


class NativeLibraryMissing(ImportError):
    pass


def load():
    """Loads libssp_b200.so and attaches the prototypes.  Raises if the library is absent."""
    if not os.path.exists(LIB_PATH):
        raise NativeLibraryMissing(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (or `make -C opencv-starry-sky-panorama-stitcher_b200/csrc`). "
            "This engine has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.ssp_abi_version() != ABI_VERSION:
        raise ImportError(f"libssp_b200.so ABI {lib.ssp_abi_version()} != binding ABI {ABI_VERSION}")
    return lib


_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        _LIB = load()
    return _LIB
