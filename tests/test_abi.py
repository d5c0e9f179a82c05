"""The C-ABI library loads and exports every symbol include/ssp_b200.h declares (no compute)."""
import ctypes
import os
import re

import pytest

from conftest import HAS_GPU, ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ssp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ssp_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_documented_surface():
    syms = declared_symbols()
    for must in ("ssp_create", "ssp_destroy", "ssp_last_error", "ssp_upload_descriptor_bank",
                 "ssp_hamming_match_pairs", "ssp_upload_star_bank", "ssp_polygon_measure",
                 "ssp_polygon_match_pairs", "ssp_star_brightness"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from ssp_b200 import _native
    assert os.path.exists(_native.LIB_PATH), "build with __graft_entry__.build()"
    lib = ctypes.CDLL(_native.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/ssp_b200.h but not exported"


def test_binding_covers_every_declared_symbol():
    from ssp_b200 import _native
    assert sorted(_native.PROTOTYPES) == declared_symbols()
    _native.load()  # attaches prototypes, checks the ABI version


def test_no_abi_signature_mentions_torch_or_cv():
    text = open(os.path.join(ROOT, "include", "ssp_b200.h")).read()
    code = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    assert "torch" not in code and "cv::" not in code and "Tensor" not in code


@pytest.mark.skipif(HAS_GPU, reason="only meaningful on a box without a GPU")
def test_engine_fails_loudly_without_gpu():
    import ssp_b200
    with pytest.raises(ssp_b200.EngineError, match="no CPU fallback"):
        ssp_b200.Engine(0)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "opencv-starry-sky-panorama-stitcher_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("# oracle", ""), f"{f} references oracle/"
