"""Import alias: the package directory is named after the reference
(`opencv-starry-sky-panorama-stitcher_b200/`), which is not a valid Python identifier, so this
module loads it under the name `ssp_b200` (submodules included: `import ssp_b200.engine`)."""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "opencv-starry-sky-panorama-stitcher_b200")
_spec = importlib.util.spec_from_file_location(
    "ssp_b200", os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["ssp_b200"] = _mod
_spec.loader.exec_module(_mod)
