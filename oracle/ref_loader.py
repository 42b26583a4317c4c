"""Import the compiled reference rasterizer (Oracle A) as the module `ref_dgr`.  TEST INFRASTRUCTURE ONLY.

oracle/_ref/ref_dgr/ holds build outputs of oracle/build_ref.py: the reference's unmodified CUDA
extension compiled for sm_100 (`_C.so`) and its Python wrapper as installed.  It needs a GPU to run
(the reference has no CPU path), so only `-m gpu` tests, bench.py --impl reference and
tests/golden/make_golden.py use it.
"""
import importlib
import os
import sys

_REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available() -> bool:
    return os.path.exists(os.path.join(_REF_DIR, "ref_dgr", "_C.so"))


def load():
    if not available():
        raise ImportError("oracle/_ref/ref_dgr/_C.so not built; run `python oracle/build_ref.py` where "
                          "/root/reference exists")
    import torch  # noqa: F401  (libtorch must be loaded before the extension)

    if _REF_DIR not in sys.path:
        sys.path.insert(0, _REF_DIR)
    return importlib.import_module("ref_dgr")


def ncc_available() -> bool:
    return os.path.exists(os.path.join(_REF_DIR, "ref_wpn", "_C.so"))


def load_ncc():
    """The compiled reference `warp_patch_ncc` module (submodules/warp-patch-ncc) as `ref_wpn`."""
    if not ncc_available():
        raise ImportError("oracle/_ref/ref_wpn/_C.so not built; run `python oracle/build_ref.py` where "
                          "/root/reference exists")
    import torch  # noqa: F401

    if _REF_DIR not in sys.path:
        sys.path.insert(0, _REF_DIR)
    return importlib.import_module("ref_wpn")
