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


APP_DIR = os.path.join(_REF_DIR, "ref_app")
_APP_MODULES = ("gaussian_renderer", "scene", "scene.gaussian_model", "scene.cameras", "scene.appearance_network",
                "utils", "utils.general_utils", "utils.graphics_utils", "utils.sh_utils", "utils.system_utils",
                "utils.loss_utils", "arguments")


def app_available() -> bool:
    return os.path.exists(os.path.join(APP_DIR, ".installed"))


class ReferenceApp:
    """The reference's own rasterizer callers (gaussian_renderer, scene.gaussian_model, scene.cameras, utils.loss_utils;
    installed verbatim by oracle/build_ref.py::build_app) imported on top of a chosen set of kernel packages.

        with ReferenceApp(diff_gaussian_rasterization=ours, warp_patch_ncc=ours_ncc, fused_ssim=ours_ssim) as app:
            app.gaussian_renderer.render(...)

    The app modules bind `diff_gaussian_rasterization` etc. at import time, so they are imported afresh for every
    instance and removed from sys.modules again on exit; modules the image does not have (plyfile, trimesh, cv2,
    simple_knn) come from tests/shims."""

    def __init__(self, shims_dir, **backends):
        self.backends = backends
        self.shims_dir = shims_dir
        self.saved = {}

    def __enter__(self):
        if not app_available():
            raise ImportError("oracle/_ref/ref_app not installed; run `python oracle/build_ref.py` where /root/reference exists")
        names = list(_APP_MODULES) + list(self.backends)
        self.saved = {n: sys.modules.pop(n) for n in names if n in sys.modules}
        self.paths = [p for p in (APP_DIR, self.shims_dir) if p]
        for p in reversed(self.paths):
            sys.path.insert(0, p)
        for name, mod in self.backends.items():
            sys.modules[name] = mod
        self.gaussian_renderer = importlib.import_module("gaussian_renderer")
        self.gaussian_model = importlib.import_module("scene.gaussian_model")
        self.cameras = importlib.import_module("scene.cameras")
        self.loss_utils = importlib.import_module("utils.loss_utils")
        return self

    def __exit__(self, *exc):
        for n in list(_APP_MODULES) + list(self.backends):
            sys.modules.pop(n, None)
        sys.modules.update(self.saved)
        for p in self.paths:
            if p in sys.path:
                sys.path.remove(p)
        return False
