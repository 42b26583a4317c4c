import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200")
for p in (ROOT, PKG, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "first_hardware_run: GPU test of code that has not run on hardware yet; such "
                            "tests are collected last so that a surprise in them cannot hide the measured ones")


def pytest_collection_modifyitems(config, items):
    def rank(item):  # stable sort: unmarked tests keep their order and run first; marked ones by their `order`
        if os.path.basename(str(item.fspath)).startswith("test_zz_"):
            return 1 << 30  # tests/test_zz_no_alloc.py stays the very last file, behind the marked ones too
        m = item.get_closest_marker("first_hardware_run")
        return 0 if m is None else 1 + int(m.kwargs.get("order", 0))

    items.sort(key=rank)


def _lib_built():
    return os.path.exists(os.path.join(PKG, "diff_gaussian_rasterization", "libggs_rasterizer.so"))


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    """The product has no fallback: build the library once if the tree is fresh."""
    if not _lib_built():
        import subprocess

        subprocess.check_call(["make", "-C", os.path.join(PKG, "csrc"), "-j", "8"])
    yield


@pytest.fixture(scope="session")
def device():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return "cuda:0"


@pytest.fixture(scope="session")
def ref_module():
    """The compiled reference rasterizer (Oracle A); None when oracle/_ref was not shipped."""
    from oracle import ref_loader

    if not ref_loader.available():
        return None
    return ref_loader.load()
