"""Build the UNMODIFIED reference rasterizer (Oracle A) into oracle/_ref/.

TEST INFRASTRUCTURE ONLY. The reference CUDA extension
(/root/reference/submodules/diff-gaussian-rasterization, setup.py:6-38) is
compiled from the sources where they lie, with the reference's exact nvcc
flags (-O3 --use_fast_math -std=c++17 --extended-lambda
--expt-relaxed-constexpr, setup.py:22-33) for sm_100.  Only build outputs
land in the repo, under the git-ignored oracle/_ref/:

    oracle/_ref/ref_dgr/_C.so        the compiled extension (pybind module "_C")
    oracle/_ref/ref_dgr/__init__.py  the reference's Python wrapper (an install
                                     artefact, like `pip install --target`)

    oracle/_ref/ref_wpn/_C.so        the reference's warp-patch-ncc extension (submodules/warp-patch-ncc,
    oracle/_ref/ref_wpn/__init__.py  setup.py:6-20: -O3 --use_fast_math), importable as `ref_wpn`

The package is importable as `ref_dgr` (see oracle/ref_loader.py) so it can
live beside the drop-in `diff_gaussian_rasterization` in one process.
Needs /root/reference; on the GPU box only the prebuilt files are used.
"""
import os
import shutil
import sys

REF = "/root/reference/submodules/diff-gaussian-rasterization"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref", "ref_dgr")
BUILD = os.environ.get("GGS_REF_BUILD_DIR", "/tmp/ggs_ref_build")


def build(force: bool = False) -> str:
    so = os.path.join(OUT, "_C.so")
    if os.path.exists(so) and not force:
        return so
    if not os.path.isdir(REF):
        raise RuntimeError("reference sources not present at " + REF)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
    from torch.utils.cpp_extension import load

    os.makedirs(BUILD, exist_ok=True)
    srcs = [
        "cuda_rasterizer/rasterizer_impl.cu",
        "cuda_rasterizer/render_forward.cu",
        "cuda_rasterizer/render_backward.cu",
        "cuda_rasterizer/sample_forward.cu",
        "cuda_rasterizer/sample_backward.cu",
        "rasterize_points.cu",
        "ext.cpp",
    ]
    load(
        name="_C",
        sources=[os.path.join(REF, s) for s in srcs],
        extra_cuda_cflags=[
            "-O3",
            "--use_fast_math",
            "-std=c++17",
            "--extended-lambda",
            "--expt-relaxed-constexpr",
            "-I" + os.path.join(REF, "third_party/glm/"),
        ],
        extra_include_paths=[REF],
        build_directory=BUILD,
        is_python_module=False,
        verbose=True,
    )
    os.makedirs(OUT, exist_ok=True)
    shutil.copyfile(os.path.join(BUILD, "_C.so"), so)
    shutil.copyfile(
        os.path.join(REF, "diff_gaussian_rasterization", "__init__.py"),
        os.path.join(OUT, "__init__.py"),
    )
    return so


REF_NCC = "/root/reference/submodules/warp-patch-ncc"
OUT_NCC = os.path.join(HERE, "_ref", "ref_wpn")


def build_ncc(force: bool = False) -> str:
    """The reference's patch-NCC extension, same recipe: its three sources, its two nvcc flags."""
    so = os.path.join(OUT_NCC, "_C.so")
    if os.path.exists(so) and not force:
        return so
    if not os.path.isdir(REF_NCC):
        raise RuntimeError("reference sources not present at " + REF_NCC)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
    from torch.utils.cpp_extension import load

    bdir = BUILD + "_ncc"
    os.makedirs(bdir, exist_ok=True)
    srcs = ["cuda_warp_patch_ncc/warp_patch_ncc_impl.cu", "warp_patch_ncc.cu", "ext.cpp"]
    load(name="_C", sources=[os.path.join(REF_NCC, s) for s in srcs], extra_cuda_cflags=["-O3", "--use_fast_math"],
         extra_include_paths=[REF_NCC], build_directory=bdir, is_python_module=False, verbose=True)
    os.makedirs(OUT_NCC, exist_ok=True)
    shutil.copyfile(os.path.join(bdir, "_C.so"), so)
    shutil.copyfile(os.path.join(REF_NCC, "warp_patch_ncc", "__init__.py"), os.path.join(OUT_NCC, "__init__.py"))
    return so


REF_ROOT = "/root/reference"
OUT_APP = os.path.join(HERE, "_ref", "ref_app")
# the reference's own callers of the rasterizer: installed verbatim (install artefacts, git-ignored like ref_dgr), so that
# tests can run gaussian_renderer.render / integrate / evaluate_sdf / sample_depth and utils.loss_utils.PatchMatch
# UNCHANGED on top of either rasterizer package (tests/test_reference_callers_gpu.py)
APP_FILES = ("gaussian_renderer/__init__.py", "scene/gaussian_model.py", "scene/cameras.py",
             "scene/appearance_network.py", "utils/general_utils.py", "utils/graphics_utils.py", "utils/sh_utils.py",
             "utils/system_utils.py", "utils/loss_utils.py", "arguments/__init__.py")


def build_app(force: bool = False) -> str:
    """Copies the reference's rasterizer callers into oracle/_ref/ref_app/.  The reference's scene/__init__.py pulls in
    its dataset readers (COLMAP, PIL, ...); the package marker written here exports the two names its callers import
    from `scene` and nothing else."""
    marker = os.path.join(OUT_APP, ".installed")
    if os.path.exists(marker) and not force:
        return OUT_APP
    if not os.path.isdir(REF_ROOT):
        raise RuntimeError("reference sources not present at " + REF_ROOT)
    for rel in APP_FILES:
        dst = os.path.join(OUT_APP, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF_ROOT, rel), dst)
    with open(os.path.join(OUT_APP, "scene", "__init__.py"), "w") as f:
        f.write("# written by oracle/build_ref.py (not the reference's scene/__init__.py, which imports its dataset readers)\n"
                "from scene.gaussian_model import GaussianModel  # noqa: F401\n"
                "from scene.cameras import Camera  # noqa: F401\n")
    with open(os.path.join(OUT_APP, "utils", "__init__.py"), "w") as f:
        f.write("")
    with open(marker, "w") as f:
        f.write("ok\n")
    return OUT_APP


if __name__ == "__main__":
    force = "--force" in sys.argv
    if "--only" in sys.argv:
        what = sys.argv[sys.argv.index("--only") + 1]
        print({"dgr": build, "ncc": build_ncc, "app": build_app}[what](force=force))
    else:
        # one process per extension: torch.utils.cpp_extension.load() renames a second module called "_C" built in
        # the same process ("_C_v1"), and the reference's PYBIND11_MODULE(TORCH_EXTENSION_NAME, ...) follows the name
        import subprocess

        for what in ("dgr", "ncc", "app"):
            subprocess.check_call([sys.executable, os.path.abspath(__file__), "--only", what] + (["--force"] if force else []))
