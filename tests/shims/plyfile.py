"""Import shim: the reference's scene/gaussian_model.py imports plyfile for save_ply / load_ply only; the rasterizer
tests build the model from tensors.  Not in this image."""


class _Unavailable:
    def __init__(self, *a, **k):
        raise NotImplementedError("plyfile is not installed in this image (import shim for the reference's GaussianModel)")

    describe = __init__


PlyData = _Unavailable
PlyElement = _Unavailable
