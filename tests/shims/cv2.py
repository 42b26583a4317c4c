"""Import shim: utils/graphics_utils.py and utils/loss_utils.py import cv2 for debug images and one remap helper that the
rasterizer callers under test never reach.  Not in this image."""
INTER_LINEAR = 1
COLORMAP_JET = 2


def __getattr__(name):
    def _missing(*a, **k):
        raise NotImplementedError(f"cv2.{name}: OpenCV is not installed in this image (import shim)")

    return _missing
