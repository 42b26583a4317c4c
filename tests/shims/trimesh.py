"""Import shim for the one trimesh call the reference's GaussianModel makes: trimesh.creation.box() in get_tetra_points
(scene/gaussian_model.py:499), whose eight vertices are the corners of the unit cube centred at the origin."""
import numpy as np


class _Box:
    def __init__(self):
        self.vertices = np.array([[x, y, z] for x in (-0.5, 0.5) for y in (-0.5, 0.5) for z in (-0.5, 0.5)], dtype=np.float64)


class creation:  # noqa: N801  (mirrors the module name trimesh.creation)
    @staticmethod
    def box():
        return _Box()
