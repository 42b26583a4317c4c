"""Import shim for simple_knn (used by GaussianModel.create_from_pcd only; not in this image)."""
