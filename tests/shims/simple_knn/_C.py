def distCUDA2(*a, **k):
    raise NotImplementedError("simple_knn is not installed in this image (import shim for the reference's GaussianModel)")
