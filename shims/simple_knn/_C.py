"""Shim: `from simple_knn._C import distCUDA2` (KNN/ext.cpp:16) resolves to the B200 implementation."""
from fisher_nerf_customized_b200.knn import distCUDA2  # noqa: F401
