"""`distCUDA2` of simple-knn on libfisher_b200.so.

Mirrors simple_knn._C.distCUDA2 (KNN/ext.cpp:16, KNN/spatial.cu:15-26): points f32[P,3] CUDA ->
f32[P], the mean of the three smallest squared distances to other points (KNN/simple_knn.cu:148-184).
"""
import torch

from . import _lib
from .rasterizer import _on_tensor_device, _stream


@_on_tensor_device
def distCUDA2(points):
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 takes a CUDA tensor (there is no CPU path)")
    L = _lib.lib()
    P = points.size(0)
    pts = points.contiguous().float()
    means = torch.zeros((P,), dtype=torch.float32, device=points.device)   # spatial.cu:19-20
    if P == 0:
        return means
    ws = torch.empty((L.fb_knn_workspace_bytes(P),), dtype=torch.uint8, device=points.device)
    _lib.check(L.fb_dist2_knn3(P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "distCUDA2")
    return means
