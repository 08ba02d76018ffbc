"""`simple_knn._C`: distCUDA2(points f32[P,3] cuda) -> f32[P], the mean squared distance of every point
to its 3 nearest other points (KNN/spatial.cu:15-26, KNN/simple_knn.cu:185-221).

Backed by csb_dist_cuda2 of libcloth_splatting_b200.so (csrc/knn.cu); no CPU fallback."""
import torch

import csb_cabi as _abi


def distCUDA2(points):
    if not points.is_cuda:
        raise ValueError("distCUDA2: points must be a CUDA tensor")
    P = points.size(0)
    pts = points.detach().contiguous().float()  # spatial.cu:17-18 (contiguous().data<float>())
    means = torch.full((P,), 0.0, dtype=torch.float32, device=points.device)
    if P == 0:
        return means
    lib = _abi.lib()
    ws = torch.empty(lib.csb_dist_cuda2_bytes(P), dtype=torch.uint8, device=points.device)
    with torch.cuda.device(points.device):
        rc = lib.csb_dist_cuda2(P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(), ws.numel(),
                                _abi.current_stream_ptr())
    _abi.check(rc, "csb_dist_cuda2")
    return means
