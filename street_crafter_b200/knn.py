"""`distCUDA2` — drop-in for `simple_knn._C.distCUDA2`
(/root/reference/submodules/simple-knn/ext.cpp:16, spatial.cu:16-25; callers
street_gaussian/models/gaussian_model.py:65, gaussian_model_actor.py:139,
data_processor/utils/render_utils.py:125)."""
from __future__ import annotations

import torch
from torch import Tensor

from ._lib import check, lib


@torch.no_grad()
def distCUDA2(points: Tensor) -> Tensor:
    """points [P,3] float32 CUDA -> [P] mean squared distance to the 3 nearest other points."""
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 needs a CUDA tensor: street_crafter_b200 has no CPU path")
    assert points.dim() == 2 and points.shape[1] == 3, points.shape
    pts = points.detach().float().contiguous()
    P = pts.shape[0]
    out = torch.zeros(P, dtype=torch.float32, device=pts.device)  # torch::full({P}, 0.0) in the reference
    if P == 0:
        return out
    ws_bytes = int(lib.sc_knn_workspace_bytes(P))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=pts.device)
    with torch.cuda.device(pts.device):
        check(lib.sc_knn3_mean_dist2(P, pts.data_ptr(), out.data_ptr(), ws.data_ptr(), ws_bytes,
                                     torch.cuda.current_stream().cuda_stream), "sc_knn3_mean_dist2")
    return out
