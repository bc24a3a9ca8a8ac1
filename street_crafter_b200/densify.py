"""Per-iteration densification statistics: the reference's
`set_max_radii2D` + `add_densification_stats`
(/root/reference/street_gaussian/models/street_gaussian_model.py:486-497, :503-519; called at
/root/reference/train.py:283-284) as ONE kernel over the Gaussians instead of ~10 boolean-index /
cat / norm / max launches.  No CPU path."""
from __future__ import annotations

import torch
from torch import Tensor

from ._lib import check, lib


@torch.no_grad()
def densification_stats(viewspace_point_tensor: Tensor, visibility_filter: Tensor, radii: Tensor, width: int,
                        height: int, xyz_gradient_accum: Tensor, denom: Tensor, max_radii2D: Tensor,
                        start: int = 0, end: int = None) -> None:
    """viewspace_point_tensor: the renderer's `viewspace_points` ([1,N,2] or [N,2]) AFTER backward, carrying
    `.grad` and `.absgrad`; visibility_filter [N] bool; radii [N] float (the renderer's `radii`);
    xyz_gradient_accum [N,2], denom [N,1] or [N], max_radii2D [N] are updated IN PLACE (contiguous fp32 —
    the arrays of ONE model of the reference's scene graph).  `start:end` selects that model's rows of the
    renderer outputs (`graph_gaussian_range[model_name]`); default: all rows."""
    g = viewspace_point_tensor.grad
    a = getattr(viewspace_point_tensor, "absgrad", None)
    if g is None or a is None:
        raise RuntimeError("viewspace_point_tensor needs .grad and .absgrad (render with absgrad=True, then backward)")
    if not g.is_cuda:
        raise RuntimeError("densification_stats needs CUDA tensors: street_crafter_b200 has no CPU path")
    g = g.reshape(-1, 2)[start:end].float().contiguous()
    a = a.reshape(-1, 2)[start:end].float().contiguous()
    N = g.shape[0]
    vis = visibility_filter.reshape(-1)[start:end]
    vis = (vis if vis.dtype == torch.bool else vis != 0).contiguous().view(torch.uint8)
    rad = radii.reshape(-1)[start:end].float().contiguous()
    assert vis.numel() == N and rad.numel() == N
    for t, n in ((xyz_gradient_accum, 2 * N), (denom, N), (max_radii2D, N)):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n, (t.shape, t.dtype)
    with torch.cuda.device(g.device):
        check(lib.sc_densification_stats(N, a.data_ptr(), g.data_ptr(), vis.data_ptr(), rad.data_ptr(), int(width),
                                         int(height), xyz_gradient_accum.data_ptr(), denom.data_ptr(),
                                         max_radii2D.data_ptr(), torch.cuda.current_stream().cuda_stream),
              "sc_densification_stats")
